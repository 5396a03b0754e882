/* feab_surface.h -- C ABI of the rows either side of the time-stepping hot path:
 * the device-resident mirror of the reference's `Surface` object (include/Utils/Surface.hpp:34-73).
 *
 *   N2  mesh ingest:     boundary-edge detection of Surface::init_from_obj        src/Utils/Surface.cpp:103-127
 *   N3  smooth normals:  shaders/FEM/smooth_normals.glsl:30-57, dispatched by
 *                        Surface::calculate_normals                                src/Utils/Surface.cpp:287-306
 *   N4  brush / picking: BVH::ray_triangle_intersection + Application::brush      src/Utils/BVH.cpp:32-40, :182-216,
 *                                                                                  src/Application.cpp:707-727
 *   N1  .ply exporter:   Surface::export_to_ply (threshold clipping, vertex
 *                        de-duplication, colour map) + ColorMap::get_color         src/Utils/Surface.cpp:138-285,
 *                                                                                  src/Utils/ColorMap.cpp:12-14
 *
 * Same conventions as feab.h: plain pointers and sizes, `int` status (0 ok, negative FEAB_E* code, positive
 * cudaError_t), no exceptions across the boundary, one handle = one Surface on one GPU, not thread-safe per handle.
 * There is no CPU fallback: feas_create fails with FEAB_ENOGPU without a CUDA device.
 * All `const float*` / `uint32_t*` arguments are HOST pointers unless the name ends in `_device`.
 */
#ifndef FEAB_SURFACE_H
#define FEAB_SURFACE_H
#include <stdint.h>

#include "feab.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct FeasHandle FeasHandle;

/* enum class MeshType (include/Utils/Surface.hpp:25-29) */
enum { FEAS_MESH_OPEN = 0, FEAS_MESH_CLOSED = 1, FEAS_MESH_MIRRORED = 2 };

int feas_create(int device, FeasHandle** out);
int feas_destroy(FeasHandle* h);
const char* feas_last_error(const FeasHandle* h);

/* Surface::vertices / normals / triangles (include/Utils/Surface.hpp:36-39).  normals == NULL: (0,1,0) for every
 * vertex, what init_from_PSLG sets (src/Utils/Surface.cpp:45).  Values are zeroed (Surface.cpp:46, :128). */
int feas_set_mesh(FeasHandle* h, int64_t num_vertices, const float* xyz, const float* normals, int64_t num_triangles,
                  const uint32_t* tri);
/* N2 -- the boundary-edge pass of Surface::init_from_obj (src/Utils/Surface.cpp:103-127): a vertex is on the boundary
 * iff it touches an edge used by exactly one triangle.  on_boundary_out: num_vertices bytes (may be NULL);
 * num_boundary_points / closed as the reference sets them (:123-129). */
int feas_detect_boundary(FeasHandle* h, uint8_t* on_boundary_out, int64_t* num_boundary_points, int* closed);
/* Host-side OBJ ingest with the semantics of Surface::init_from_obj (src/Utils/Surface.cpp:58-101): `v` lines in file
 * order, faces fan-triangulated, per-vertex normals = normalised sum of the face-corner `vn`s; files without
 * normals are rejected like the reference does (:73-74).  Then feas_set_mesh + feas_detect_boundary. */
int feas_load_obj(FeasHandle* h, const char* path);
int feas_get_counts(const FeasHandle* h, int64_t* num_vertices, int64_t* num_triangles, int64_t* num_boundary_points);
/* any output may be NULL */
int feas_get_mesh(FeasHandle* h, float* xyz, float* normals, uint32_t* tri, uint8_t* on_boundary);

/* Surface::values (include/Utils/Surface.hpp:38) */
int feas_set_values(FeasHandle* h, const float* values);
int feas_get_values(FeasHandle* h, float* values_out);
/* Zero-copy hand-over from the solver: copies num_vertices floats device-to-device from `values_device`
 * (see feab_values_device below) on the Surface's stream. */
int feas_set_values_device(FeasHandle* h, const float* values_device);
/* Device address of the solver's `surface->values` mirror (length num_nodes); valid until the next feab_set_mesh. */
int feab_values_device(FeabHandle* h, const float** values_device);

/* ColorMap::coeffs (include/Utils/ColorMap.hpp:14): 7 x rgb.  Default: Viridis (src/Utils/AssetStorage.cpp:141-152). */
int feas_set_color_map(FeasHandle* h, const float* coeffs21);
/* ColorMap::get_color for `count` values (src/Utils/ColorMap.cpp:12-14), Horner form, bit-exact fp32 */
int feas_color_map(FeasHandle* h, int64_t count, const float* t, float* rgb_out);

/* N3 -- smooth_normals.glsl: out[v] = sum over incident triangles of cross(b' - a', c' - a') with
 * p' = p + min(1, vertex_extrusion * max(0, value)) * normal.  Un-normalised, like the shader's output buffer.
 * Float atomics: the summation order is arbitrary, exactly as in the shader.  normals_out: num_vertices x 3. */
int feas_smooth_normals(FeasHandle* h, float vertex_extrusion, float* normals_out);

/* N1 -- Surface::export_to_ply (src/Utils/Surface.cpp:138-285).
 * feas_ply_build runs the clipping, de-duplication (first occurrence wins, in the reference's insertion order) and
 * colouring on the device and leaves the result there; feas_ply_fetch copies it out; feas_export_ply = build + fetch +
 * the ASCII writer (std::to_chars shortest form == std::format("{}")), byte-identical to the reference's file. */
int feas_ply_build(FeasHandle* h, float vertex_extrusion, float threshold, int mesh_type, int64_t* num_out_vertices,
                   int64_t* num_out_triangles);
/* positions / colors: num_out_vertices x 3 float; values: num_out_vertices; tri: num_out_triangles x 3; any may be NULL */
int feas_ply_fetch(FeasHandle* h, float* positions, float* values, float* colors, uint32_t* tri);
int feas_export_ply(FeasHandle* h, const char* path, float vertex_extrusion, float threshold, int mesh_type);
/* The host half alone (no GPU involved, no handle): the ASCII writer of Surface::export_to_ply (src/Utils/Surface.cpp:263-285)
 * on arrays that are already clipped / de-duplicated / coloured.  positions, colors: nv x 3; tri: nt x 3.  0 or FEAB_EINVAL. */
int feas_write_ply(const char* path, int64_t nv, const float* positions, const float* colors, int64_t nt, const uint32_t* tri);
/* The OBJ reader alone (host only): counts first (outputs NULL), then the arrays.  Semantics of Surface::init_from_obj :58-101.
 * Returns FEAB_EINVAL for unreadable files, files without `v` or without `vn` lines (the reference throws). */
int feas_parse_obj(const char* path, int64_t* num_vertices, int64_t* num_triangles, float* xyz, float* normals, uint32_t* tri);

/* N4 -- BVH::ray_triangle_intersection (src/Utils/BVH.cpp:32-40): nearest triangle hit by the ray with distance > 0,
 * by the reference's own plane + three-edge test (:182-216), evaluated for every triangle on the device (no hierarchy).
 * tri_idx = -1 when nothing is hit (point / distance untouched).  Exact distance ties go to the lowest triangle index. */
int feas_pick(FeasHandle* h, const float* origin3, const float* direction3, int* tri_idx, float* point3, float* distance);
/* Application::brush (src/Application.cpp:707-727): pick, then set Surface::values[node] = value for the corner of the hit
 * triangle closest to the hit point (distance > 0).  node_out = that node or -1.  The caller forwards the write to the
 * solver with feab_brush(solver, &node, 1, value), like Application does through surface->values. */
int feas_brush(FeasHandle* h, const float* origin3, const float* direction3, float value, int* node_out);

/* N2 once more, for callers that must not allocate GPU memory.  No counterpart in the reference (its Surface owns its own
 * std::map and its own GL buffers); added for host programs that cannot afford feas_detect_boundary's per-call
 * cudaMalloc/cudaFree of the edge table: a render loop that the device-wide synchronisation of cudaMalloc would stall, a
 * plugin inside an application whose memory pool already owns all of the GPU, code that is forbidden to enter the CUDA
 * allocator after start-up.  The entry point allocates NOTHING (no cudaMalloc/cudaFree, no allocating library call, no
 * handle): the caller passes device buffers it already owns, and the edge table lives in a fixed 24 MiB array inside
 * libfeab's device image, which every GPU that loads the library holds until the process exits.
 *   device               CUDA ordinal the two buffers live on (made current for the call, the caller's device is restored)
 *   stream               a cudaStream_t, or NULL for the default stream; the call WAITS for it before returning
 *   tri_device           num_triangles x 3 uint32, device memory (may be NULL when num_triangles == 0)
 *   on_boundary_device   num_vertices bytes, device memory; overwritten with 0 / 1 like feas_detect_boundary's output
 *   num_boundary_points  host, optional: how many flags are set (0 = closed mesh)
 *   errbuf / errbuf_len  host, optional: receives the message when the return value is not 0 (there is no handle to ask)
 * Limits and rules: num_triangles <= FEAS_NOALLOC_MAX_TRIANGLES, larger meshes are refused with FEAB_EINVAL (use
 * feas_detect_boundary); a triangle index >= num_vertices is reported as FEAB_EINVAL instead of being written out of
 * bounds.  All calls on a device share the one table, which still holds the previous call's edges: each call wipes what it
 * needs first, and two calls running on the GPU at the same time would corrupt each other -- the library serialises its
 * callers with a lock and the final wait, so expect no concurrency from it.  The first call on a device also looks up the
 * table's device address (one-time set-up): make that call before the process begins any stream capture. */
#define FEAS_NOALLOC_MAX_TRIANGLES 349525   /* 6 x triangles <= 2^21 table slots (load factor <= 0.5) */
int feas_detect_boundary_noalloc(int device, void* stream, int64_t num_vertices, int64_t num_triangles, const uint32_t* tri_device,
                                 uint8_t* on_boundary_device, int64_t* num_boundary_points, char* errbuf, int64_t errbuf_len);

/* measurement: device milliseconds of the last feas_ply_build / feas_detect_boundary / feas_smooth_normals / feas_pick (CUDA events
 * on the Surface's stream), and the number of kernels this handle has launched */
int feas_last_ms(const FeasHandle* h, double* ms, int64_t* kernel_launches);

#ifdef __cplusplus
}
#endif
#endif /* FEAB_SURFACE_H */
