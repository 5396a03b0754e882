// TEST INFRASTRUCTURE ONLY -- never linked into, imported by or executed from the product path.
//
// Headless C-ABI harness around the UNMODIFIED reference sources
//   /root/reference/src/Utils/Surface.cpp    (OBJ ingest + boundary-edge detection :58-133, .ply exporter :138-285,
//                                             PSLG triangulation hand-off :25-51, :431-486)
//   /root/reference/src/Utils/ColorMap.cpp   (6th-order polynomial colour map :12-14)
//   /root/reference/src/Utils/BVH.cpp        (median-split BVH + ray / triangle test used by the brush :8-216)
//   /root/reference/src/Utils/PSLG.cpp, Shader.cpp, Mesh.cpp, lib/glad/src/glad.c, lib/triangle/triangle.cpp
// compiled where they lie (oracle/Makefile, target _ref/libfea_surface_ref.so).  Nothing from the reference is copied.
//
// Two things are supplied from outside the reference:
//   * <tinyobjloader/tiny_obj_loader.h>: the submodule is empty in the snapshot; oracle/stubs/ holds an API-compatible
//     stand-in (see its header for what that does and does not pin);
//   * OpenGL: there is no graphics stack.  glad.c defines every gl* entry point as a NULL function pointer; the few
//     that Surface / PSLG call on the paths used here (buffer creation in load_buffers) are pointed at typed no-ops
//     below.  No GL library is loaded, no context is created; the data members the hot path and the exporter read
//     (vertices, normals, values, triangles, on_boundary) are plain std::vectors.
#include <glad/glad.h>
// PSLG.cpp references stb_image (stencil images, never used here); the reference instantiates it in src/main.cpp
#define STB_IMAGE_IMPLEMENTATION
#include <stb/stb_image.h>

#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <type_traits>
#include <vector>

#include "Utils/BVH.hpp"
#include "Utils/ColorMap.hpp"
#include "Utils/PSLG.hpp"
#include "Utils/Surface.hpp"

namespace {

template <class F>
struct Noop;
template <class R, class... A>
struct Noop<R (*)(A...)> {
    static R fn(A...) {
        if constexpr (!std::is_void_v<R>) return R();
    }
};
#define GL_NOOP(name) glad_##name = Noop<decltype(glad_##name)>::fn

void install_gl_noops() {
    static bool done = false;
    if (done) return;
    done = true;
    GL_NOOP(glBindBuffer);
    GL_NOOP(glBindBufferBase);
    GL_NOOP(glBindVertexArray);
    GL_NOOP(glBufferData);
    GL_NOOP(glClearBufferData);
    GL_NOOP(glEnableVertexAttribArray);
    GL_NOOP(glGenBuffers);
    GL_NOOP(glGenVertexArrays);
    GL_NOOP(glGetBufferSubData);
    GL_NOOP(glMemoryBarrier);
    GL_NOOP(glVertexAttribPointer);
    GL_NOOP(glDeleteBuffers);
    GL_NOOP(glDeleteVertexArrays);
}

struct SrfHandle {
    std::shared_ptr<Surface> surface;
    std::shared_ptr<BVH> bvh;
    std::string err;
};

}  // namespace

extern "C" {

void* srf_create() {
    install_gl_noops();
    auto* h = new SrfHandle();
    h->surface = std::make_shared<Surface>();
    // Viridis, the first map AssetStorage registers (src/Utils/AssetStorage.cpp:141-152); srf_set_colormap replaces it
    h->surface->color_map = std::make_shared<ColorMap>(
        "Viridis", std::array<glm::vec3, 7>({glm::vec3(0.274344, 0.004462, 0.331359), glm::vec3(0.108915, 1.397291, 1.388110),
                                             glm::vec3(-0.319631, 0.243490, 0.156419), glm::vec3(-4.629188, -5.882803, -19.646115),
                                             glm::vec3(6.181719, 14.388598, 57.442181), glm::vec3(4.876952, -13.955112, -66.125783),
                                             glm::vec3(-5.513165, 4.709245, 26.582180)}));
    return h;
}

void srf_destroy(void* p) { delete static_cast<SrfHandle*>(p); }
const char* srf_last_error(void* p) { return static_cast<SrfHandle*>(p)->err.c_str(); }

void srf_set_colormap(void* p, const float* c21) {
    auto* h = static_cast<SrfHandle*>(p);
    std::array<glm::vec3, 7> c;
    for (int i = 0; i < 7; i++) c[i] = glm::vec3(c21[3 * i], c21[3 * i + 1], c21[3 * i + 2]);
    h->surface->color_map = std::make_shared<ColorMap>("custom", c);
}

// ColorMap::get_color for n values
void srf_get_color(void* p, int64_t n, const float* t, float* rgb) {
    auto* h = static_cast<SrfHandle*>(p);
    for (int64_t i = 0; i < n; i++) {
        glm::vec3 c = h->surface->color_map->get_color(t[i]);
        rgb[3 * i] = c.r;
        rgb[3 * i + 1] = c.g;
        rgb[3 * i + 2] = c.b;
    }
}

// Surface::init_from_obj (src/Utils/Surface.cpp:58-133), exceptions -> -1 + srf_last_error
int srf_init_from_obj(void* p, const char* path) {
    auto* h = static_cast<SrfHandle*>(p);
    try {
        h->surface->init_from_obj(path);
    } catch (const std::exception& e) {
        h->err = e.what();
        return -1;
    }
    return 0;
}

// Surface::init_from_PSLG (src/Utils/Surface.cpp:25-51) on a closed polygon given as XZ points; one closed loop
int srf_init_from_pslg(void* p, int64_t nv, const float* xz, float triangle_area) {
    auto* h = static_cast<SrfHandle*>(p);
    try {
        PSLG pslg;
        for (int64_t i = 0; i < nv; i++) {
            pslg.vertices.push_back(glm::vec3(xz[2 * i], 0.0f, xz[2 * i + 1]));
            pslg.indices.push_back((unsigned)i);
            pslg.indices.push_back((unsigned)((i + 1) % nv));
        }
        pslg.section_start_idx = (int)pslg.vertices.size();  // what PSLG::finalize leaves (closed() == true)
        pslg.triangle_area = triangle_area;
        h->surface->init_from_PSLG(pslg);
    } catch (const std::exception& e) {
        h->err = e.what();
        return -1;
    }
    return 0;
}

// Fill the Surface data members directly (include/Utils/Surface.hpp:36-40)
void srf_set_mesh(void* p, int64_t V, const float* xyz, const float* nrm, int64_t T, const uint32_t* tri, const uint8_t* onb) {
    auto* h = static_cast<SrfHandle*>(p);
    Surface& s = *h->surface;
    s.vertices.resize(V);
    s.normals.resize(V);
    for (int64_t i = 0; i < V; i++) {
        s.vertices[i] = glm::vec3(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
        s.normals[i] = nrm ? glm::vec3(nrm[3 * i], nrm[3 * i + 1], nrm[3 * i + 2]) : glm::vec3(0.0f, 1.0f, 0.0f);
    }
    s.triangles.resize(T);
    for (int64_t k = 0; k < T; k++) {
        s.triangles[k].idx_a = tri[3 * k];
        s.triangles[k].idx_b = tri[3 * k + 1];
        s.triangles[k].idx_c = tri[3 * k + 2];
    }
    s.on_boundary.assign(V, false);
    s.num_boundary_points = 0;
    if (onb)
        for (int64_t i = 0; i < V; i++) {
            s.on_boundary[i] = onb[i] != 0;
            s.num_boundary_points += onb[i] != 0;
        }
    s.values.assign(V, 0.0f);
    s.closed = s.num_boundary_points == 0;
    s.initialized = true;
}

void srf_counts(void* p, int64_t* V, int64_t* T, int64_t* nb, int* closed) {
    Surface& s = *static_cast<SrfHandle*>(p)->surface;
    *V = (int64_t)s.vertices.size();
    *T = (int64_t)s.triangles.size();
    *nb = s.num_boundary_points;
    *closed = s.closed ? 1 : 0;
}

void srf_get_mesh(void* p, float* xyz, float* nrm, uint32_t* tri, uint8_t* onb) {
    Surface& s = *static_cast<SrfHandle*>(p)->surface;
    for (size_t i = 0; i < s.vertices.size(); i++) {
        if (xyz) std::memcpy(xyz + 3 * i, &s.vertices[i], 12);
        if (nrm && i < s.normals.size()) std::memcpy(nrm + 3 * i, &s.normals[i], 12);
        if (onb) onb[i] = s.on_boundary[i] ? 1 : 0;
    }
    if (tri)
        for (size_t k = 0; k < s.triangles.size(); k++) {
            tri[3 * k] = s.triangles[k].idx_a;
            tri[3 * k + 1] = s.triangles[k].idx_b;
            tri[3 * k + 2] = s.triangles[k].idx_c;
        }
}

void srf_set_values(void* p, const float* v) {
    Surface& s = *static_cast<SrfHandle*>(p)->surface;
    s.values.assign(v, v + s.vertices.size());
}

// Surface::export_to_ply (src/Utils/Surface.cpp:138-285); mesh_type: 0 Open, 1 Closed, 2 Mirrored
int srf_export_to_ply(void* p, const char* path, float vertex_extrusion, float threshold, int mesh_type) {
    auto* h = static_cast<SrfHandle*>(p);
    try {
        h->surface->export_to_ply(path, vertex_extrusion, threshold, static_cast<MeshType>(mesh_type));
    } catch (const std::exception& e) {
        h->err = e.what();
        return -1;
    }
    return 0;
}

// BVH::BVH(surface, max_depth) (src/Utils/BVH.cpp:8-16); Application uses settings.bvh_depth = 10 (include/Application.hpp:43)
void srf_bvh_build(void* p, int max_depth) {
    auto* h = static_cast<SrfHandle*>(p);
    h->bvh = std::make_shared<BVH>(h->surface, max_depth);
}

// BVH::ray_triangle_intersection (src/Utils/BVH.cpp:32-40)
int srf_pick(void* p, const float* o, const float* d, float* point3, float* dist) {
    auto* h = static_cast<SrfHandle*>(p);
    RayTriangleIntersection r = h->bvh->ray_triangle_intersection(glm::vec3(o[0], o[1], o[2]), glm::vec3(d[0], d[1], d[2]));
    if (r.tri_idx != -1) {
        std::memcpy(point3, &r.point, 12);
        *dist = r.distance;
    }
    return r.tri_idx;
}

// Application::brush (src/Application.cpp:707-727) RESTATED: Application.cpp itself drags in the whole GUI and is not
// compiled here.  Same statements, on the reference's own BVH result and Surface members.
int srf_brush(void* p, const float* o, const float* d, float value) {
    auto* h = static_cast<SrfHandle*>(p);
    Surface& surface = *h->surface;
    RayTriangleIntersection intersection = h->bvh->ray_triangle_intersection(glm::vec3(o[0], o[1], o[2]), glm::vec3(d[0], d[1], d[2]));
    int point_idx = -1;
    if (intersection.tri_idx != -1) {
        float min_point_dist = 0.0f;
        for (int i = 0; i < 3; i++) {
            float dist = glm::distance(surface.vertices[surface.triangles[intersection.tri_idx][i]], intersection.point);
            if ((point_idx == -1 || dist < min_point_dist) && dist > 0.0f) {
                min_point_dist = dist;
                point_idx = i;
            }
        }
        if (point_idx != -1) surface.values[surface.triangles[intersection.tri_idx][point_idx]] = value;
    }
    return point_idx != -1 ? (int)surface.triangles[intersection.tri_idx][point_idx] : -1;
}

void srf_get_values(void* p, float* out) {
    Surface& s = *static_cast<SrfHandle*>(p)->surface;
    std::memcpy(out, s.values.data(), s.values.size() * sizeof(float));
}

}  // extern "C"
