/* feab.h -- C ABI of libfeab.so, the B200-native (sm_100a) implementation of fea-visualizer's
 * time-stepping hot path: P1 triangle mass / stiffness / advection assembly into a CSR system and the
 * per-timestep Jacobi-PCG (or BiCGSTAB) solve of (M/dt + c K [- Adv]) x = b for the heat, wave,
 * advection-diffusion and Gray-Scott solvers.
 *
 * Plain pointers and sizes only; every pointer argument is a HOST pointer unless it says "device".
 * Every entry point returns 0 on success, a negative FEAB_E* code or a positive cudaError_t otherwise;
 * feab_last_error() gives the text.  No exceptions cross this boundary.  A handle is not thread-safe
 * (the reference's solvers are single-threaded and non-reentrant, src/Application.cpp:506-565).
 *
 * Each entry point cites the reference interface it replaces (paths relative to the reference root).
 * The host-side C++ mirror of the reference classes that calls this ABI is include/FEM/CUDASolver.hpp;
 * INTEGRATION.md shows the binding a maintainer of the reference would add.
 */
#ifndef FEAB_H
#define FEAB_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct FeabHandle FeabHandle;

/* include/FEM/FEMContext.hpp:9-14 (enum class Equation) */
enum { FEAB_EQ_HEAT = 0, FEAB_EQ_WAVE = 1, FEAB_EQ_ADVECTION_DIFFUSION = 2, FEAB_EQ_REACTION_DIFFUSION = 3 };
/* include/FEM/FEMContext.hpp:16-19 (enum class BoundaryCondition) */
enum { FEAB_BC_DIRICHLET = 0, FEAB_BC_NEUMANN = 1 };
/* FEMContext's three private matrices, include/FEM/FEMContext.hpp:47-49 */
enum { FEAB_MAT_MASS = 0, FEAB_MAT_STIFFNESS = 1, FEAB_MAT_ADVECTION = 2 };
/* Linear solver for the (non-symmetric) advection-diffusion operator.
 * PARITY_CG reproduces the reference: Eigen::ConjugateGradient<.., Lower|Upper> multiplies by A^T
 * (lib/eigen/Eigen/src/IterativeLinearSolvers/ConjugateGradient.h:192-208, src/FEM/CPUSolver.cpp:65,98-99).
 * BICGSTAB solves the same system correctly (Jacobi right-preconditioned, after BiCGSTAB.h:31-119).
 * The three symmetric equations always use the Jacobi-PCG. */
enum { FEAB_SOLVER_PARITY_CG = 0, FEAB_SOLVER_BICGSTAB = 1 };
/* Assembly algorithm.  SCATTER: one thread per triangle, local 3x3 matrices, precomputed CSR scatter map,
 * atomic adds (diagonal sums are order-dependent in the last bit).  ORDERED: one thread per row summing its
 * incident triangles in ascending element order -- Eigen's setFromTriplets order
 * (lib/eigen/Eigen/src/SparseCore/SparseMatrix.h:1475-1511), deterministic. */
enum { FEAB_ASSEMBLY_SCATTER = 0, FEAB_ASSEMBLY_ORDERED = 1 };
/* Solve-loop driver.  PERSISTENT (default): the whole PCG solve is ONE cooperative kernel that stays resident and
 * separates the phases of an iteration by grid barriers (no launches, no host round trips inside a solve).
 * GRAPH_WHILE: one CUDA graph per step with a device-side WHILE node around three kernels per iteration.
 * HOST_BATCHED: the same three kernels launched in batches with a flag poll per batch.  BiCGSTAB uses GRAPH_WHILE
 * when PERSISTENT is selected. */
/* feab_step's sync_values.  ROOT only differs from FULL on a partitioned (multi-GPU) handle: surface->values lives on the
 * host of rank 0 (the viewer process); rank 0 uploads it and forwards it to the peers over NVLink, the peers ignore their
 * values_inout and touch no PCIe, and only rank 0 receives the result. */
enum { FEAB_SYNC_NONE = 0, FEAB_SYNC_FULL = 1, FEAB_SYNC_ROOT = 2 };
enum { FEAB_LOOP_GRAPH_WHILE = 0, FEAB_LOOP_HOST_BATCHED = 1, FEAB_LOOP_PERSISTENT = 2 };

enum {
    FEAB_OK = 0,
    FEAB_EINVAL = -1,   /* bad argument / call order */
    FEAB_ENOMESH = -2,  /* no mesh set */
    FEAB_ENOGPU = -3,   /* no CUDA device: there is no CPU fallback */
    FEAB_ENOMEM = -4,
    FEAB_ECOMM = -5     /* partitioned solve: a peer never answered (bounded device-side wait expired) */
};

/* include/FEM/EquationParameters.hpp:4-49.  Scalars keep the reference's float type; they are promoted to
 * the handle's arithmetic type exactly like the reference's values would be.
 *   Heat:   c[0] = conductivity              Wave: c[0] = c
 *   AdvDiff: c[0] = c                        ReactionDiffusion: c[0..3] = Du, Dv, feed_rate, kill_rate */
typedef struct {
    float time_step;
    float c[4];
} FeabParams;

/* Solver controls; the defaults are Eigen's (IterativeSolverBase.h:251, :368). */
typedef struct {
    int solver;        /* FEAB_SOLVER_*; default PARITY_CG */
    double tolerance;  /* on |r|/|b|; < 0 = epsilon of the arithmetic type (Eigen default) */
    int64_t max_iter;  /* < 0 = 2 n (Eigen default) */
    int loop_mode;     /* FEAB_LOOP_* */
    int assembly;      /* FEAB_ASSEMBLY_* */
} FeabSolverOptions;

typedef struct {
    int64_t iterations[2];     /* per solve of the last step (Gray-Scott has two) */
    double rel_residual[2];    /* sqrt(|r|^2/|b|^2) as Eigen reports it (ConjugateGradient.h:83) */
    double last_call_ms;       /* device time of the last feab_step call (CUDA events on the solver stream) */
    int64_t last_call_steps;
    int64_t kernel_launches;   /* kernels launched by this handle since creation (graph nodes counted per execution) */
    double solve_ms;           /* PERSISTENT loop: device time of the PCG kernel(s) of the LAST step of the last call; else 0 */
    int64_t shared_iterations; /* Gray-Scott: SpMV passes of the last step in which the u- and the v-solve ran in lock-step
                                  from one index stream (pcg_dual.cuh); 0 when the two solves ran one after the other */
    int64_t spmv_passes[2];    /* SpMV (+ update) passes each solve of the last step really ran: Eigen's iterations() is one less
                                  for a solve that stopped on its residual test (ConjugateGradient.h:72-73 breaks before counting) */
} FeabStats;

/* -- lifetime ------------------------------------------------------------------------------- */
/* real_bytes: 4 = fp32 (the reference's arithmetic), 8 = fp64.  device: CUDA ordinal.
 * Replaces the constructors CPUSolver::CPUSolver / FEMContext::FEMContext (src/FEM/CPUSolver.cpp:6-8,
 * src/FEM/FEMContext.cpp:10-15). */
int feab_create(int device, int real_bytes, FeabHandle** out);
int feab_destroy(FeabHandle* h);
const char* feab_last_error(const FeabHandle* h);

/* -- FEMContext ----------------------------------------------------------------------------- */
/* FEMContext::init_from_surface (src/FEM/FEMContext.cpp:21-27): takes the Surface data members the
 * path reads (include/Utils/Surface.hpp:36-40): vertices V x 3 float, triangles T x 3 uint32,
 * on_boundary V bytes.  Builds idx_map, the CSR pattern, the scatter map, and assembles M, K, Adv. */
int feab_set_mesh(FeabHandle* h, int64_t num_vertices, const float* xyz, int64_t num_triangles, const uint32_t* tri,
                  const uint8_t* on_boundary);
/* The same with the context's boundary condition and advection velocity (NULL = keep) taken in the same call: a surface load
 * (src/Application.cpp:593-596, :620-623: init_from_surface with whatever fem_ctx->boundary_condition and velocity hold)
 * costs ONE pattern build and ONE assembly instead of one per setter. */
int feab_set_mesh_ex(FeabHandle* h, int64_t num_vertices, const float* xyz, int64_t num_triangles, const uint32_t* tri,
                     const uint8_t* on_boundary, int bc, const float* velocity3);
/* fem_ctx->boundary_condition = bc; update_boundary_conditions() (src/FEM/FEMContext.cpp:224-240).
 * Does NOT touch solver state; callers follow with feab_clear like Application does (src/Application.cpp:313-316). */
int feab_set_boundary_condition(FeabHandle* h, int bc);
/* parameters[Advection_Diffusion]->velocity = v; assemble_matrices() (src/FEM/FEMContext.cpp:57-64,
 * called from src/Application.cpp:393-394).  velocity == NULL keeps the current one (default (1,0,0)). */
int feab_assemble(FeabHandle* h, const float* velocity3);
/* num_nodes / num_unknowns / num_max_nonzeros_per_row (src/FEM/FEMContext.cpp:33-52) and nnz of the shared pattern */
int feab_get_counts(const FeabHandle* h, int64_t* num_nodes, int64_t* num_unknowns, int64_t* nnz, int* max_row_nonzeros);
/* FEMContext::idx_map (include/FEM/FEMContext.hpp:29) */
int feab_get_idx_map(FeabHandle* h, int32_t* out_num_nodes);
/* The assembled matrices, for parity tests: the reference keeps them private (FEMContext.hpp:47-49).
 * CSR with ascending columns; for the symmetric patterns this equals Eigen's CSC arrays index for index.
 * Call with all-NULL outputs to get only *nnz.  val has the handle's real type. */
int feab_get_csr(FeabHandle* h, int which, int64_t* nnz, int32_t* rowptr, int32_t* col, void* val);

/* -- Solver --------------------------------------------------------------------------------- */
/* fem_ctx->equation = eq (src/Application.cpp:639-641) */
int feab_set_equation(FeabHandle* h, int equation);
/* fem_ctx->parameters[eq] fields (edited live by the GUI, src/Application.cpp:373-423); read at every step */
int feab_set_params(FeabHandle* h, int equation, const FeabParams* params);
int feab_get_params(const FeabHandle* h, int equation, FeabParams* params);
int feab_set_solver_options(FeabHandle* h, const FeabSolverOptions* opt);
int feab_get_solver_options(const FeabHandle* h, FeabSolverOptions* opt);
/* Solver::clear_values (src/FEM/CPUSolver.cpp:54-59) plus Surface::clear_values as Application::clear_solver
 * does (src/Application.cpp:579-583): u, v <- 0 (resized to num_unknowns), values <- 0 */
int feab_clear(FeabHandle* h);
/* surface->values (include/Utils/Surface.hpp:38), device-resident here */
int feab_set_values(FeabHandle* h, const float* values_num_nodes);
int feab_get_values(FeabHandle* h, float* values_num_nodes);
/* Page-locks the host array behind surface->values (std::vector<float>, include/Utils/Surface.hpp:38) so that feab_step's
 * copies of values_inout run at full PCIe speed and asynchronously: cudaHostRegister(values, count * 4, Portable).  The
 * binding calls it whenever surface->values.data() or its size changes -- that only happens on a surface load
 * (src/Application.cpp:593-596, :620-623) -- and feab_step is then given that same pointer.  values == NULL or a new
 * registration undoes the previous one; feab_destroy undoes the last.  Without it feab_step still works from pageable
 * memory (staged by the driver, slower).  The caller keeps the array alive and unmoved while it is registered. */
int feab_register_host_values(FeabHandle* h, float* values, int64_t count);
/* GPUSolver::brush(idx, strength) (include/FEM/GPUSolver.hpp:41, src/FEM/GPUSolver.cpp:44-51) /
 * Application::brush's write surface->values[node] = strength (src/Application.cpp:722) for many nodes */
int feab_brush(FeabHandle* h, const int64_t* nodes, int64_t count, float strength);
/* Solver::advance_time (src/FEM/CPUSolver.cpp:64-147), nsteps times.
 * values_inout != NULL: copied host->device before the first step (the brush may have written it) and, when
 * sync_values != 0, device->host after the last one -- the reference's data contract through surface->values.
 * values_inout == NULL keeps everything device-resident (benchmark mode).  sync_values: FEAB_SYNC_*. */
int feab_step(FeabHandle* h, float* values_inout, int nsteps, int sync_values);
/* Solver::has_numerical_instability (src/FEM/CPUSolver.cpp:41-48): any u_i > 1e4 or v_i > 1e4 */
int feab_unstable(FeabHandle* h, int* out);
int feab_last_stats(FeabHandle* h, FeabStats* out);
/* CPUSolver's private u (which = 0) / v (which = 1), length num_unknowns, handle's real type */
int feab_get_state(FeabHandle* h, int which, void* out);
int feab_set_state(FeabHandle* h, int which, const void* in);

/* -- partitioned (multi-GPU) solve: SURVEY.md 8(e) ------------------------------------------------
 * One handle per GPU ("rank"), all holding the SAME mesh; rows are partitioned over the ranks, every rank
 * assembles and solves its own rows, halo values and dot products travel through peer memory over NVLink inside
 * the solver's kernels.  Call order per rank:
 *   feab_create; feab_mg_config(rank, world); [feab_mg_set_partition]; feab_set_mesh (all ranks, same arrays);
 *   feab_mg_export -> exchange the FeabMgExport blobs between the ranks (any transport; bytes) ->
 *   feab_mg_connect for every peer; then feab_step as usual, all ranks calling it with the same arguments.
 * feab_set_mesh / feab_set_boundary_condition re-allocate the exported arena: export + connect again afterwards.
 * The reference has no counterpart (single process, single thread); the data contract through surface->values is
 * kept: after feab_step(values, n, sync=1) every rank's values array is complete. */
typedef struct {
    unsigned char ipc_handle[64]; /* cudaIpcMemHandle_t of the arena */
    uint64_t base_ptr;            /* device address in the exporting process (same-process connection) */
    uint64_t arena_bytes;
    uint64_t off_vec[5];          /* byte offsets of p[0], p[1], u, v, values inside the arena */
    int64_t n_own, n_ghost, ghost_base, num_nodes;
    int32_t rank, world, device, pid;
} FeabMgExport;

/* Must precede feab_set_mesh.  world <= 8. */
int feab_mg_config(FeabHandle* h, int rank, int world);
/* Optional: part[v] in [0, world) for every mesh vertex, identical on all ranks.  Default: recursive coordinate
 * bisection of the vertex coordinates (feab_partition_rcb), computed inside feab_set_mesh. */
int feab_mg_set_partition(FeabHandle* h, const int32_t* part, int64_t num_vertices);
int feab_partition_rcb(int64_t num_vertices, const float* xyz, int parts, int32_t* part_out);
int feab_mg_export(FeabHandle* h, FeabMgExport* out);
int feab_mg_connect(FeabHandle* h, int peer_rank, const FeabMgExport* peer, int same_process);
/* Local layout: rows [0, n_interior) touch no ghost column; columns are [rows | padding | ghosts], n_cols in all. */
int feab_mg_get_layout(FeabHandle* h, int64_t* n_own, int64_t* n_interior, int64_t* n_ghost, int64_t* n_cols);
/* row_node[n_own]: mesh node of every local row; col_unknown[n_cols]: global unknown id (idx_map numbering) of every
 * local column, -1 for padding.  With these, feab_get_csr / feab_get_state of a rank can be placed in the global system. */
int feab_mg_get_maps(FeabHandle* h, int32_t* row_node, int32_t* col_unknown);
/* part[] actually in use (length num_vertices): the array given to feab_mg_set_partition, or the default recursive coordinate
 * bisection, which feab_set_mesh computes on the device and which equals feab_partition_rcb (host) element for element. */
int feab_mg_get_partition(FeabHandle* h, int32_t* part_out);
/* Test driver: advances `count` handles that live in THIS process on ONE device ("virtual ranks") in lock-step,
 * phase by phase on one stream, so that no kernel ever waits for a kernel that has not been launched yet.  Exercises
 * the partitioned data path (numbering, halo pushes, mailbox all-reduce) without `count` GPUs. */
int feab_mg_group_step(FeabHandle** handles, int count, int nsteps);

/* -- measurement hooks (used by bench.py; not part of the reference interface) ---------------- */
/* Runs `reps` launches of one hot kernel on the current operator and returns the mean device time per
 * launch (CUDA events on the solver stream) and the algorithmic bytes one launch moves.
 * kernel: 0 = SpMV+dot, 1 = PCG vector update, 2 = PCG direction update, 3 = RHS build for the current equation */
int feab_time_kernel(FeabHandle* h, int kernel, int reps, double* ms_per_launch, double* algorithmic_bytes);
/* With FEAB_PCG_PROF=1 (pcg_persistent, pcg_cluster) or 2 (pcg_dual; 3: plus the per-CTA sync arrival spread on stderr) in the environment at feab_create: nanoseconds CTA 0 of the
 * persistent PCG kernel spent in [0] SpMV phase, [1] grid sync merged with the all-reduce of p.Ap, [3] update phase,
 * [4] grid sync merged with the all-reduce of |r|^2, r.z, [6] direction phase (incl. halo stores), [7] grid sync that
 * raises the halo flags; [2], [5], [8] unused (they were separate all-reduce / push stages before the syncs were merged);
 * [9] = iterations counted.  Zeros when profiling is off.
 * FEAB_STEP_PROF=1: [10..15] = device ms of the pieces of the LAST step of the last feab_step call (events on the solver
 * stream): upload (+ root push), gather + RHS kernel, solve kernels, epilogue, value broadcast, download. */
int feab_pcg_profile(FeabHandle* h, double* out16, int reset);
int feab_device_info(const FeabHandle* h, int* sm_count, int64_t* total_mem_bytes, char* name64);
/* Which kernel(s) ran the solve(s) of the LAST step (diagnostics; bench.py labels its roofline entry with it): "pcg_cluster",
 * "pcg_dual + pcg_persistent" (Gray-Scott, lock-step), "pcg_persistent" / "pcg_persistent x2",
 * "bicg_persistent", "graph" (WHILE-graph driver), "host" (host-batched driver), "" before the first step. */
int feab_last_solve_kernel(FeabHandle* h, char* name64);

#ifdef __cplusplus
}
#endif
#endif /* FEAB_H */
