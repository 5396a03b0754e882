// TEST INFRASTRUCTURE ONLY -- never linked into, imported by or executed from the product path.
//
// Headless C-ABI harness around the UNMODIFIED reference sources
//   /root/reference/src/FEM/FEMContext.cpp   (assembly, idx_map)
//   /root/reference/src/FEM/CPUSolver.cpp    (Eigen Jacobi-PCG time stepping, fp32)
// which are compiled where they lie (see oracle/Makefile, target _ref) and linked with this
// file into oracle/_ref/libfea_ref*.so.  Nothing from the reference is copied into the repo.
//
// The harness fills the plain data members of `Surface` (include/Utils/Surface.hpp:36-52) the way
// Application::init_surface_from_obj does (src/Application.cpp:615-630) and then drives the same
// call sequence as Application::run (src/Application.cpp:530-540): write values -> advance_time()
// -> has_numerical_instability().
//
// Private members (FEMContext matrices, CPUSolver u/v) are read by building THIS translation unit
// with -fno-access-control; the reference translation units are built without it.
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#include "FEM/CPUSolver.hpp"
#include "FEM/FEMContext.hpp"

#ifdef _OPENMP
#include <omp.h>
#endif

namespace {
struct RefHandle {
    std::shared_ptr<Surface> surface;
    std::shared_ptr<FEMContext> ctx;
    std::shared_ptr<CPUSolver> solver;
};

const Eigen::SparseMatrix<float>& pick(RefHandle* h, int which) {
    switch (which) {
        case 0: return h->ctx->mass_matrix;
        case 1: return h->ctx->stiffness_matrix;
        default: return h->ctx->advection_matrix;
    }
}
}  // namespace

extern "C" {

// bc: 0 Dirichlet, 1 Neumann (include/FEM/FEMContext.hpp:16-19)
void* ref_create(int64_t V, const float* xyz, int64_t T, const uint32_t* tri, const uint8_t* on_boundary, int bc) {
    auto* h = new RefHandle();
    h->surface = std::make_shared<Surface>();
    Surface& s = *h->surface;
    s.vertices.resize(V);
    for (int64_t i = 0; i < V; i++) s.vertices[i] = glm::vec3(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
    s.triangles.resize(T);
    for (int64_t k = 0; k < T; k++) {
        s.triangles[k].idx_a = tri[3 * k];
        s.triangles[k].idx_b = tri[3 * k + 1];
        s.triangles[k].idx_c = tri[3 * k + 2];
    }
    s.on_boundary.assign(V, false);
    s.num_boundary_points = 0;
    for (int64_t i = 0; i < V; i++) {
        s.on_boundary[i] = on_boundary[i] != 0;
        if (on_boundary[i]) s.num_boundary_points++;
    }
    s.values.assign(V, 0.0f);
    s.closed = s.num_boundary_points == 0;
    s.initialized = true;

    h->ctx = std::make_shared<FEMContext>();
    h->ctx->boundary_condition = static_cast<BoundaryCondition>(bc);
    h->solver = std::make_shared<CPUSolver>(h->ctx);
    h->ctx->init_from_surface(h->surface);  // -> update_boundary_conditions -> assemble_matrices
    h->solver->clear_values();
    return h;
}

void ref_destroy(void* p) { delete static_cast<RefHandle*>(p); }

void ref_set_boundary_condition(void* p, int bc) {
    auto* h = static_cast<RefHandle*>(p);
    h->ctx->boundary_condition = static_cast<BoundaryCondition>(bc);
    h->ctx->update_boundary_conditions();
    h->solver->clear_values();
}

// eq: 0 Heat, 1 Wave, 2 Advection_Diffusion, 3 Reaction_Diffusion (include/FEM/FEMContext.hpp:9-14)
void ref_set_equation(void* p, int eq) { static_cast<RefHandle*>(p)->ctx->equation = static_cast<Equation>(eq); }

// Parameter block, same order for every equation: {dt, c0, c1, c2, c3}
//  Heat: c0 = conductivity; Wave: c0 = c; AdvDiff: c0 = c; ReactionDiffusion: c0..c3 = Du, Dv, feed, kill
void ref_set_params(void* p, int eq, const float* q) {
    auto* h = static_cast<RefHandle*>(p);
    auto e = static_cast<Equation>(eq);
    h->ctx->parameters[e]->time_step = q[0];
    switch (e) {
        case Equation::Heat: std::static_pointer_cast<HeatParameters>(h->ctx->parameters[e])->conductivity = q[1]; break;
        case Equation::Wave: std::static_pointer_cast<WaveParameters>(h->ctx->parameters[e])->c = q[1]; break;
        case Equation::Advection_Diffusion:
            std::static_pointer_cast<AdvectionDiffusionParameters>(h->ctx->parameters[e])->c = q[1];
            break;
        case Equation::Reaction_Diffusion: {
            auto r = std::static_pointer_cast<ReactionDiffusionParameters>(h->ctx->parameters[e]);
            r->Du = q[1];
            r->Dv = q[2];
            r->feed_rate = q[3];
            r->kill_rate = q[4];
        } break;
    }
}

// velocity change re-assembles all three matrices, as src/Application.cpp:393-394 does
void ref_set_velocity(void* p, const float* vel) {
    auto* h = static_cast<RefHandle*>(p);
    std::static_pointer_cast<AdvectionDiffusionParameters>(h->ctx->parameters[Equation::Advection_Diffusion])->velocity =
        Eigen::Vector3f(vel[0], vel[1], vel[2]);
    h->ctx->assemble_matrices();
}

void ref_assemble(void* p) { static_cast<RefHandle*>(p)->ctx->assemble_matrices(); }

int64_t ref_num_nodes(void* p) { return static_cast<RefHandle*>(p)->ctx->num_nodes(); }
int64_t ref_num_unknowns(void* p) { return static_cast<RefHandle*>(p)->ctx->num_unknowns(); }
int ref_max_row_nonzeros(void* p) { return static_cast<RefHandle*>(p)->ctx->num_max_nonzeros_per_row(); }
void ref_get_idx_map(void* p, int32_t* out) {
    auto* h = static_cast<RefHandle*>(p);
    std::memcpy(out, h->ctx->idx_map.data(), h->ctx->idx_map.size() * sizeof(int));
}

// which: 0 mass, 1 stiffness, 2 advection.  Eigen storage is compressed column-major (CSC).
int64_t ref_nnz(void* p, int which) { return pick(static_cast<RefHandle*>(p), which).nonZeros(); }
void ref_get_csc(void* p, int which, int32_t* outer, int32_t* inner, float* val) {
    const auto& m = pick(static_cast<RefHandle*>(p), which);
    std::memcpy(outer, m.outerIndexPtr(), (m.outerSize() + 1) * sizeof(int));
    std::memcpy(inner, m.innerIndexPtr(), m.nonZeros() * sizeof(int));
    std::memcpy(val, m.valuePtr(), m.nonZeros() * sizeof(float));
}

void ref_clear(void* p) {
    auto* h = static_cast<RefHandle*>(p);
    h->solver->clear_values();
    std::fill(h->surface->values.begin(), h->surface->values.end(), 0.0f);  // Surface::clear_values
}
void ref_set_values(void* p, const float* v) {
    auto* h = static_cast<RefHandle*>(p);
    std::memcpy(h->surface->values.data(), v, h->surface->values.size() * sizeof(float));
}
void ref_get_values(void* p, float* v) {
    auto* h = static_cast<RefHandle*>(p);
    std::memcpy(v, h->surface->values.data(), h->surface->values.size() * sizeof(float));
}
// which: 0 = u, 1 = v (CPUSolver private state, include/FEM/CPUSolver.hpp:17-18)
void ref_get_state(void* p, int which, float* out) {
    auto* h = static_cast<RefHandle*>(p);
    const Eigen::VectorXf& x = which == 0 ? h->solver->u : h->solver->v;
    std::memcpy(out, x.data(), x.size() * sizeof(float));
}
void ref_step(void* p, int nsteps) {
    auto* h = static_cast<RefHandle*>(p);
    for (int i = 0; i < nsteps; i++) h->solver->advance_time();
}
int ref_unstable(void* p) { return static_cast<RefHandle*>(p)->solver->has_numerical_instability() ? 1 : 0; }

// Eigen only parallelises the row-major SpMV (SparseDenseProduct.h:47-58) and only with -fopenmp
int ref_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) {
        omp_set_num_threads(n);
        Eigen::setNbThreads(n);
    }
    return Eigen::nbThreads();
#else
    (void)n;
    return 1;
#endif
}
}
