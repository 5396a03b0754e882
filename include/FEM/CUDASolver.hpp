// CUDASolver -- drop-in implementation of the reference's `Solver` interface (reference include/FEM/Solver.hpp:8-15)
// on top of the C ABI of libfeab.so (include/feab.h).  Host code stays C++; everything numerical happens on the GPU.
//
// It is written against the PUBLIC surface of FEMContext / Surface only, so the same header compiles
//   * inside the reference tree, next to CPUSolver / GPUSolver (reference include/FEM/CPUSolver.hpp:9-23), and
//   * headless in this repository against the mirror headers include/FEM/FEMContext.hpp + include/Utils/Surface.hpp.
// Usage is CPUSolver's (reference src/Application.cpp:61-67, :530-540, :579-583):
//     auto solver = std::make_shared<CUDASolver>(fem_ctx);          // throws std::runtime_error without a GPU
//     solver->clear_values();                                        // after (re)loading a surface
//     solver->advance_time();  if (solver->has_numerical_instability()) ...
// Data contract (reference src/FEM/CPUSolver.cpp:13-36): fem_ctx->surface->values is read at entry of advance_time()
// and overwritten at exit; Dirichlet nodes are forced to 0.  Parameters, equation, boundary condition and velocity
// are read fresh at every call, like the reference does.
#pragma once
#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

#include "FEM/Solver.hpp"
#include "feab.h"

namespace feab_detail {
// The mirror FEMContext carries `revision` / `mesh_revision` counters; the reference's does not.  Without them the
// solver falls back to a content signature (sizes, boundary count, BC, velocity) plus the explicit init() call that
// Application already makes for GPUSolver after loading a surface (reference src/Application.cpp:593-596).
template <typename C>
auto mesh_revision_of(C& c, int) -> decltype((unsigned long long)c.mesh_revision) {
    return (unsigned long long)c.mesh_revision;
}
template <typename C>
unsigned long long mesh_revision_of(C&, long) {
    return 0;
}
// FEMContext::max_row_nonzeros is public in the mirror, private in the reference (there CUDASolver would be a friend,
// see INTEGRATION.md); when it is not accessible the value stays available through feab_get_counts.
template <typename C>
auto set_max_row(C& c, int v, int) -> decltype(c.max_row_nonzeros = v, void()) {
    c.max_row_nonzeros = v;
}
template <typename C>
void set_max_row(C&, int, long) {}
}  // namespace feab_detail

class CUDASolver : public Solver {
   public:
    // real_bytes: 8 = fp64 (default; the tolerances of the port), 4 = fp32 (the reference's arithmetic)
    explicit CUDASolver(std::shared_ptr<FEMContext> ctx, int device = 0, int real_bytes = 8) {
        this->fem_ctx = ctx;
        const int rc = feab_create(device, real_bytes, &h_);
        if (rc != FEAB_OK) throw std::runtime_error(std::string("CUDASolver: ") + feab_last_error(nullptr));
    }
    virtual ~CUDASolver() {
        if (h_) feab_destroy(h_);
    }
    CUDASolver(const CUDASolver&) = delete;
    CUDASolver& operator=(const CUDASolver&) = delete;

    // Same role as GPUSolver::init() (reference include/FEM/GPUSolver.hpp:40): call after the surface was replaced.
    void init() {
        mesh_sig_.clear();
        sync_context();
    }
    // GPUSolver::brush(idx, strength) (reference include/FEM/GPUSolver.hpp:41); the CPU path writes surface->values
    // directly (reference src/Application.cpp:722), which also works here because advance_time() uploads the values.
    void brush(int vertex_idx, float strength) {
        if (vertex_idx >= 0 && (size_t)vertex_idx < fem_ctx->surface->values.size()) fem_ctx->surface->values[vertex_idx] = strength;
    }

    void advance_time() override { advance(1); }
    // nsteps > 1 keeps the fields on the device between steps (one upload, one download)
    void advance(int nsteps) {
        if (!fem_ctx->surface || !fem_ctx->surface->initialized) return;
        sync_context();
        push_parameters();
        check(feab_step(h_, fem_ctx->surface->values.data(), nsteps, 1), "feab_step");
    }
    void clear_values() override {  // reference src/FEM/CPUSolver.cpp:54-59
        if (!fem_ctx->surface || !fem_ctx->surface->initialized) return;
        sync_context();
        check(feab_clear(h_), "feab_clear");
    }
    bool has_numerical_instability() override {  // reference src/FEM/CPUSolver.cpp:41-48
        if (mesh_sig_.empty()) return false;
        int f = 0;
        check(feab_unstable(h_, &f), "feab_unstable");
        return f != 0;
    }

    void set_solver_options(const FeabSolverOptions& o) { check(feab_set_solver_options(h_, &o), "feab_set_solver_options"); }
    FeabSolverOptions solver_options() const {
        FeabSolverOptions o;
        feab_get_solver_options(h_, &o);
        return o;
    }
    FeabStats last_stats() {
        FeabStats s;
        check(feab_last_stats(h_, &s), "feab_last_stats");
        return s;
    }
    FeabHandle* handle() { return h_; }

   private:
    FeabHandle* h_ = nullptr;
    const float* reg_values_ = nullptr;   // what feab_register_host_values last page-locked
    size_t reg_count_ = 0;
    std::vector<unsigned long long> mesh_sig_, asm_sig_;

    void check(int rc, const char* what) {
        if (rc != FEAB_OK) throw std::runtime_error(std::string("CUDASolver: ") + what + ": " + feab_last_error(h_));
    }

    template <typename P>
    std::shared_ptr<P> params(Equation eq) {
        return std::static_pointer_cast<P>(fem_ctx->parameters[eq]);
    }

    static unsigned long long bits(float f) {
        uint32_t u;
        std::memcpy(&u, &f, 4);
        return u;
    }

    // Upload the mesh / re-assemble exactly when the reference would (init_from_surface, update_boundary_conditions,
    // assemble_matrices after a velocity change: reference src/Application.cpp:313-316, :393-394, :593-596).
    void sync_context() {
        FEMContext& c = *fem_ctx;
        Surface& s = *c.surface;
        // surface->values is the array every feab_step copies from and to: page-lock it whenever the vector moved or changed
        // size (a surface load); an empty vector undoes the registration
        if (s.values.data() != reg_values_ || s.values.size() != reg_count_) {
            check(feab_register_host_values(h_, s.values.empty() ? nullptr : s.values.data(), (int64_t)s.values.size()), "feab_register_host_values");
            reg_values_ = s.values.data();
            reg_count_ = s.values.size();
        }
        static_assert(sizeof(s.vertices[0]) == 3 * sizeof(float), "Surface::vertices must hold 3 packed floats");
        static_assert(sizeof(s.triangles[0]) == 3 * sizeof(unsigned int), "Triangle must hold 3 packed indices");
        std::vector<unsigned long long> msig = {(unsigned long long)(uintptr_t)&s, s.vertices.size(), s.triangles.size(),
                                                (unsigned long long)s.num_boundary_points};
        msig.push_back(feab_detail::mesh_revision_of(c, 0));
        auto vel = params<AdvectionDiffusionParameters>(Equation::Advection_Diffusion)->velocity;
        const float v3[3] = {vel[0], vel[1], vel[2]};
        std::vector<unsigned long long> asig = {(unsigned long long)c.boundary_condition, bits(v3[0]), bits(v3[1]), bits(v3[2])};
        if (msig != mesh_sig_) {
            std::vector<uint8_t> onb(s.on_boundary.size());
            for (size_t i = 0; i < onb.size(); i++) onb[i] = s.on_boundary[i] ? 1 : 0;
            // boundary condition and velocity travel with the mesh: one pattern build, one assembly per surface load
            check(feab_set_mesh_ex(h_, (int64_t)s.vertices.size(), reinterpret_cast<const float*>(s.vertices.data()), (int64_t)s.triangles.size(),
                                   reinterpret_cast<const uint32_t*>(s.triangles.data()), onb.data(), (int)c.boundary_condition, v3),
                  "feab_set_mesh_ex");
            mesh_sig_ = msig;
            asm_sig_ = asig;
            publish_counts();
            return;
        }
        if (asig != asm_sig_) {
            if (asig[0] != asm_sig_[0]) check(feab_set_boundary_condition(h_, (int)c.boundary_condition), "feab_set_boundary_condition");
            check(feab_assemble(h_, v3), "feab_assemble");
            asm_sig_ = asig;
            publish_counts();
        }
    }
    void publish_counts() {
        int max_row = 0;
        check(feab_get_counts(h_, nullptr, nullptr, nullptr, &max_row), "feab_get_counts");
        feab_detail::set_max_row(*fem_ctx, max_row, 0);
    }

    // parameters are read fresh on every call (reference src/FEM/CPUSolver.cpp:73, :91, :109, :128)
    void push_parameters() {
        FEMContext& c = *fem_ctx;
        FeabParams p;
        std::memset(&p, 0, sizeof(p));
        switch (c.equation) {
            case Equation::Heat: {
                auto q = params<HeatParameters>(Equation::Heat);
                p.time_step = q->time_step;
                p.c[0] = q->conductivity;
            } break;
            case Equation::Wave: {
                auto q = params<WaveParameters>(Equation::Wave);
                p.time_step = q->time_step;
                p.c[0] = q->c;
            } break;
            case Equation::Advection_Diffusion: {
                auto q = params<AdvectionDiffusionParameters>(Equation::Advection_Diffusion);
                p.time_step = q->time_step;
                p.c[0] = q->c;
            } break;
            default: {
                auto q = params<ReactionDiffusionParameters>(Equation::Reaction_Diffusion);
                p.time_step = q->time_step;
                p.c[0] = q->Du;
                p.c[1] = q->Dv;
                p.c[2] = q->feed_rate;
                p.c[3] = q->kill_rate;
            } break;
        }
        check(feab_set_equation(h_, (int)c.equation), "feab_set_equation");
        check(feab_set_params(h_, (int)c.equation, &p), "feab_set_params");
    }
};
