// Headless C++ host program written the way the reference's Application drives a solver (reference
// src/Application.cpp:61-67 construct, :593-596 load surface, :722 brush, :530-540 advance_time + instability check),
// but with CUDASolver (include/FEM/CUDASolver.hpp -> C ABI -> libfeab.so).  Used by tests/test_cpp_host.py:
//   cuda_solver_driver mesh.bin <equation 0..3> <bc 0|1> <nsteps> <real_bytes 4|8> out.bin [export.ply threshold mesh_type]
// With the three optional arguments the final field is also exported the way Application::export_to_ply does (reference
// src/Application.cpp:680-690), through CUDASurface, taking the values from the solver device-to-device.
// mesh.bin: int64 V, int64 T, int64 S, float xyz[3V], uint32 tri[3T], int64 seeds[S]
// out.bin : float values[nsteps][V] (surface->values after every step), then int32 unstable flag per step
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <vector>

#include "FEM/CUDASolver.hpp"
#include "Utils/CUDASurface.hpp"

static void die(const char* m) {
    std::fprintf(stderr, "cuda_solver_driver: %s\n", m);
    std::exit(2);
}

int main(int argc, char** argv) {
    if (argc != 7 && argc != 10) die("usage: mesh.bin equation bc nsteps real_bytes out.bin [export.ply threshold mesh_type]");
    const int eq = std::atoi(argv[2]), bc = std::atoi(argv[3]), nsteps = std::atoi(argv[4]), real_bytes = std::atoi(argv[5]);
    FILE* f = std::fopen(argv[1], "rb");
    if (!f) die("cannot open mesh file");
    int64_t hdr[3];
    if (std::fread(hdr, 8, 3, f) != 3) die("short header");
    const int64_t V = hdr[0], T = hdr[1], S = hdr[2];
    auto surface = std::make_shared<Surface>();
    surface->vertices.resize(V);
    surface->triangles.resize(T);
    std::vector<int64_t> seeds(S);
    if ((int64_t)std::fread(surface->vertices.data(), 12, V, f) != V) die("short vertices");
    if ((int64_t)std::fread(surface->triangles.data(), 12, T, f) != T) die("short triangles");
    if ((int64_t)std::fread(seeds.data(), 8, S, f) != S) die("short seeds");
    std::fclose(f);
    surface->finish_init();  // boundary flags, zero values, initialized = true

    auto fem_ctx = std::make_shared<FEMContext>();
    fem_ctx->boundary_condition = static_cast<BoundaryCondition>(bc);
    fem_ctx->equation = static_cast<Equation>(eq);
    std::shared_ptr<Solver> solver;
    try {
        solver = std::make_shared<CUDASolver>(fem_ctx, 0, real_bytes);
    } catch (const std::exception& e) {
        die(e.what());
    }
    fem_ctx->init_from_surface(surface);
    solver->clear_values();
    surface->clear_values();

    FILE* out = std::fopen(argv[6], "wb");
    if (!out) die("cannot open output");
    std::vector<int32_t> unstable;
    try {
        for (int s = 0; s < nsteps; s++) {
            if (s == 0 || (eq == 3 && s < 5))
                for (int64_t k : seeds) surface->values[k] = 1.0f;  // the brush (reference src/Application.cpp:722)
            solver->advance_time();
            unstable.push_back(solver->has_numerical_instability() ? 1 : 0);
            std::fwrite(surface->values.data(), 4, V, out);
        }
    } catch (const std::exception& e) {
        die(e.what());
    }
    std::fwrite(unstable.data(), 4, unstable.size(), out);
    std::fclose(out);
    if (argc == 10) {
        try {
            const std::vector<bool> host_flags = surface->on_boundary;
            CUDASurface dev(surface);          // normals empty -> (0,1,0), what init_from_PSLG sets
            dev.detect_boundary();             // device pass must agree with the host mirror's finish_init()
            if (surface->on_boundary != host_flags) die("device boundary flags differ from the host mirror");
            dev.export_to_ply_from(std::static_pointer_cast<CUDASolver>(solver)->handle(), argv[7], 0.25f, (float)std::atof(argv[8]),
                                   std::atoi(argv[9]));
        } catch (const std::exception& e) {
            die(e.what());
        }
    }
    std::printf("ok V=%lld unknowns=%u max_row=%u\n", (long long)V, fem_ctx->num_unknowns(), fem_ctx->num_max_nonzeros_per_row());
    return 0;
}
