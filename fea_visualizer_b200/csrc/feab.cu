// libfeab.so -- host driver and C ABI (include/feab.h) around the sm_100a kernels in pattern.cuh,
// assembly.cuh and solver.cuh.  One handle = one mesh on one GPU; everything stays resident in HBM.
// There is no CPU fallback: without a CUDA device feab_create fails with FEAB_ENOGPU.
#include "../../include/feab.h"
#include "../../include/feab_surface.h"

#include <unistd.h>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "assembly.cuh"
#include "bicg_persistent.cuh"
#include "bicgstab.cuh"
#include "common.cuh"
#include "partition.cuh"
#include "pattern.cuh"
#include "pcg_cluster.cuh"
#include "pcg_persistent.cuh"
#include "pcg_dual.cuh"
#include "rhs_tma.cuh"
#include "solver.cuh"
#include "spmv_tma.cuh"

namespace feab {

struct Err {
    int code;
    std::string msg;
};
#define CK(call)                                                                                          \
    do {                                                                                                  \
        cudaError_t e__ = (call);                                                                         \
        if (e__ != cudaSuccess)                                                                           \
            throw Err{(int)e__, std::string(#call) + ": " + cudaGetErrorString(e__) + " (" __FILE__ ":" + \
                                    std::to_string(__LINE__) + ")"};                                      \
    } while (0)
#define REQUIRE(cond, code, text) \
    do {                          \
        if (!(cond)) throw Err{code, text}; \
    } while (0)

template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t count = 0;
    bool owner = true;
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    ~DevBuf() { release(); }
    void release() {
        if (p && owner) cudaFree(p);
        p = nullptr;
        count = 0;
        owner = true;
    }
    void view(void* ptr, size_t n) {  // non-owning window into another allocation (the comm arena)
        release();
        p = static_cast<T*>(ptr);
        count = n;
        owner = false;
    }
    void alloc(size_t n) {
        if (n == count && p) return;
        release();
        if (n == 0) n = 1;
        // kTmaPad spare elements: 1-D bulk TMA copies round their sizes up to 16 bytes
        cudaError_t e = cudaMalloc(&p, (n + kTmaPad) * sizeof(T));
        if (e != cudaSuccess) throw Err{FEAB_ENOMEM, std::string("cudaMalloc of ") + std::to_string(n * sizeof(T)) + " bytes: " + cudaGetErrorString(e)};
        count = n;
    }
};

static inline int cdiv64(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }


struct SolverBase {
    virtual ~SolverBase() {}
    virtual void set_mesh(int64_t V, const float* xyz, int64_t T, const uint32_t* tri, const uint8_t* onb) = 0;
    virtual void set_mesh_ex(int64_t V, const float* xyz, int64_t T, const uint32_t* tri, const uint8_t* onb, int bc, const float* vel) = 0;
    virtual void set_bc(int bc) = 0;
    virtual void assemble(const float* vel) = 0;
    virtual void get_counts(int64_t* V, int64_t* n, int64_t* nnz, int* maxrow) = 0;
    virtual void get_idx_map(int32_t* out) = 0;
    virtual void get_csr(int which, int64_t* nnz, int32_t* rowptr, int32_t* col, void* val) = 0;
    virtual void set_equation(int eq) = 0;
    virtual void set_params(int eq, const FeabParams* p) = 0;
    virtual void get_params(int eq, FeabParams* p) = 0;
    virtual void set_options(const FeabSolverOptions* o) = 0;
    virtual void get_options(FeabSolverOptions* o) = 0;
    virtual void clear() = 0;
    virtual void set_values(const float* v) = 0;
    virtual void get_values(float* v) = 0;
    virtual const float* values_device() = 0;
    virtual void brush(const int64_t* nodes, int64_t count, float strength) = 0;
    virtual void step(float* values, int nsteps, int sync) = 0;
    virtual int unstable() = 0;
    virtual void stats(FeabStats* s) = 0;
    virtual void get_state(int which, void* out) = 0;
    virtual void set_state(int which, const void* in) = 0;
    virtual void time_kernel(int kernel, int reps, double* ms, double* bytes) = 0;
    virtual void pcg_profile(double* out16, int reset) = 0;
    virtual void mg_config(int rank, int world) = 0;
    virtual void mg_set_partition(const int32_t* part, int64_t V) = 0;
    virtual void mg_export(FeabMgExport* out) = 0;
    virtual void mg_connect(int peer, const FeabMgExport* e, int same_process) = 0;
    virtual void mg_get_layout(int64_t* n_own, int64_t* n_int, int64_t* n_ghost, int64_t* n_cols) = 0;
    virtual void mg_get_maps(int32_t* row_node, int32_t* col_unknown) = 0;
    virtual void mg_get_partition(int32_t* part) = 0;
    // lock-step driving (feab_mg_group_step)
    enum { LS_PUSH_STATE, LS_BEGIN, LS_COLLECT, LS_CG_BEGIN, LS_SPMV, LS_UPDATE, LS_DIRECTION, LS_FINISH };
    virtual cudaStream_t ls_prepare(cudaStream_t shared) = 0;
    virtual void ls_phase(int phase, int slot) = 0;
    virtual int ls_num_solves() = 0;
    virtual int ls_poll_done(int slot) = 0;
    virtual int64_t ls_maxit() = 0;
    virtual void ls_finish_call(int nsteps) = 0;
    virtual int device() const = 0;
    virtual const char* last_solve_kernel() const = 0;
};

template <typename R>
class SolverT final : public SolverBase {
   public:
    explicit SolverT(int device) : device_(device) {
        CK(cudaSetDevice(device_));
        CK(cudaGetDeviceProperties(&prop_, device_));
        CK(cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&stream2_, cudaStreamNonBlocking));
        CK(cudaEventCreate(&ev0_));
        CK(cudaEventCreate(&ev1_));
        CK(cudaMallocHost(&h_flag_, 64));
        std::memset(h_flag_, 0, 64);  // [8]: pcg_cluster's inter-cluster barrier gave up (written by the device)
        sms_ = prop_.multiProcessorCount;
        // opt in to > 48 KB dynamic shared memory for the tile kernels
        const int sm1 = kTileCap * (int)sizeof(R), sm2 = 2 * kTileCap * (int)sizeof(R);
        CK(cudaFuncSetAttribute(spmv_dot<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm1));
        CK(cudaFuncSetAttribute(rhs_init<R, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm2));
        CK(cudaFuncSetAttribute(rhs_init<R, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm2));
        CK(cudaFuncSetAttribute(rhs_init<R, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm2));
        CK(cudaFuncSetAttribute(rhs_init<R, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm2));
        CK(cudaFuncSetAttribute(spmv_plain<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm1));
        // include/FEM/EquationParameters.hpp defaults
        par_[FEAB_EQ_HEAT] = {0.01f, {0.05f, 0, 0, 0}};
        par_[FEAB_EQ_WAVE] = {0.05f, {0.05f, 0, 0, 0}};
        par_[FEAB_EQ_ADVECTION_DIFFUSION] = {0.001f, {0.25f, 0, 0, 0}};
        par_[FEAB_EQ_REACTION_DIFFUSION] = {0.001f, {0.08f, 0.04f, 0.035f, 0.06f}};
        opt_.solver = FEAB_SOLVER_PARITY_CG;
        opt_.tolerance = -1.0;
        opt_.max_iter = -1;
        opt_.loop_mode = FEAB_LOOP_GRAPH_WHILE;
        opt_.assembly = FEAB_ASSEMBLY_SCATTER;
        CK(cudaMalloc(&d_unstable_, sizeof(int)));
        CK(cudaMemset(d_unstable_, 0, sizeof(int)));
        CK(cudaMalloc(&d_xbar_, sizeof(unsigned long long)));
        CK(cudaMemset(d_xbar_, 0, sizeof(unsigned long long)));
        CK(cudaMalloc(&d_st_, 2 * sizeof(CgState<R>)));
        CK(cudaMemset(d_st_, 0, 2 * sizeof(CgState<R>)));
        CK(cudaMalloc(&d_bst_, sizeof(BicgState<R>)));
        CK(cudaMemset(d_bst_, 0, sizeof(BicgState<R>)));
        d_scan_total_.alloc(1);
        d_rhs_tot_.alloc(8);
        CK(cudaMalloc(&d_loc_, sizeof(CommLocal)));
        CK(cudaMemset(d_loc_, 0, sizeof(CommLocal)));
        cm_.loc = d_loc_;
        if (const char* v = getenv("FEAB_SPMV")) spmv_variant_ = atoi(v);
        if (const char* v = getenv("FEAB_PREFETCH_X")) prefetch_x_ = atoi(v);
        if (const char* v = getenv("FEAB_RHS_TMA")) rhs_tma_ = atoi(v) != 0;
        {
            using RCfg = TmaCfg<R, kRhsRows, kRhsCpr, kRhsStages>;
            CK(cudaFuncSetAttribute(rhs_init_tma<R, 0, kRhsRows, kRhsCpr, kRhsStages, kRhsMinB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)RCfg::kSmemBytes));
            CK(cudaFuncSetAttribute(rhs_init_tma<R, 2, kRhsRows, kRhsCpr, kRhsStages, kRhsMinB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)RCfg::kSmemBytes));
            CK(cudaFuncSetAttribute(rhs_init_tma<R, 3, kRhsRows, kRhsCpr, kRhsStages, kRhsMinB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)RCfg::kSmemBytes));
        }
        setup_spmv_variants();
        // persistent whole-solve kernel (pcg_persistent.cuh): same tile geometry as SpMV variant 4
        using PCfg = TmaCfg<R, kPcgRows, kPcgCpr, kPcgStages>;
        CK(cudaFuncSetAttribute(pcg_persistent<R, kPcgRows, kPcgCpr, kPcgStages, kPcgMinB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)PCfg::kSmemBytes));
        int occ = 0;
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, pcg_persistent<R, kPcgRows, kPcgCpr, kPcgStages, kPcgMinB>, kPcgRows,
                                                         PCfg::kSmemBytes));
        int coop = 0;
        CK(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, device_));
        pcg_max_grid_ = coop ? occ * sms_ : 0;
        if (const char* v = getenv("FEAB_PCG_CTAS_PER_SM")) pcg_max_grid_ = std::min(pcg_max_grid_, std::max(1, atoi(v)) * sms_);
        for (int i = 0; i < 4; i++) CK(cudaEventCreate(&ev_solve_[i]));
        if (const char* v = getenv("FEAB_STEP_PROF")) {
            step_prof_ = atoi(v) != 0;
            for (int i = 0; i < 8; i++) CK(cudaEventCreate(&ev_ph_[i]));
        }
        if (const char* v = getenv("FEAB_PCG_PROF")) {  // 1: stamps of pcg_persistent, 2: stamps of pcg_dual, 3: + every CTA's sync arrival times
            prof_dual_ = atoi(v) >= 2;
            prof_arrivals_ = atoi(v) == 3;
            CK(cudaMalloc(&d_prof_, 16 * sizeof(unsigned long long)));
            CK(cudaMemset(d_prof_, 0, 16 * sizeof(unsigned long long)));
        }
        if (pcg_max_grid_ > 0) {
            opt_.loop_mode = FEAB_LOOP_PERSISTENT;
            d_ppart_.alloc((size_t)4 * std::max(pcg_max_grid_, 1024));  // pcg_persistent: 3 x grid; bicg_persistent: 4 x grid
        }
        // Gray-Scott: both solves of a step in lock-step from one index stream (pcg_dual.cuh); FEAB_DUAL=0/1 overrides
        using DCfg = DualCfg<R, kDualRows, kPcgCpr, kPcgStages>;
        CK(cudaFuncSetAttribute(pcg_dual<R, kDualRows, kPcgCpr, kPcgStages, kPcgMinB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DCfg::kSmemBytes));
        int docc = 0;
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&docc, pcg_dual<R, kDualRows, kPcgCpr, kPcgStages, kPcgMinB>, 2 * kDualRows, DCfg::kSmemBytes));
        dual_max_grid_ = coop ? docc * sms_ : 0;
        if (dual_max_grid_ > 0) d_dpart_.alloc((size_t)6 * dual_max_grid_);
        if (const char* v = getenv("FEAB_DUAL")) dual_ = atoi(v) != 0;
        if (const char* v = getenv("FEAB_ZIGZAG")) zigzag_ = std::max(0, std::min(2, atoi(v)));
        if (const char* v = getenv("FEAB_COOP")) coop_launch_ = atoi(v) != 0;
        if (const char* v = getenv("FEAB_FAULT_GSYNC")) fault_gsync_call_ = atoi(v);  // test hook, see step()
        // non-symmetric mode: the whole BiCGSTAB solve as one cooperative kernel (bicg_persistent.cuh), same tile geometry
        CK(cudaFuncSetAttribute(bicg_persistent<R, kPcgRows, kPcgCpr, kPcgStages, kPcgMinB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)PCfg::kSmemBytes));
        int bocc = 0;
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bocc, bicg_persistent<R, kPcgRows, kPcgCpr, kPcgStages, kPcgMinB>, kPcgRows,
                                                         PCfg::kSmemBytes));
        bicg_max_grid_ = coop ? bocc * sms_ : 0;
        // small meshes: whole time steps inside one thread-block cluster (pcg_cluster.cuh); FEAB_CLUSTER=0 switches it off
        {
            const int smem = kClusterSmemBudget;
            if (const char* v = getenv("FEAB_CLUSTER_THREADS")) cluster_threads_ = atoi(v) == 1024 ? 1024 : 512;
            bool ok = true;
            for (const void* f : {(const void*)pcg_cluster<R, 1024>, (const void*)pcg_cluster<R, 512>}) {
                ok = ok && cudaFuncSetAttribute(f, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) == cudaSuccess;
                ok = ok && cudaFuncSetAttribute(f, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess;
            }
            for (int cs = kClusterMaxCtas; ok && cs >= 2 && cluster_ctas_ == 0; cs /= 2) {
                cudaLaunchConfig_t cfg = {};
                cfg.gridDim = dim3(cs);
                cfg.blockDim = dim3(1024);
                cfg.dynamicSmemBytes = smem;
                cudaLaunchAttribute at[1];
                at[0].id = cudaLaunchAttributeClusterDimension;
                at[0].val.clusterDim.x = cs;
                at[0].val.clusterDim.y = 1;
                at[0].val.clusterDim.z = 1;
                cfg.attrs = at;
                cfg.numAttrs = 1;
                int nclusters = 0;
                if (cudaOccupancyMaxActiveClusters(&nclusters, pcg_cluster<R, 1024>, &cfg) == cudaSuccess && nclusters >= 1) {
                    cluster_ctas_ = cs;
                    cluster_pair_ = nclusters >= 2;  // Gray-Scott: one cluster per system, side by side (they meet at a spin barrier: both must be resident)
                }
            }
            if (const char* v = getenv("FEAB_CLUSTER_PAIR")) cluster_pair_ = cluster_pair_ && atoi(v) != 0;
            cudaGetLastError();
            if (const char* v = getenv("FEAB_CLUSTER"))
                if (atoi(v) == 0) cluster_ctas_ = 0;
        }
        // CUDA loads kernels lazily, on their first launch, and that load may wait for the device to drain.  A partitioned
        // step launches kernels while kernels of other ranks SPIN on flags this rank has yet to raise; if the launch that
        // raises them is a first launch, the load and the spin wait for each other (seen with several ranks in one process).
        // Touch every kernel of the stepping path once, here, while nothing is running.
        preload_kernels();
        // several handles whose persistent kernels must be co-resident on ONE device (the threaded virtual-rank tests)
        if (const char* v = getenv("FEAB_PCG_MAX_GRID")) {
            const int g = std::max(1, atoi(v));
            pcg_max_grid_ = std::min(pcg_max_grid_, g);
            dual_max_grid_ = std::min(dual_max_grid_, g);
            bicg_max_grid_ = std::min(bicg_max_grid_, g);
            grid_cap_ = g;  // also for the RHS kernel, whose CTAs wait for the neighbours' state flags
        }
    }
    static constexpr int kPcgRows = 512, kPcgCpr = 8, kPcgStages = 2, kPcgMinB = 2;
    static constexpr int kDualRows = 256;

    template <typename F>
    static void touch(F* f) {
        cudaFuncAttributes at;
        cudaFuncGetAttributes(&at, (const void*)f);
        cudaGetLastError();
    }
    void preload_kernels() {
        touch(wait_kernel);
        touch(root_signal);
        touch(root_wait);
        touch(values_broadcast);
        touch(pull_cols);
        touch(brush_nodes);
        touch(gather_values<R>);
        touch(gather_cols<R>);
        touch(finish_step<R>);
        touch(finish_rows<R>);
        touch(state_push<R>);
        touch(halo_push<R>);
        touch(build_operator<R>);
        touch(inv_diag<R>);
        touch(rhs_init<R, 0>);
        touch(rhs_init<R, 1>);
        touch(rhs_init<R, 2>);
        touch(rhs_init<R, 3>);
        touch(rhs_init_tma<R, 0, kRhsRows, kRhsCpr, kRhsStages, kRhsMinB>);
        touch(rhs_init_tma<R, 2, kRhsRows, kRhsCpr, kRhsStages, kRhsMinB>);
        touch(rhs_init_tma<R, 3, kRhsRows, kRhsCpr, kRhsStages, kRhsMinB>);
        touch(ar_collect<R, 2>);
        touch(ar_collect<R, 4>);
        touch(cg_begin<R>);
        touch(cg_update<R>);
        touch(cg_direction<R>);
        touch(spmv_dot<R>);
        touch(spmv_dot_tma<R, 512, 8, 2, 2>);
        touch(pcg_persistent<R, kPcgRows, kPcgCpr, kPcgStages, kPcgMinB>);
        touch(pcg_dual<R, kDualRows, kPcgCpr, kPcgStages, kPcgMinB>);
        touch(bicg_persistent<R, kPcgRows, kPcgCpr, kPcgStages, kPcgMinB>);
        touch(pcg_cluster<R, 1024>);
        touch(pcg_cluster<R, 512>);
    }

    // ---- SpMV variants: 0 = LDG/CSR-stream (solver.cuh), 1.. = TMA pipelines (spmv_tma.cuh) ----
    template <int ROWS, int CPR, int STAGES, int MINB>
    void setup_tma() {
        using Cfg = TmaCfg<R, ROWS, CPR, STAGES>;
        CK(cudaFuncSetAttribute(spmv_dot_tma<R, ROWS, CPR, STAGES, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::kSmemBytes));
    }
    void setup_spmv_variants() {
        setup_tma<256, 10, 3, 2>();
        setup_tma<512, 9, 3, 1>();
        setup_tma<256, 10, 2, 3>();
        setup_tma<512, 8, 2, 2>();
        setup_tma<1024, 8, 2, 1>();
    }
    template <int ROWS, int CPR, int STAGES, int MINB>
    void launch_tma(cudaStream_t s, const R* A, const R* x, R* y, CgState<R>* st) {
        using Cfg = TmaCfg<R, ROWS, CPR, STAGES>;
        const int tiles = std::max(1, cdiv64(n_, ROWS));
        const int grid = std::min(tiles, sms_ * MINB);
        spmv_dot_tma<R, ROWS, CPR, STAGES, MINB><<<grid, ROWS, Cfg::kSmemBytes, s>>>(d_rowptr_.p, d_col_.p, A, x, y, n_, tiles, st, part_pap(), prefetch_x_,
                                                                                     cm_, hp_, first_halo_tile(ROWS));
    }
    template <int ROWS, int MINB>
    int grid_tma() const {
        return std::min(std::max(1, cdiv64(n_, ROWS)), sms_ * MINB);
    }
    // number of p.Ap partials the SpMV of the current variant writes
    // the LDG variant has no halo wait: a partitioned handle always runs a TMA pipeline
    int spmv_variant() const { return (world_ > 1 && spmv_variant_ == 0) ? 4 : spmv_variant_; }
    int spmv_parts() const {
        switch (spmv_variant()) {
            case 1: return grid_tma<256, 2>();
            case 2: return grid_tma<512, 1>();
            case 3: return grid_tma<256, 3>();
            case 4: return grid_tma<512, 2>();
            case 5: return grid_tma<1024, 1>();
            default: return g_tile_;
        }
    }
    void launch_spmv(cudaStream_t s, const R* A, const R* x, R* y, CgState<R>* st) {
        switch (spmv_variant()) {
            case 1: launch_tma<256, 10, 3, 2>(s, A, x, y, st); break;
            case 2: launch_tma<512, 9, 3, 1>(s, A, x, y, st); break;
            case 3: launch_tma<256, 10, 2, 3>(s, A, x, y, st); break;
            case 4: launch_tma<512, 8, 2, 2>(s, A, x, y, st); break;
            case 5: launch_tma<1024, 8, 2, 1>(s, A, x, y, st); break;
            default:
                spmv_dot<R><<<g_tile_, kBlock, kTileCap * sizeof(R), s>>>(d_rowptr_.p, d_col_.p, A, x, y, n_, num_tiles_, st, part_pap());
                break;
        }
    }
    ~SolverT() override {
        cudaSetDevice(device_);
        cudaDeviceSynchronize();
        drop_graphs();
        if (d_unstable_) cudaFree(d_unstable_);
        if (d_xbar_) cudaFree(d_xbar_);
        if (d_st_) cudaFree(d_st_);
        if (d_bst_) cudaFree(d_bst_);
        if (d_loc_) cudaFree(d_loc_);
        if (d_prof_) cudaFree(d_prof_);
        for (int i = 0; i < 4; i++)
            if (ev_solve_[i]) cudaEventDestroy(ev_solve_[i]);
        for (int i = 0; i < 8; i++)
            if (ev_ph_[i]) cudaEventDestroy(ev_ph_[i]);
        close_peers();
        if (h_flag_) cudaFreeHost(h_flag_);
        cudaEventDestroy(ev0_);
        cudaEventDestroy(ev1_);
        cudaStreamDestroy(stream_);
        cudaStreamDestroy(stream2_);
    }

    // ---------------------------------------------------------------- mesh / symbolic / assembly
    void set_mesh(int64_t V, const float* xyz, int64_t T, const uint32_t* tri, const uint8_t* onb) override {
        CK(cudaSetDevice(device_));
        REQUIRE(V > 0 && T >= 0 && xyz && onb && (T == 0 || tri), FEAB_EINVAL, "set_mesh: bad arguments");
        REQUIRE(V < (int64_t)INT32_MAX && 9 * T < (int64_t)INT32_MAX, FEAB_EINVAL, "set_mesh: mesh too large for int32 indexing");
        if (world_ > 1 && part_given_) REQUIRE((int64_t)h_part_.size() == V, FEAB_EINVAL, "set_mesh: partition array length != num_vertices");
        V_ = V;
        T_ = T;
        d_xyz_.alloc(3 * V);
        d_tri_.alloc(3 * std::max<int64_t>(T, 1));
        d_onb_.alloc(V);
        CK(cudaMemcpyAsync(d_xyz_.p, xyz, 3 * V * sizeof(float), cudaMemcpyHostToDevice, stream_));
        if (T) CK(cudaMemcpyAsync(d_tri_.p, tri, 3 * T * sizeof(uint32_t), cudaMemcpyHostToDevice, stream_));
        CK(cudaMemcpyAsync(d_onb_.p, onb, V, cudaMemcpyHostToDevice, stream_));
        // validate the triangle indices (the reference would index out of bounds) on the device copy: no host pass over 3 T indices
        CK(cudaMemsetAsync(d_unstable_, 0, sizeof(int), stream_));
        if (T) validate_indices<<<std::max(1, std::min(cdiv64(3 * T, kBlock), sms_ * 8)), kBlock, 0, stream_>>>(d_tri_.p, 3 * T, (uint32_t)V, d_unstable_);
        int bad_index = 0;
        CK(cudaMemcpyAsync(&bad_index, d_unstable_, sizeof(int), cudaMemcpyDeviceToHost, stream_));
        CK(cudaStreamSynchronize(stream_));
        if (bad_index) {
            have_mesh_ = false;
            CK(cudaMemsetAsync(d_unstable_, 0, sizeof(int), stream_));
            throw Err{FEAB_EINVAL, "set_mesh: triangle index out of range"};
        }
        if (world_ > 1) {
            d_part_of_.alloc(V);
            if (part_given_)
                CK(cudaMemcpyAsync(d_part_of_.p, h_part_.data(), V * sizeof(int), cudaMemcpyHostToDevice, stream_));
            else  // default partition: recursive coordinate bisection on the device, identical on every rank (partition.cuh)
                rcb_device(world_);
        }
        CK(cudaStreamSynchronize(stream_));
        have_mesh_ = true;
        symbolic();
        assemble(nullptr);
        clear();
    }

    // d_part_of_ = recursive coordinate bisection of d_xyz_ into `parts` pieces; == rcb_partition() on the host, bit for bit
    void rcb_device(int parts) {
        const int64_t V = V_;
        struct Seg {
            int label, parts;
            long long count;
        };
        std::vector<Seg> segs = {{0, parts, (long long)V}};
        CK(cudaMemsetAsync(d_part_of_.p, 0, V * sizeof(int), stream_));
        DevBuf<unsigned> box, hist;
        DevBuf<int> flag, rank;
        box.alloc(kMaxRanks * 6);
        hist.alloc(kMaxRanks / 2 * 256);
        flag.alloc(V);
        rank.alloc(V + 1);
        const int gS = std::max(1, std::min(cdiv64(V, kBlock), sms_ * 8)), gV = cdiv64(V, kBlock);
        while (true) {
            RcbLevel lv;
            std::memset(&lv, 0, sizeof(lv));
            for (int q = 0; q < kMaxRanks; q++) lv.seg_of_label[q] = -1;
            std::vector<int> act;
            for (int i = 0; i < (int)segs.size() && (int)act.size() < kMaxRanks / 2; i++)
                if (segs[i].parts > 1) act.push_back(i);
            if (act.empty()) break;
            lv.nseg = (int)act.size();
            long long k[kMaxRanks] = {}, nl[kMaxRanks] = {};
            for (int a = 0; a < lv.nseg; a++) {
                const Seg& sg = segs[act[a]];
                lv.label[a] = sg.label;
                lv.seg_of_label[sg.label] = a;
                lv.right_label[a] = sg.label + sg.parts / 2;
                nl[a] = sg.count * (sg.parts / 2) / sg.parts;
                k[a] = nl[a] - 1;
            }
            // widest axis of every segment (ties: the lowest axis, like the host's strict comparison)
            unsigned hbox[kMaxRanks * 6];
            for (int i = 0; i < lv.nseg * 6; i++) hbox[i] = (i % 6) < 3 ? 0xffffffffu : 0u;
            CK(cudaMemcpyAsync(box.p, hbox, lv.nseg * 6 * sizeof(unsigned), cudaMemcpyHostToDevice, stream_));
            rcb_bbox<<<gS, kBlock, 0, stream_>>>(d_xyz_.p, d_part_of_.p, V, lv, box.p);
            CK(cudaMemcpyAsync(hbox, box.p, lv.nseg * 6 * sizeof(unsigned), cudaMemcpyDeviceToHost, stream_));
            CK(cudaStreamSynchronize(stream_));
            for (int a = 0; a < lv.nseg; a++) {
                int axis = 0;
                float ext[3];
                for (int d = 0; d < 3; d++) ext[d] = rcb_unkey(hbox[a * 6 + 3 + d]) - rcb_unkey(hbox[a * 6 + d]);
                for (int d = 1; d < 3; d++)
                    if (ext[d] > ext[axis]) axis = d;
                lv.axis[a] = axis;
            }
            // radix select of the (nl - 1)-th smallest coordinate key of every segment, 8 bits per pass
            for (int shift = 24; shift >= 0; shift -= 8) {
                CK(cudaMemsetAsync(hist.p, 0, lv.nseg * 256 * sizeof(unsigned), stream_));
                rcb_hist<<<gS, kBlock, 0, stream_>>>(d_xyz_.p, d_part_of_.p, V, lv, shift, hist.p);
                unsigned hh[kMaxRanks / 2 * 256];
                CK(cudaMemcpyAsync(hh, hist.p, lv.nseg * 256 * sizeof(unsigned), cudaMemcpyDeviceToHost, stream_));
                CK(cudaStreamSynchronize(stream_));
                for (int a = 0; a < lv.nseg; a++) {
                    if (nl[a] <= 0) continue;
                    long long cum = 0;
                    int d = 0;
                    for (; d < 255; d++) {
                        if (cum + hh[a * 256 + d] > k[a]) break;
                        cum += hh[a * 256 + d];
                    }
                    k[a] -= cum;
                    lv.prefix[a] |= (unsigned)d << shift;
                }
            }
            for (int a = 0; a < lv.nseg; a++) {
                lv.key[a] = nl[a] > 0 ? lv.prefix[a] : 0u;
                lv.take[a] = nl[a] > 0 ? (int)(k[a] + 1) : 0;   // ties at the selected coordinate that stay on the left, by vertex index
            }
            for (int a = 0; a < lv.nseg; a++) {
                rcb_tie_flags<<<gV, kBlock, 0, stream_>>>(d_xyz_.p, d_part_of_.p, V, lv, a, flag.p);
                exclusive_scan(flag.p, V, rank.p, nullptr);
                rcb_assign<<<gV, kBlock, 0, stream_>>>(d_xyz_.p, d_part_of_.p, V, lv, a, rank.p);
            }
            launches_ += 6 + 2 * lv.nseg;
            CK(cudaGetLastError());
            for (int a = 0; a < lv.nseg; a++) {
                const Seg sg = segs[act[a]];
                const int pl = sg.parts / 2;
                segs[act[a]] = {sg.label, pl, nl[a]};
                segs.push_back({sg.label + pl, sg.parts - pl, sg.count - nl[a]});
            }
        }
    }

    void set_mesh_ex(int64_t V, const float* xyz, int64_t T, const uint32_t* tri, const uint8_t* onb, int bc, const float* vel) override {
        REQUIRE(bc == 0 || bc == 1, FEAB_EINVAL, "set_mesh_ex: bc must be 0 or 1");
        bc_ = bc;  // taken by the symbolic pass / the assembly that set_mesh runs
        if (vel) std::memcpy(vel_, vel, 3 * sizeof(float));
        set_mesh(V, xyz, T, tri, onb);
    }

    void set_bc(int bc) override {
        REQUIRE(bc == 0 || bc == 1, FEAB_EINVAL, "set_boundary_condition: bc must be 0 or 1");
        bc_ = bc;
        if (!have_mesh_) return;
        CK(cudaSetDevice(device_));
        // update_boundary_conditions(): new idx_map, then assemble_matrices() (FEMContext.cpp:224-240).
        // Like the reference this leaves the solver vectors to the caller (clear_values()).
        symbolic();
        assemble(nullptr);
        resize_state();
    }

    void exclusive_scan(const int* in, int64_t n, int* out, long long* total_host) {
        int nb = std::max(1, cdiv64(n, kScanTile));
        d_scan_bsum_.alloc(nb);
        scan_block_sums<<<nb, 1024, 0, stream_>>>(in, n, d_scan_bsum_.p);
        scan_block_offsets<<<1, 1024, 0, stream_>>>(d_scan_bsum_.p, nb, d_scan_total_.p);
        scan_apply<<<nb, 1024, 0, stream_>>>(in, n, d_scan_bsum_.p, d_scan_total_.p, out);
        launches_ += 3;
        CK(cudaGetLastError());
        if (total_host) {
            CK(cudaMemcpyAsync(total_host, d_scan_total_.p, sizeof(long long), cudaMemcpyDeviceToHost, stream_));
            CK(cudaStreamSynchronize(stream_));
        }
    }

    // Local numbering of a partition (partition.cuh): rows, ghost columns, send lists.  Everything is derived from
    // the shared part[] array, identically on every rank.
    void number_partition() {
        const int64_t V = V_, T = T_, ng = nglob_;
        const int gV = cdiv64(V, kBlock), gT = std::max(1, cdiv64(T, kBlock)), gU = std::max(1, cdiv64(ng, kBlock));
        DevBuf<unsigned> mask;
        DevBuf<int> flag, excl, lrow, lcol, cnt;
        mask.alloc(std::max<int64_t>(ng, 1));
        flag.alloc(ng + 1);
        excl.alloc(ng + 2);
        lrow.alloc(std::max<int64_t>(ng, 1));
        lcol.alloc(std::max<int64_t>(ng, 1));
        cnt.alloc(kMaxRanks * kMaxRanks);
        CK(cudaMemsetAsync(mask.p, 0, std::max<int64_t>(ng, 1) * sizeof(unsigned), stream_));
        CK(cudaMemsetAsync(lrow.p, 0xff, std::max<int64_t>(ng, 1) * sizeof(int), stream_));
        CK(cudaMemsetAsync(cnt.p, 0, kMaxRanks * kMaxRanks * sizeof(int), stream_));
        if (T) ghost_mask_build<<<gT, kBlock, 0, stream_>>>(d_tri_.p, T, d_idx_.p, d_part_of_.p, mask.p);
        long long tot = 0;
        // rows: interior first, boundary last
        flag_unknowns<<<gU, kBlock, 0, stream_>>>(d_node_of_.p, d_part_of_.p, mask.p, ng, rank_, 0, 0, flag.p);
        exclusive_scan(flag.p, ng, excl.p, &tot);
        n_int_ = tot;
        assign_ids<<<gU, kBlock, 0, stream_>>>(flag.p, excl.p, ng, 0, lrow.p);
        flag_unknowns<<<gU, kBlock, 0, stream_>>>(d_node_of_.p, d_part_of_.p, mask.p, ng, rank_, 1, 0, flag.p);
        exclusive_scan(flag.p, ng, excl.p, &tot);
        assign_ids<<<gU, kBlock, 0, stream_>>>(flag.p, excl.p, ng, (int)n_int_, lrow.p);
        n_ = n_int_ + tot;
        const int64_t line = 128 / (int64_t)sizeof(R);  // ghosts start on their own 128-byte line (comm.cuh)
        ghost_base_ = (n_ + line - 1) / line * line;
        CK(cudaMemcpyAsync(lcol.p, lrow.p, std::max<int64_t>(ng, 1) * sizeof(int), cudaMemcpyDeviceToDevice, stream_));
        // unknowns owned per rank (for the ghost base of every peer) and the ghost-pair counts
        int64_t n_own_of[kMaxRanks] = {};
        for (int q = 0; q < world_; q++) {
            flag_unknowns<<<gU, kBlock, 0, stream_>>>(d_node_of_.p, d_part_of_.p, mask.p, ng, q, 3, 0, flag.p);
            exclusive_scan(flag.p, ng, excl.p, &tot);
            n_own_of[q] = tot;
        }
        count_ghost_pairs<<<std::min(gU, sms_ * 8), kBlock, 0, stream_>>>(d_node_of_.p, d_part_of_.p, mask.p, ng, cnt.p);
        int h_cnt[kMaxRanks * kMaxRanks];
        CK(cudaMemcpyAsync(h_cnt, cnt.p, sizeof(h_cnt), cudaMemcpyDeviceToHost, stream_));
        CK(cudaStreamSynchronize(stream_));
        // ghosts grouped by owner
        int64_t goff = ghost_base_;
        for (int q = 0; q < world_; q++) {
            if (q == rank_ || h_cnt[q * kMaxRanks + rank_] == 0) continue;
            flag_unknowns<<<gU, kBlock, 0, stream_>>>(d_node_of_.p, d_part_of_.p, mask.p, ng, q, 2, rank_, flag.p);
            exclusive_scan(flag.p, ng, excl.p, &tot);
            REQUIRE(tot == h_cnt[q * kMaxRanks + rank_], FEAB_EINVAL, "partition: ghost count mismatch");
            assign_ids<<<gU, kBlock, 0, stream_>>>(flag.p, excl.p, ng, (int)goff, lcol.p);
            goff += tot;
        }
        n_cols_ = goff;
        n_ghost_ = goff - ghost_base_;
        // send lists and where they land on each neighbour
        hp_ = HaloPlan();
        int64_t total_send = 0;
        for (int q = 0; q < world_; q++)
            if (q != rank_) total_send += h_cnt[rank_ * kMaxRanks + q];
        d_send_rows_.alloc(std::max<int64_t>(total_send, 1));
        d_send_nodes_.alloc(std::max<int64_t>(total_send, 1));
        int64_t soff = 0;
        for (int q = 0; q < world_; q++) {
            const int c = h_cnt[rank_ * kMaxRanks + q];
            if (q == rank_ || c == 0) continue;
            flag_unknowns<<<gU, kBlock, 0, stream_>>>(d_node_of_.p, d_part_of_.p, mask.p, ng, rank_, 2, q, flag.p);
            exclusive_scan(flag.p, ng, excl.p, &tot);
            REQUIRE(tot == c, FEAB_EINVAL, "partition: send count mismatch");
            gather_list<<<gU, kBlock, 0, stream_>>>(flag.p, excl.p, ng, lrow.p, d_node_of_.p, d_send_rows_.p + soff, d_send_nodes_.p + soff);
            const int b = hp_.nnb++;
            hp_.nb_rank[b] = q;
            hp_.send_off[b] = (int)soff;
            const int64_t gb_q = (n_own_of[q] + line - 1) / line * line;
            int64_t before = 0;
            for (int o = 0; o < rank_; o++)
                if (o != q) before += h_cnt[o * kMaxRanks + q];
            hp_.dst_col[b] = (int)(gb_q + before);
            soff += c;
            hp_.send_off[b + 1] = (int)soff;
        }
        hp_.send_rows = d_send_rows_.p;
        hp_.send_nodes = d_send_nodes_.p;
        hp_.n_int = (int)n_int_;
        {
            const size_t cells = (size_t)std::max<int64_t>((n_ - n_int_) * std::max(hp_.nnb, 1), 1);
            d_push_tab_.alloc(cells);
            CK(cudaMemsetAsync(d_push_tab_.p, 0xff, cells * sizeof(int), stream_));
            if (total_send > 0) build_push_table<<<std::max(1, cdiv64(total_send, kBlock)), kBlock, 0, stream_>>>(hp_, d_push_tab_.p);
            hp_.push_tab = d_push_tab_.p;
        }
        // per-node maps and their inverses
        d_rowmap_.alloc(V);
        d_colmap_.alloc(V);
        d_node_of_col_.alloc(std::max<int64_t>(n_cols_, 1));
        d_glob_of_col_.alloc(std::max<int64_t>(n_cols_, 1));
        CK(cudaMemsetAsync(d_node_of_col_.p, 0xff, std::max<int64_t>(n_cols_, 1) * sizeof(int), stream_));
        CK(cudaMemsetAsync(d_glob_of_col_.p, 0xff, std::max<int64_t>(n_cols_, 1) * sizeof(int), stream_));
        make_local_maps<<<gV, kBlock, 0, stream_>>>(d_idx_.p, lrow.p, lcol.p, V, d_rowmap_.p, d_colmap_.p, d_node_of_col_.p, d_glob_of_col_.p);
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(stream_));
        rowmap_p_ = d_rowmap_.p;
        colmap_p_ = d_colmap_.p;
        node_of_col_p_ = d_node_of_col_.p;
    }

    void symbolic() {
        drop_graphs();
        cluster_ent_key_ = -1;
        op_valid_ = false;
        close_peers();
        const int64_t V = V_, T = T_;
        const int gV = cdiv64(V, kBlock), gT = std::max(1, cdiv64(T, kBlock));
        // --- idx_map (exclusive scan of the unknown flags) ---
        DevBuf<int> flag, excl;
        flag.alloc(V);
        excl.alloc(V + 1);
        unknown_flags<<<gV, kBlock, 0, stream_>>>(d_onb_.p, V, bc_, flag.p);
        long long n_ll = 0;
        exclusive_scan(flag.p, V, excl.p, &n_ll);
        nglob_ = n_ll;
        d_idx_.alloc(V);
        d_node_of_.alloc(std::max<int64_t>(nglob_, 1));
        make_idx_map<<<gV, kBlock, 0, stream_>>>(flag.p, excl.p, V, d_idx_.p, d_node_of_.p);
        launches_ += 2;
        n_dir_ = V - nglob_;
        // epilogues that walk rows instead of nodes (partitioned finish_rows, pcg_cluster) need the Dirichlet nodes as a list
        d_dir_nodes_.alloc(std::max<int64_t>(n_dir_, 1));
        make_dirichlet_list<<<gV, kBlock, 0, stream_>>>(flag.p, excl.p, V, d_dir_nodes_.p);
        launches_++;
        if (world_ > 1) {
            number_partition();
        } else {
            n_ = n_int_ = ghost_base_ = n_cols_ = nglob_;
            n_ghost_ = 0;
            rowmap_p_ = colmap_p_ = d_idx_.p;
            node_of_col_p_ = d_node_of_.p;
            hp_ = HaloPlan();
        }
        const int64_t n = n_;
        const int gN = std::max(1, cdiv64(n, kBlock));
        // --- row candidates -> sorted unique rows -> rowptr / col ---
        DevBuf<int> cnt, cptr, fill, cand, rowlen;
        cnt.alloc(n + 1);
        cptr.alloc(n + 2);
        fill.alloc(n + 1);
        rowlen.alloc(n + 1);
        CK(cudaMemsetAsync(cnt.p, 0, (n + 1) * sizeof(int), stream_));
        CK(cudaMemsetAsync(fill.p, 0, (n + 1) * sizeof(int), stream_));
        if (T) count_row_candidates<<<gT, kBlock, 0, stream_>>>(d_tri_.p, T, rowmap_p_, colmap_p_, cnt.p);
        long long ncand = 0;
        exclusive_scan(cnt.p, n, cptr.p, &ncand);
        REQUIRE(ncand < (long long)INT32_MAX, FEAB_EINVAL, "symbolic: candidate list exceeds int32");
        cand.alloc(std::max<long long>(ncand, 1));
        if (T) fill_row_candidates<<<gT, kBlock, 0, stream_>>>(d_tri_.p, T, rowmap_p_, colmap_p_, cptr.p, fill.p, cand.p);
        sort_unique_rows<<<gN, kBlock, 0, stream_>>>(cptr.p, cand.p, n, rowlen.p);
        d_rowptr_.alloc(n + 1);
        long long nnz_ll = 0;
        exclusive_scan(rowlen.p, n, d_rowptr_.p, &nnz_ll);
        nnz_ = nnz_ll;
        d_col_.alloc(std::max<int64_t>(nnz_, 1));
        d_diag_.alloc(std::max<int64_t>(n, 1));
        DevBuf<int> maxlen;
        maxlen.alloc(1);
        CK(cudaMemsetAsync(maxlen.p, 0, sizeof(int), stream_));
        compact_rows<<<gN, kBlock, 0, stream_>>>(cptr.p, cand.p, d_rowptr_.p, n, d_col_.p, d_diag_.p, maxlen.p);
        d_advp_.alloc(std::max<int64_t>(nnz_, 1));
        mark_adv_present<<<gN, kBlock, 0, stream_>>>(d_rowptr_.p, d_col_.p, n, node_of_col_p_, d_onb_.p, d_advp_.p);
        launches_ += 5;
        CK(cudaMemcpyAsync(&max_row_, maxlen.p, sizeof(int), cudaMemcpyDeviceToHost, stream_));
        // --- scatter map (scatter assembly) ---
        d_smap_.alloc(std::max<int64_t>(9 * T, 1));
        if (T) build_scatter_map<<<gT, kBlock, 0, stream_>>>(d_tri_.p, T, rowmap_p_, colmap_p_, d_rowptr_.p, d_col_.p, d_smap_.p);
        // --- ordered corner adjacency (ordered assembly) ---
        d_cptr_.alloc(n + 1);
        CK(cudaMemsetAsync(cnt.p, 0, (n + 1) * sizeof(int), stream_));
        CK(cudaMemsetAsync(fill.p, 0, (n + 1) * sizeof(int), stream_));
        if (T) count_row_corners<<<gT, kBlock, 0, stream_>>>(d_tri_.p, T, rowmap_p_, cnt.p);
        long long ncorner = 0;
        exclusive_scan(cnt.p, n, d_cptr_.p, &ncorner);
        d_corner_.alloc(std::max<long long>(ncorner, 1));
        if (T) fill_row_corners<<<gT, kBlock, 0, stream_>>>(d_tri_.p, T, rowmap_p_, d_cptr_.p, fill.p, d_corner_.p);
        sort_row_corners<<<gN, kBlock, 0, stream_>>>(d_cptr_.p, d_corner_.p, n);
        launches_ += 4;
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(stream_));
        // --- numeric storage ---
        const int64_t nz = std::max<int64_t>(nnz_, 1), nn = std::max<int64_t>(n, 1);
        d_M_.alloc(nz);
        d_K_.alloc(nz);
        d_Adv_.alloc(nz);
        d_AdvT_.alloc(nz);
        d_Md_.alloc(nz);
        d_A0_.alloc(nz);
        d_A1_.alloc(nz);
        d_dinv0_.alloc(nn);
        d_dinv1_.alloc(nn);
        d_y_.alloc(nn);
        for (int s = 0; s < 2; s++) d_r_[s].alloc(nn);
        for (int s = 0; s < 5; s++) d_w_[s].alloc(nn);
        num_tiles_ = std::max(1, cdiv64(n, kTileRows));
        g_tile_ = std::min(num_tiles_, sms_ * 8);
        g_vec_ = std::max(1, std::min(cdiv64(std::max<int64_t>(n / Vec<R>::N, 1), kBlock), sms_ * 8));
        g_part_ = std::max(std::max(g_tile_, g_vec_), sms_ * 8);
        d_part_.alloc((size_t)16 * g_part_);
        alloc_arena();
    }

    // The vectors peers write into (ghost tails of p, u, v; the surface values) live in ONE allocation that starts
    // with the CommCtrl mailboxes, so a peer needs a single mapping (feab_mg_export / feab_mg_connect).
    void alloc_arena() {
        const size_t nc = (size_t)std::max<int64_t>(n_cols_, 1) + kTmaPad;
        auto pad = [](size_t b) { return (b + 255) / 256 * 256; };
        size_t off = kCtrlBytes;
        for (int v = 0; v < 4; v++) {
            off_vec_[v] = off;
            off += pad(nc * sizeof(R));
        }
        off_vec_[kVecValues] = off;
        off += pad(((size_t)V_ + kTmaPad) * sizeof(float));
        d_arena_.release();
        d_arena_.alloc(off);
        CK(cudaMemsetAsync(d_arena_.p, 0, off, stream_));
        unsigned char* base = d_arena_.p;
        d_p_[0].view(base + off_vec_[kVecP0], nc);
        d_p_[1].view(base + off_vec_[kVecP1], nc);
        d_u_.view(base + off_vec_[kVecU], nc);
        d_v_.view(base + off_vec_[kVecV], nc);
        d_values_.view(base + off_vec_[kVecValues], (size_t)V_);
        CK(cudaMemsetAsync(d_loc_, 0, sizeof(CommLocal), stream_));
        CK(cudaStreamSynchronize(stream_));
        cm_ = Comm();
        cm_.rank = rank_;
        cm_.world = world_;
        cm_.loc = d_loc_;
        cm_.ctrl[rank_] = reinterpret_cast<CommCtrl*>(base);
        hp_.base[rank_] = reinterpret_cast<char*>(base);
        for (int v = 0; v < 5; v++) hp_.off_vec[v][rank_] = off_vec_[v];
        for (int q = 0; q < kMaxRanks; q++) connected_[q] = (q == rank_);
    }

    void close_peers() {
        for (int q = 0; q < kMaxRanks; q++) {
            if (ipc_opened_[q]) cudaIpcCloseMemHandle(ipc_opened_[q]);
            ipc_opened_[q] = nullptr;
            connected_[q] = false;
        }
    }

    void resize_state() {
        const int64_t nc = std::max<int64_t>(n_cols_, 1);
        CK(cudaMemsetAsync(d_u_.p, 0, nc * sizeof(R), stream_));
        CK(cudaMemsetAsync(d_v_.p, 0, nc * sizeof(R), stream_));
        CK(cudaStreamSynchronize(stream_));
        values_dirty_ = true;
    }

    void assemble(const float* vel) override {
        REQUIRE(have_mesh_, FEAB_ENOMESH, "assemble: no mesh");
        CK(cudaSetDevice(device_));
        if (vel) std::memcpy(vel_, vel, 3 * sizeof(float));
        op_valid_ = false;
        const int64_t nz = std::max<int64_t>(nnz_, 1);
        if (opt_.assembly == FEAB_ASSEMBLY_ORDERED) {
            if (n_ > 0) {
                assemble_ordered<R><<<cdiv64(n_, kBlock), kBlock, 0, stream_>>>(d_xyz_.p, d_tri_.p, n_, d_onb_.p, colmap_p_, d_cptr_.p,
                                                                                  d_corner_.p, d_rowptr_.p, d_col_.p, (R)vel_[0], (R)vel_[1],
                                                                                  (R)vel_[2], d_M_.p, d_K_.p, d_Adv_.p, d_AdvT_.p);
                launches_++;
            }
        } else {
            CK(cudaMemsetAsync(d_M_.p, 0, nz * sizeof(R), stream_));
            CK(cudaMemsetAsync(d_K_.p, 0, nz * sizeof(R), stream_));
            CK(cudaMemsetAsync(d_Adv_.p, 0, nz * sizeof(R), stream_));
            CK(cudaMemsetAsync(d_AdvT_.p, 0, nz * sizeof(R), stream_));
            if (T_ > 0) {
                assemble_scatter<R><<<cdiv64(T_, kBlock), kBlock, 0, stream_>>>(d_xyz_.p, d_tri_.p, T_, d_onb_.p, d_smap_.p, (R)vel_[0],
                                                                                  (R)vel_[1], (R)vel_[2], d_M_.p, d_K_.p, d_Adv_.p, d_AdvT_.p);
                launches_++;
            }
        }
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(stream_));
    }

    void get_counts(int64_t* V, int64_t* n, int64_t* nnz, int* maxrow) override {
        REQUIRE(have_mesh_, FEAB_ENOMESH, "get_counts: no mesh");
        if (V) *V = V_;
        if (n) *n = n_;
        if (nnz) *nnz = nnz_;
        if (maxrow) *maxrow = max_row_;
    }
    void get_idx_map(int32_t* out) override {
        REQUIRE(have_mesh_, FEAB_ENOMESH, "get_idx_map: no mesh");
        CK(cudaSetDevice(device_));
        CK(cudaMemcpy(out, d_idx_.p, V_ * sizeof(int), cudaMemcpyDeviceToHost));
    }
    void get_csr(int which, int64_t* nnz_out, int32_t* rowptr, int32_t* col, void* val) override {
        REQUIRE(have_mesh_, FEAB_ENOMESH, "get_csr: no mesh");
        REQUIRE(which >= 0 && which <= 2, FEAB_EINVAL, "get_csr: which must be 0..2");
        CK(cudaSetDevice(device_));
        std::vector<int32_t> rp(n_ + 1), cl(std::max<int64_t>(nnz_, 1));
        CK(cudaMemcpy(rp.data(), d_rowptr_.p, (n_ + 1) * sizeof(int), cudaMemcpyDeviceToHost));
        if (which != FEAB_MAT_ADVECTION) {
            if (nnz_out) *nnz_out = nnz_;
            if (rowptr) std::memcpy(rowptr, rp.data(), (n_ + 1) * sizeof(int));
            if (col) CK(cudaMemcpy(col, d_col_.p, nnz_ * sizeof(int), cudaMemcpyDeviceToHost));
            if (val) CK(cudaMemcpy(val, which == FEAB_MAT_MASS ? d_M_.p : d_K_.p, nnz_ * sizeof(R), cudaMemcpyDeviceToHost));
            return;
        }
        // the advection matrix keeps only entries between two non-boundary nodes (FEMContext.cpp:188)
        std::vector<uint8_t> pres(std::max<int64_t>(nnz_, 1));
        CK(cudaMemcpy(pres.data(), d_advp_.p, nnz_, cudaMemcpyDeviceToHost));
        int64_t cnt = 0;
        for (int64_t k = 0; k < nnz_; k++) cnt += pres[k];
        if (nnz_out) *nnz_out = cnt;
        if (!rowptr && !col && !val) return;
        std::vector<R> vv(std::max<int64_t>(nnz_, 1));
        CK(cudaMemcpy(cl.data(), d_col_.p, nnz_ * sizeof(int), cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(vv.data(), d_Adv_.p, nnz_ * sizeof(R), cudaMemcpyDeviceToHost));
        int64_t w = 0;
        for (int64_t r = 0; r < n_; r++) {
            if (rowptr) rowptr[r] = (int32_t)w;
            for (int k = rp[r]; k < rp[r + 1]; k++)
                if (pres[k]) {
                    if (col) col[w] = cl[k];
                    if (val) static_cast<R*>(val)[w] = vv[k];
                    w++;
                }
        }
        if (rowptr) rowptr[n_] = (int32_t)w;
    }

    // ---------------------------------------------------------------- parameters / state
    void set_equation(int eq) override {
        REQUIRE(eq >= 0 && eq <= 3, FEAB_EINVAL, "set_equation: equation must be 0..3");
        if (eq != eq_) values_dirty_ = true;  // another vector takes surface->values now
        eq_ = eq;
    }
    void set_params(int eq, const FeabParams* p) override {
        REQUIRE(eq >= 0 && eq <= 3 && p, FEAB_EINVAL, "set_params: bad arguments");
        par_[eq] = *p;
    }
    void get_params(int eq, FeabParams* p) override {
        REQUIRE(eq >= 0 && eq <= 3 && p, FEAB_EINVAL, "get_params: bad arguments");
        *p = par_[eq];
    }
    void set_options(const FeabSolverOptions* o) override {
        REQUIRE(o, FEAB_EINVAL, "set_solver_options: null");
        REQUIRE(o->solver == 0 || o->solver == 1, FEAB_EINVAL, "set_solver_options: solver");
        REQUIRE(o->loop_mode >= 0 && o->loop_mode <= 2, FEAB_EINVAL, "set_solver_options: loop_mode");
        REQUIRE(o->loop_mode != FEAB_LOOP_PERSISTENT || pcg_max_grid_ > 0, FEAB_EINVAL, "set_solver_options: device cannot launch cooperative kernels");
        REQUIRE(o->assembly == 0 || o->assembly == 1, FEAB_EINVAL, "set_solver_options: assembly");
        bool reassemble = o->assembly != opt_.assembly;
        opt_ = *o;
        drop_graphs();
        op_valid_ = false;
        if (reassemble && have_mesh_) assemble(nullptr);
    }
    void get_options(FeabSolverOptions* o) override { *o = opt_; }

    void clear() override {
        REQUIRE(have_mesh_, FEAB_ENOMESH, "clear: no mesh");
        CK(cudaSetDevice(device_));
        resize_state();
        CK(cudaMemsetAsync(d_values_.p, 0, V_ * sizeof(float), stream_));
        CK(cudaMemsetAsync(d_unstable_, 0, sizeof(int), stream_));
        CK(cudaStreamSynchronize(stream_));
        values_dirty_ = true;
    }
    void set_values(const float* v) override {
        REQUIRE(have_mesh_ && v, FEAB_ENOMESH, "set_values: no mesh");
        CK(cudaSetDevice(device_));
        CK(cudaMemcpyAsync(d_values_.p, v, V_ * sizeof(float), cudaMemcpyHostToDevice, stream_));
        CK(cudaStreamSynchronize(stream_));
        values_dirty_ = true;
    }
    const float* values_device() override {
        REQUIRE(have_mesh_, FEAB_ENOMESH, "values_device: no mesh");
        CK(cudaSetDevice(device_));
        CK(cudaStreamSynchronize(stream_));
        return d_values_.p;
    }
    void get_values(float* v) override {
        REQUIRE(have_mesh_ && v, FEAB_ENOMESH, "get_values: no mesh");
        CK(cudaSetDevice(device_));
        CK(cudaMemcpyAsync(v, d_values_.p, V_ * sizeof(float), cudaMemcpyDeviceToHost, stream_));
        CK(cudaStreamSynchronize(stream_));
    }
    void brush(const int64_t* nodes, int64_t count, float strength) override {
        REQUIRE(have_mesh_, FEAB_ENOMESH, "brush: no mesh");
        if (count <= 0) return;
        CK(cudaSetDevice(device_));
        d_brush_.alloc(count);
        CK(cudaMemcpyAsync(d_brush_.p, nodes, count * sizeof(long long), cudaMemcpyHostToDevice, stream_));
        brush_nodes<<<cdiv64(count, kBlock), kBlock, 0, stream_>>>(d_values_.p, d_brush_.p, count, V_, strength);
        launches_++;
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(stream_));
        values_dirty_ = true;
    }
    void get_state(int which, void* out) override {
        REQUIRE(have_mesh_ && out && (which == 0 || which == 1), FEAB_EINVAL, "get_state: bad arguments");
        CK(cudaSetDevice(device_));
        CK(cudaMemcpy(out, which == 0 ? d_u_.p : d_v_.p, n_ * sizeof(R), cudaMemcpyDeviceToHost));
    }
    void set_state(int which, const void* in) override {
        REQUIRE(have_mesh_ && in && (which == 0 || which == 1), FEAB_EINVAL, "set_state: bad arguments");
        CK(cudaSetDevice(device_));
        CK(cudaMemcpy(which == 0 ? d_u_.p : d_v_.p, in, n_ * sizeof(R), cudaMemcpyHostToDevice));
        values_dirty_ = true;  // as before: the next step re-reads the vector surface->values feeds from values
    }
    int unstable() override {
        REQUIRE(have_mesh_, FEAB_ENOMESH, "unstable: no mesh");
        CK(cudaSetDevice(device_));
        int f = 0;
        CK(cudaMemcpy(&f, d_unstable_, sizeof(int), cudaMemcpyDeviceToHost));
        // finish_step stamps the flag with the step number (atomicMax keeps the latest unstable step): true when ANY step of the
        // last feab_step call went unstable -- the reference tests after every advance_time (src/Application.cpp:532-533)
        return f != 0 && call_first_stamp_ != 0 && f >= call_first_stamp_;
    }

    // ---------------------------------------------------------------- operator cache (K2)
    StepScalars scalars() const {
        const FeabParams& p = par_[eq_];
        // promote the reference's floats exactly
        return StepScalars{(double)p.time_step, (double)p.c[0], (double)p.c[1], (double)p.c[2], (double)p.c[3]};
    }
    bool use_bicgstab() const { return eq_ == FEAB_EQ_ADVECTION_DIFFUSION && opt_.solver == FEAB_SOLVER_BICGSTAB; }
    // BiCGSTAB through the multi-kernel WHILE graph / host-batched loop (bicgstab.cuh) instead of the persistent kernel
    bool use_bicgstab_legacy() const { return use_bicgstab() && (opt_.loop_mode != FEAB_LOOP_PERSISTENT || bicg_max_grid_ <= 0); }

    void ensure_operator() { ensure_operator_on(stream_); }
    void ensure_operator_on(cudaStream_t s) {
        const FeabParams& p = par_[eq_];
        const int transposed = (eq_ == FEAB_EQ_ADVECTION_DIFFUSION && !use_bicgstab()) ? 1 : 0;
        if (op_valid_ && op_eq_ == eq_ && op_transposed_ == transposed && std::memcmp(&op_par_, &p, sizeof(p)) == 0) return;
        REQUIRE(p.time_step != 0.0f, FEAB_EINVAL, "time_step must be non-zero");
        if (nnz_ > 0) {
            int g = std::min(cdiv64(nnz_, kBlock), sms_ * 16);
            build_operator<R><<<g, kBlock, 0, s>>>(eq_, transposed, nnz_, d_M_.p, d_K_.p, d_Adv_.p, d_AdvT_.p, scalars(), d_Md_.p, d_A0_.p,
                                                         d_A1_.p);
            inv_diag<R><<<cdiv64(n_, kBlock), kBlock, 0, s>>>(d_A0_.p, d_diag_.p, n_, d_dinv0_.p);
            launches_ += 2;
            if (eq_ == FEAB_EQ_REACTION_DIFFUSION) {
                inv_diag<R><<<cdiv64(n_, kBlock), kBlock, 0, s>>>(d_A1_.p, d_diag_.p, n_, d_dinv1_.p);
                launches_++;
            }
            CK(cudaGetLastError());
        }
        op_valid_ = true;
        op_eq_ = eq_;
        op_transposed_ = transposed;
        op_par_ = p;
        // the step graph bakes the scalars into its kernel nodes
        drop_graphs();
    }

    // ---------------------------------------------------------------- one time step
    R tol() const { return opt_.tolerance >= 0 ? (R)opt_.tolerance : Ops<R>::eps(); }
    int maxit() const {
        int64_t m = opt_.max_iter >= 0 ? opt_.max_iter : 2 * n_;
        return (int)std::min<int64_t>(m, INT32_MAX - 2);
    }

    R* part_pap() { return d_part_.p; }
    R* part_rr() { return d_part_.p + (size_t)g_part_; }
    R* part_rz() { return d_part_.p + (size_t)2 * g_part_; }
    R* part_rhs() { return d_part_.p + (size_t)3 * g_part_; }   // 4 x g_tile_
    R* part_bi() { return d_part_.p + (size_t)8 * g_part_; }    // 8 x g_part_ for BiCGSTAB

    // ---- phases of one time step (the same code is captured into the step graph, launched directly in
    //      host-batched mode, and driven in lock-step over several handles by feab_mg_group_step) ----
    int first_halo_tile(int rows_per_tile) const { return world_ > 1 ? (int)(n_int_ / rows_per_tile) : INT32_MAX; }

    // K6: RHS + PCG start.  Heat / advection-diffusion / Gray-Scott stream M/dt through the TMA ring (rhs_tma.cuh); the wave
    // equation (two value arrays) and FEAB_RHS_TMA=0 use the LDG kernel of solver.cuh.  Same bits either way.
    static constexpr int kRhsRows = 512, kRhsCpr = 8, kRhsStages = 2, kRhsMinB = 2;
    // publish: the kernel's last CTA all-reduces |b|^2, b.D^-1 b into every rank's mailbox (the persistent solve kernels and
    // every partitioned handle start from there) and, on a partition, the kernel stores the first search directions of the
    // boundary rows into the neighbours' ghost tails and raises their halo flags.
    template <int EQ>
    void launch_rhs_tma(cudaStream_t s, const StepScalars& sc, R* prt, const Comm& cm, int publish) {
        using Cfg = TmaCfg<R, kRhsRows, kRhsCpr, kRhsStages>;
        const int nt = std::max(1, cdiv64(n_, kRhsRows));
        const int grid = std::max(1, std::min({nt, kRhsMinB * sms_, g_tile_, grid_cap_}));
        rhs_init_tma<R, EQ, kRhsRows, kRhsCpr, kRhsStages, kRhsMinB><<<grid, kRhsRows, Cfg::kSmemBytes, s>>>(
            d_rowptr_.p, d_col_.p, d_Md_.p, d_u_.p, d_v_.p, d_dinv0_.p, d_dinv1_.p, n_, nt, sc, d_r_[0].p, d_p_[0].p, d_r_[1].p, d_p_[1].p, prt,
            g_tile_, cm, hp_, publish);
    }
    void launch_rhs(cudaStream_t s, const StepScalars& sc, R* prt, const Comm& cm, int publish) {
        const size_t sm2 = 2 * kTileCap * sizeof(R);
#define FEAB_RHS(EQ)                                                                                                             \
    rhs_init<R, EQ><<<g_tile_, kBlock, sm2, s>>>(d_rowptr_.p, d_col_.p, d_Md_.p, d_K_.p, d_u_.p, d_v_.p, d_dinv0_.p, d_dinv1_.p, n_, \
                                                 num_tiles_, sc, d_r_[0].p, d_p_[0].p, d_r_[1].p, d_p_[1].p, prt, cm, hp_, publish)
        switch (eq_) {
            case FEAB_EQ_HEAT:
                if (rhs_tma_) launch_rhs_tma<0>(s, sc, prt, cm, publish); else FEAB_RHS(0);
                break;
            case FEAB_EQ_WAVE: FEAB_RHS(1); break;
            case FEAB_EQ_ADVECTION_DIFFUSION:
                if (rhs_tma_) launch_rhs_tma<2>(s, sc, prt, cm, publish); else FEAB_RHS(2);
                break;
            default:
                if (rhs_tma_) launch_rhs_tma<3>(s, sc, prt, cm, publish); else FEAB_RHS(3);
                break;
        }
#undef FEAB_RHS
    }

    void ph_push_state(cudaStream_t s) {  // ghosts of u, v and the surface values of boundary rows -> neighbours
        if (world_ == 1) return;
        state_push<R><<<1, 1024, 0, s>>>(cm_, hp_, d_u_.p, d_v_.p, d_values_.p);
        launches_++;
    }
    // true when the solves of this step start inside a persistent kernel (from the RHS kernel's published sums)
    bool persistent_solves(bool capturing) const { return !capturing && !lockstep_ && opt_.loop_mode == FEAB_LOOP_PERSISTENT && !use_bicgstab_legacy(); }
    void ph_step_begin(cudaStream_t s, bool capturing) {
        const int gV = cdiv64(V_, kBlock);
        const StepScalars sc = scalars();
        // surface->values -> solver vector only when somebody else wrote values since the last step (brush, set_values, host
        // upload, ...): the solver's own result is re-read through float by the RHS kernel (solver.cuh through_float).  A
        // captured graph and the lock-step test driver always gather.
        const bool gather = values_dirty_ || capturing || lockstep_;
        if (gather) {
            R* target = eq_ == FEAB_EQ_REACTION_DIFFUSION ? d_v_.p : d_u_.p;
            if (world_ > 1) wait_kernel<<<1, 32, 0, s>>>(cm_, hp_, 1);  // ghost values: the neighbours' end-of-step pushes have landed
            if (world_ > 1)
                gather_cols<R><<<std::max(1, std::min({cdiv64(n_cols_, kBlock), sms_ * 4, grid_cap_})), kBlock, 0, s>>>(d_values_.p, node_of_col_p_, n_cols_,
                                                                                                                    target, cm_);
            else
                gather_values<R><<<gV, kBlock, 0, s>>>(d_values_.p, colmap_p_, V_, target, cm_, hp_);
        }
        const int publish = (world_ > 1 || persistent_solves(capturing)) ? 1 : 0;
        launch_rhs(s, sc, part_rhs(), cm_, publish);
        if (!capturing) {
            launches_ += gather ? (world_ > 1 ? 3 : 2) : 1;
            values_dirty_ = false;
        }
        CK(cudaGetLastError());
    }
    // partitioned, three-kernel loop drivers only: totals of the RHS dot products for cg_begin.
    // A separate phase: it consumes what EVERY rank's rhs_init published.
    void ph_step_collect(cudaStream_t s, bool capturing) {
        if (world_ > 1 && !persistent_solves(capturing)) {
            const bool gs = eq_ == FEAB_EQ_REACTION_DIFFUSION;
            if (gs)
                ar_collect<R, 4><<<1, 32, 0, s>>>(cm_, d_rhs_tot_.p);
            else
                ar_collect<R, 2><<<1, 32, 0, s>>>(cm_, d_rhs_tot_.p);
            if (!capturing) launches_ += 1;
        }
        CK(cudaGetLastError());
    }
    void ph_cg_begin(cudaStream_t s, int slot, R* x, cudaGraphConditionalHandle cond, int use_cond) {
        CgState<R>* st = d_st_ + slot;
        CK(cudaMemsetAsync(x, 0, n_ * sizeof(R), s));
        if (world_ > 1) {
            const R* tot = d_rhs_tot_.p + 2 * slot;
            cg_begin<R><<<1, kBlock, 0, s>>>(st, tot, tot + 1, 1, tol(), maxit(), cond, use_cond);
        } else {
            const R* rhs = part_rhs() + (size_t)(slot * 2) * g_tile_;
            cg_begin<R><<<1, kBlock, 0, s>>>(st, rhs, rhs + g_tile_, g_tile_, tol(), maxit(), cond, use_cond);
        }
    }
    void ph_spmv(cudaStream_t s, int slot, const R* A) { launch_spmv(s, A, d_p_[slot].p, d_y_.p, d_st_ + slot); }
    void ph_update(cudaStream_t s, int slot, const R* dinv, R* x) {
        cg_update<R><<<g_vec_, kBlock, 0, s>>>(x, d_r_[slot].p, d_p_[slot].p, d_y_.p, dinv, n_, d_st_ + slot, part_pap(), spmv_parts(), part_rr(),
                                               part_rz(), cm_);
    }
    void ph_direction(cudaStream_t s, int slot, const R* dinv, cudaGraphConditionalHandle cond, int use_cond) {
        cg_direction<R><<<g_vec_, kBlock, 0, s>>>(d_r_[slot].p, d_p_[slot].p, dinv, n_, d_st_ + slot, part_rr(), part_rz(), g_vec_, cond, use_cond,
                                                  cm_);
        if (world_ > 1) halo_push<R><<<1, 1024, 0, s>>>(cm_, hp_, d_p_[slot].p, slot == 0 ? kVecP0 : kVecP1, (const R*)nullptr, 0);
    }
    int kernels_per_iteration() const { return world_ > 1 ? 4 : 3; }
    // epilogue: values <- solution, wave u += v dt, instability stamp; on a partition the owners of boundary rows also store
    // u, v and the value into the neighbours' ghosts and the last CTA raises the state flags (no separate push kernel)
    void ph_finish(cudaStream_t s, bool capturing) {
        const int gV = cdiv64(V_, kBlock);
        // a captured graph replays one stamp: it keeps the sticky semantics of a flag that is reset by feab_clear only
        // a captured graph replays one stamp; step() clears the flag once before it launches the graphs of a call
        if (!capturing && step_stamp_ >= INT32_MAX - 1) {  // only the lock-step driver gets here: step() wraps at the start of a call
            CK(cudaMemsetAsync(d_unstable_, 0, sizeof(int), s));
            step_stamp_ = 0;
            call_first_stamp_ = 0;
        }
        const int stamp = capturing ? 1 : ++step_stamp_;
        if (!capturing) {
            last_stamp_ = stamp;
            if (!call_first_stamp_) call_first_stamp_ = stamp;
        }
        if (world_ > 1)
            finish_rows<R><<<std::max(1, std::min({cdiv64(n_, kBlock), sms_ * 4, grid_cap_})), kBlock, 0, s>>>(
                d_values_.p, node_of_col_p_, n_, d_dir_nodes_.p, n_dir_, d_u_.p, d_v_.p, eq_ == FEAB_EQ_WAVE ? 1 : 0, (R)scalars().dt,
                eq_ == FEAB_EQ_REACTION_DIFFUSION ? 1 : 0, d_unstable_, stamp, cm_, hp_);
        else
            finish_step<R><<<gV, kBlock, 0, s>>>(d_values_.p, d_idx_.p, rowmap_p_, V_, d_u_.p, d_v_.p, eq_ == FEAB_EQ_WAVE ? 1 : 0, (R)scalars().dt,
                                                 eq_ == FEAB_EQ_REACTION_DIFFUSION ? 1 : 0, d_unstable_, stamp, cm_, hp_);
        if (!capturing) launches_ += 1;
        CK(cudaGetLastError());
    }
    // the solves of the current equation: (operator, inverse diagonal, solution vector)
    int solves(const R* A[2], const R* dinv[2], R* x[2]) {
        A[0] = d_A0_.p;
        dinv[0] = d_dinv0_.p;
        x[0] = eq_ == FEAB_EQ_WAVE ? d_v_.p : d_u_.p;
        A[1] = d_A1_.p;
        dinv[1] = d_dinv1_.p;
        x[1] = d_v_.p;
        return eq_ == FEAB_EQ_REACTION_DIFFUSION ? 2 : 1;
    }

    // Launch of a persistent kernel whose CTAs sync with each other: the grid never exceeds what the device holds at once
    // (occupancy x SMs, computed at creation).  A cooperative launch makes the driver check exactly that; FEAB_COOP=0 uses a
    // plain launch of the same grid (an experiment switch: the co-residency argument is the same, the check is skipped).
    cudaError_t launch_resident(const void* f, dim3 grid, dim3 block, void** args, size_t smem, cudaStream_t s) {
        return coop_launch_ ? cudaLaunchCooperativeKernel(f, grid, block, args, smem, s) : cudaLaunchKernel(f, grid, block, args, smem, s);
    }

    void launch_iteration(cudaStream_t s, int slot, const R* A, const R* dinv, R* x, cudaGraphConditionalHandle cond, int use_cond) {
        ph_spmv(s, slot, A);
        ph_update(s, slot, dinv, x);
        ph_direction(s, slot, dinv, cond, use_cond);
    }

    // PCG solve of slot `slot` into x.  capturing: build WHILE node into the graph being captured on stream_.
    bool dual_applies(bool capturing) const {
        return dual_ && !capturing && opt_.loop_mode == FEAB_LOOP_PERSISTENT && eq_ == FEAB_EQ_REACTION_DIFFUSION && dual_max_grid_ > 0 &&
               !use_bicgstab();
    }
    // Order of phase C of the persistent PCG kernels (PcgArgs::zigzag).  Measured (profiles/r02_v17_*, r02_v18_*, r02_v19_*, r02_v20_*):
    // plain ascending is the fastest everywhere -- one GPU 53.95 vs 52.3 steps/s at 16 M rows against descending, N = 2 103.5
    // vs 100.1, N = 8 342.2 vs 325.0 against the order that starts on the boundary rows.
    int phase_c_order() const { return zigzag_ >= 0 ? zigzag_ : 0; }
    // Gray-Scott: u- and v-solve advance together until the first of them stops (pcg_dual.cuh); the single-system kernel
    // then finishes whichever is still running (it picks the system from CgState; both done -> it returns at once).
    // Two launches per step; the start of both solves (thresholds, early-outs) is taken inside pcg_dual.
    void enqueue_dual(const R* const* A, const R* const* dinv, R* const* x) {
        PcgDualArgs<R> da;
        da.rowptr = d_rowptr_.p;
        da.col = d_col_.p;
        for (int s = 0; s < 2; s++) {
            da.A[s] = A[s];
            da.dinv[s] = dinv[s];
            da.x[s] = x[s];
            da.r[s] = d_r_[s].p;
            da.p[s] = d_p_[s].p;
        }
        da.y[0] = d_y_.p;
        da.y[1] = d_w_[0].p;
        da.n = n_;
        da.num_tiles = std::max(1, cdiv64(n_, kDualRows));
        da.st = d_st_;
        da.part = d_dpart_.p;
        da.first_halo_tile = first_halo_tile(kDualRows);
        da.tol = tol();
        da.maxit = maxit();
        da.zigzag = phase_c_order();
        da.prof = prof_dual_ ? d_prof_ : nullptr;
        da.arrivals = nullptr;
        if (prof_arrivals_) {
            d_arrivals_.alloc((size_t)6 * std::max(dual_max_grid_, 1));
            da.arrivals = d_arrivals_.p;
            arrivals_grid_ = std::min(dual_max_grid_, da.num_tiles);
        }
        const int grid = std::min(dual_max_grid_, da.num_tiles);
        using DCfg = DualCfg<R, kDualRows, kPcgCpr, kPcgStages>;
        void* kargs[] = {(void*)&da, (void*)&cm_, (void*)&hp_};
        if (time_solves_) CK(cudaEventRecord(ev_solve_[0], stream_));
        CK(launch_resident((const void*)pcg_dual<R, kDualRows, kPcgCpr, kPcgStages, kPcgMinB>, dim3(grid), dim3(2 * kDualRows), kargs,
                                       DCfg::kSmemBytes, stream_));
        launches_ += 1;
        launch_pcg_persistent(-1, 0, A, dinv, x);
        if (time_solves_) CK(cudaEventRecord(ev_solve_[1], stream_));
    }

    // slot >= 0: that system; -1: whichever pcg_dual left running.  begin: number of systems whose start this kernel takes
    // from the RHS kernel's published sums (0: CgState is already set by an earlier kernel of the step).
    void launch_pcg_persistent(int slot, int begin, const R* const* A, const R* const* dinv, R* const* x) {
        PcgArgs<R> pa;
        pa.rowptr = d_rowptr_.p;
        pa.col = d_col_.p;
        for (int s = 0; s < 2; s++) {
            pa.A[s] = A[s];
            pa.dinv[s] = dinv[s];
            pa.x[s] = x[s];
            pa.r[s] = d_r_[s].p;
            pa.p[s] = d_p_[s].p;
        }
        pa.y = d_y_.p;
        pa.n = n_;
        pa.num_tiles = std::max(1, cdiv64(n_, kPcgRows));
        pa.first_halo_tile = first_halo_tile(kPcgRows);
        pa.st = d_st_;
        pa.part = d_ppart_.p;
        pa.slot = slot;
        pa.begin = begin;
        pa.tol = tol();
        pa.maxit = maxit();
        pa.zigzag = phase_c_order();
        pa.prof = prof_dual_ ? nullptr : d_prof_;
        const int grid = std::min(pcg_max_grid_, pa.num_tiles);
        using PCfg = TmaCfg<R, kPcgRows, kPcgCpr, kPcgStages>;
        void* kargs[] = {(void*)&pa, (void*)&cm_, (void*)&hp_};
        CK(launch_resident((const void*)pcg_persistent<R, kPcgRows, kPcgCpr, kPcgStages, kPcgMinB>, dim3(grid), dim3(kPcgRows), kargs,
                                       PCfg::kSmemBytes, stream_));
        launches_ += 1;
    }

    // Non-symmetric mode, persistent loop driver: the whole BiCGSTAB solve is one cooperative launch (bicg_persistent.cuh).
    // Vectors: r = d_r_[0] (holds b from the RHS kernel), y = d_p_[0] and z = d_p_[1] (the arena vectors with ghost tails: the
    // two vectors the SpMV gathers from), x = d_u_, p = d_w_[2], v = d_w_[1], r0 = d_w_[0], s = d_r_[1], t = d_y_, b = d_w_[4].
    void launch_bicg_persistent() {
        BicgArgs<R> ba;
        ba.rowptr = d_rowptr_.p;
        ba.col = d_col_.p;
        ba.A = d_A0_.p;
        ba.dinv = d_dinv0_.p;
        ba.x = d_u_.p;
        ba.r = d_r_[0].p;
        ba.r0 = d_w_[0].p;
        ba.v = d_w_[1].p;
        ba.p = d_w_[2].p;
        ba.y = d_p_[0].p;
        ba.s = d_r_[1].p;
        ba.z = d_p_[1].p;
        ba.t = d_y_.p;
        ba.b = d_w_[4].p;
        ba.n = n_;
        ba.num_tiles = std::max(1, cdiv64(n_, kPcgRows));
        ba.first_halo_tile = first_halo_tile(kPcgRows);
        ba.st = d_bst_;
        ba.part = d_ppart_.p;
        ba.tol = tol();
        ba.maxit = maxit();
        ba.prof = d_prof_;
        const int grid = std::min(bicg_max_grid_, ba.num_tiles);
        using PCfg = TmaCfg<R, kPcgRows, kPcgCpr, kPcgStages>;
        void* kargs[] = {(void*)&ba, (void*)&cm_, (void*)&hp_};
        if (time_solves_) CK(cudaEventRecord(ev_solve_[0], stream_));
        CK(launch_resident((const void*)bicg_persistent<R, kPcgRows, kPcgCpr, kPcgStages, kPcgMinB>, dim3(grid), dim3(kPcgRows), kargs,
                                       PCfg::kSmemBytes, stream_));
        if (time_solves_) CK(cudaEventRecord(ev_solve_[1], stream_));
        launches_ += 1;
    }

    // Small meshes on one GPU with the persistent loop driver: all steps of the call in one cluster launch (pcg_cluster.cuh)
    bool use_cluster() const {
        return cluster_ctas_ > 0 && world_ == 1 && !lockstep_ && opt_.loop_mode == FEAB_LOOP_PERSISTENT && !use_bicgstab() && n_ > 0 &&
               n_ <= (int64_t)cluster_ctas_ * kClusterMaxRowsPerCta;
    }
    void launch_cluster_steps(int nsteps) {
        const R* A[2];
        const R* dinv[2];
        R* x[2];
        solves(A, dinv, x);
        if (values_dirty_) {  // somebody else wrote surface->values since the last step
            R* target = eq_ == FEAB_EQ_REACTION_DIFFUSION ? d_v_.p : d_u_.p;
            gather_values<R><<<cdiv64(V_, kBlock), kBlock, 0, stream_>>>(d_values_.p, colmap_p_, V_, target, cm_, hp_);
            launches_++;
            values_dirty_ = false;
        }
        const int C = (int)std::min<int64_t>(cluster_ctas_, std::max<int64_t>(1, (n_ + 511) / 512));
        const int rpc = (int)((n_ + C - 1) / C);
        if (cluster_ent_key_ != n_ * 32 + C) {  // most entries any CTA's row block holds: sizes the shared-memory copy of the matrix
            std::vector<int> rp(C + 1);
            for (int q = 0; q <= C; q++)
                CK(cudaMemcpyAsync(&rp[q], d_rowptr_.p + std::min<int64_t>(n_, (int64_t)q * rpc), sizeof(int), cudaMemcpyDeviceToHost, stream_));
            CK(cudaStreamSynchronize(stream_));
            cluster_ent_max_ = 0;
            for (int q = 0; q < C; q++) cluster_ent_max_ = std::max(cluster_ent_max_, rp[q + 1] - rp[q]);
            cluster_ent_max_ = (cluster_ent_max_ + 1) & ~1;
            cluster_ent_key_ = n_ * 32 + C;
        }
        const size_t base = (size_t)5 * rpc * sizeof(R) + (size_t)(rpc + 2) * sizeof(int);
        int col_cap = 0, val_cap = 0;
        if (base + (size_t)cluster_ent_max_ * (sizeof(int) + sizeof(R)) <= (size_t)kClusterSmemBudget)
            col_cap = val_cap = cluster_ent_max_;
        else if (base + (size_t)cluster_ent_max_ * sizeof(int) <= (size_t)kClusterSmemBudget)
            col_cap = cluster_ent_max_;
        ClusterArgs<R> ca;
        ca.rowptr = d_rowptr_.p;
        ca.col = d_col_.p;
        ca.Md = d_Md_.p;
        ca.Kv = d_K_.p;
        for (int s = 0; s < 2; s++) {
            ca.A[s] = A[s];
            ca.dinv[s] = dinv[s];
            ca.r[s] = d_r_[s].p;
        }
        ca.u = d_u_.p;
        ca.v = d_v_.p;
        ca.p[0] = d_p_[0].p;
        ca.p[1] = d_p_[1].p;
        const bool pair = cluster_pair_ && eq_ == FEAB_EQ_REACTION_DIFFUSION;
        ca.pair = pair ? 1 : 0;
        ca.xbar = d_xbar_;
        ca.xbar_base = xbar_count_;
        if (pair) xbar_count_ += 4ull * (unsigned long long)nsteps;  // two meetings per step, two arrivals per meeting
        ca.loc = cm_.loc;
        ca.xerr = h_flag_ + 8;
        ca.values = d_values_.p;
        ca.node_of = node_of_col_p_;
        ca.dir_nodes = d_dir_nodes_.p;
        ca.n_dir = n_dir_;
        ca.n = n_;
        ca.eq = eq_;
        ca.sc = scalars();
        ca.tol = tol();
        ca.maxit = maxit();
        ca.nsteps = nsteps;
        ca.st = d_st_;
        ca.unstable = d_unstable_;
        ca.stamp0 = step_stamp_ + 1;  // step() made room for nsteps stamps
        if (!call_first_stamp_) call_first_stamp_ = ca.stamp0;
        ca.col_cap = col_cap;
        ca.prof = prof_dual_ ? nullptr : d_prof_;
        ca.val_cap = val_cap;
        step_stamp_ += nsteps;
        last_stamp_ = step_stamp_;
        cudaLaunchConfig_t cfg = {};
        const int threads = cluster_threads_;
        cfg.gridDim = dim3(pair ? 2 * C : C);
        cfg.blockDim = dim3(threads);
        cfg.dynamicSmemBytes = base + (size_t)val_cap * sizeof(R) + (size_t)col_cap * sizeof(int);
        cfg.stream = stream_;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = C;
        at[0].val.clusterDim.y = 1;
        at[0].val.clusterDim.z = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        CK(cudaEventRecord(ev_solve_[0], stream_));
        if (threads == 512)
            CK(cudaLaunchKernelEx(&cfg, pcg_cluster<R, 512>, ca));
        else
            CK(cudaLaunchKernelEx(&cfg, pcg_cluster<R, 1024>, ca));
        CK(cudaEventRecord(ev_solve_[1], stream_));
        launches_ += 1;
        cluster_steps_last_ = nsteps;
        solve_kernel_ = "pcg_cluster";
        dual_last_ = false;
    }

    // num_solves: solves of this step (the first one takes the start of all of them, see launch_pcg_persistent)
    void enqueue_pcg(int slot, const R* const* Aall, const R* const* dinvall, R* const* xall, bool capturing, int num_solves = 1) {
        const R* A = Aall[slot];
        const R* dinv = dinvall[slot];
        R* x = xall[slot];
        CgState<R>* st = d_st_ + slot;
        if (!capturing && opt_.loop_mode == FEAB_LOOP_PERSISTENT && !lockstep_) {
            // the whole solve is one cooperative launch: no WHILE node, no host polling (pcg_persistent.cuh)
            if (time_solves_) CK(cudaEventRecord(ev_solve_[2 * slot], stream_));
            launch_pcg_persistent(slot, slot == 0 ? num_solves : 0, Aall, dinvall, xall);
            if (time_solves_) CK(cudaEventRecord(ev_solve_[2 * slot + 1], stream_));
            return;
        }
        if (capturing) {
            cudaStreamCaptureStatus status;
            cudaGraph_t graph;
            const cudaGraphNode_t* deps;
            size_t ndeps;
            CK(cudaStreamGetCaptureInfo(stream_, &status, nullptr, &graph, &deps, &ndeps));
            cudaGraphConditionalHandle cond;
            CK(cudaGraphConditionalHandleCreate(&cond, graph, 0, 0));
            ph_cg_begin(stream_, slot, x, cond, 1);
            CK(cudaStreamGetCaptureInfo(stream_, &status, nullptr, &graph, &deps, &ndeps));
            cudaGraphNodeParams np = {cudaGraphNodeTypeConditional};
            np.type = cudaGraphNodeTypeConditional;
            np.conditional.handle = cond;
            np.conditional.type = cudaGraphCondTypeWhile;
            np.conditional.size = 1;
            cudaGraphNode_t node;
            CK(cudaGraphAddNode(&node, graph, deps, ndeps, &np));
            cudaGraph_t body = np.conditional.phGraph_out[0];
            CK(cudaStreamUpdateCaptureDependencies(stream_, &node, 1, cudaStreamSetCaptureDependencies));
            CK(cudaStreamBeginCaptureToGraph(stream2_, body, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal));
            launch_iteration(stream2_, slot, A, dinv, x, cond, 1);
            CK(cudaStreamEndCapture(stream2_, nullptr));
        } else {
            cudaGraphConditionalHandle none = 0;
            ph_cg_begin(stream_, slot, x, none, 0);
            launches_ += 1;
            const int batch = 8;
            const int64_t mi = maxit();
            for (int64_t done_iters = 0; done_iters < mi + batch;) {
                for (int b = 0; b < batch; b++) launch_iteration(stream_, slot, A, dinv, x, none, 0);
                launches_ += kernels_per_iteration() * batch;
                done_iters += batch;
                CK(cudaMemcpyAsync(h_flag_, &st->done, sizeof(int), cudaMemcpyDeviceToHost, stream_));
                CK(cudaStreamSynchronize(stream_));
                if (*h_flag_) break;
            }
        }
        CK(cudaGetLastError());
    }

    void enqueue_bicgstab(const R* A, const R* dinv, R* x, bool capturing);  // bicgstab driver below

    void enqueue_step(bool capturing) {
        if (!capturing) mark(1);
        ph_step_begin(stream_, capturing);
        ph_step_collect(stream_, capturing);
        if (!capturing) mark(2);
        if (use_bicgstab()) {
            solve_kernel_ = persistent_solves(capturing) ? "bicg_persistent" : capturing ? "graph" : "host";
            if (persistent_solves(capturing))
                launch_bicg_persistent();
            else
                enqueue_bicgstab(d_A0_.p, d_dinv0_.p, d_u_.p, capturing);
        } else {
            const R* A[2];
            const R* dinv[2];
            R* x[2];
            const int ns = solves(A, dinv, x);
            dual_last_ = ns == 2 && dual_applies(capturing);
            solve_kernel_ = capturing ? "graph" : dual_last_ ? "pcg_dual + pcg_persistent"
                            : opt_.loop_mode == FEAB_LOOP_PERSISTENT && !lockstep_ ? (ns == 2 ? "pcg_persistent x2" : "pcg_persistent") : "host";
            if (dual_last_)
                enqueue_dual(A, dinv, x);
            else
                for (int slot = 0; slot < ns; slot++) enqueue_pcg(slot, A, dinv, x, capturing, ns);
        }
        if (!capturing) mark(3);
        ph_finish(stream_, capturing);
        if (!capturing) mark(4);
    }

    void drop_graphs() {
        if (graph_exec_) cudaGraphExecDestroy(graph_exec_);
        if (graph_) cudaGraphDestroy(graph_);
        graph_exec_ = nullptr;
        graph_ = nullptr;
    }

    void build_graph() {
        drop_graphs();
        CK(cudaStreamBeginCapture(stream_, cudaStreamCaptureModeThreadLocal));
        try {
            enqueue_step(true);
        } catch (...) {
            cudaGraph_t g = nullptr;
            cudaStreamEndCapture(stream_, &g);
            if (g) cudaGraphDestroy(g);
            cudaGetLastError();
            throw;
        }
        CK(cudaStreamEndCapture(stream_, &graph_));
        CK(cudaGraphInstantiate(&graph_exec_, graph_, 0));
        graph_eq_ = eq_;
    }

    void require_ready_for_step() {
        REQUIRE(have_mesh_, FEAB_ENOMESH, "step: no mesh");
        if (world_ > 1) {
            for (int q = 0; q < world_; q++) REQUIRE(connected_[q], FEAB_EINVAL, "step: partitioned handle is not connected to every peer (feab_mg_connect)");
            REQUIRE(!use_bicgstab_legacy(), FEAB_EINVAL,
                    "step: on a partitioned mesh BiCGSTAB needs the persistent loop driver (FEAB_LOOP_PERSISTENT); the WHILE-graph / host-batched "
                    "BiCGSTAB kernels have no halo exchange");
            REQUIRE(n_ > 0, FEAB_EINVAL, "step: this rank owns no unknowns");
        }
    }
    void check_comm_error() {
        // One GPU too: the grid syncs of the persistent kernels are bounded spins of their own (comm.cuh kind 6) and fail closed
        // exactly like a lost peer does -- without this check a timed-out solve would leave the handle frozen and silent.
        CommLocal l;
        CK(cudaMemcpy(&l, d_loc_, sizeof(l), cudaMemcpyDeviceToHost));
        if (l.error) {
            // the grid-sync / ticket counters of the kernels that gave up are meaningless now: the monotone one-GPU counter
            // stopped short of a sync, so it restarts from zero together with its base (the next kernel reads gs_base first)
            l.error = 0;
            l.ticket = 0;
            l.gs_arrive = 0;
            l.gs_count = 0;
            l.gs_base = 0;
            const int kind = l.err_kind, peer = l.err_peer;
            const unsigned long long want = l.err_want;
            l.err_kind = 0;
            CK(cudaMemcpy(d_loc_, &l, sizeof(l), cudaMemcpyHostToDevice));
            static const char* names[] = {"?", "halo flag", "state flag", "full-values flag", "root flag", "all-reduce mailbox word", "grid sync of this rank"};
            throw Err{FEAB_ECOMM, std::string("step: rank ") + std::to_string(rank_) + " timed out waiting for " + names[kind >= 0 && kind <= 6 ? kind : 0] +
                                      " #" + std::to_string(want) + " of rank " + std::to_string(peer) +
                                      "; the solve was abandoned, values were not written back (feab_clear / reconnect before stepping again)"};
        }
    }
    // surface->values complete on every rank: owners broadcast, everybody waits for everybody
    // root_only (FEAB_SYNC_ROOT): the owners store into rank 0's mirror only and only rank 0 waits
    void ph_broadcast_values(cudaStream_t s, bool root_only) {
        if (world_ == 1) return;
        const int g = std::max(1, std::min(cdiv64(n_, kBlock), sms_ * 4));
        values_broadcast<<<g, kBlock, 0, s>>>(cm_, hp_, node_of_col_p_, n_, d_values_.p, root_only ? 1 : 0);
        launches_ += 1;
        if (!root_only || rank_ == 0) {
            wait_kernel<<<1, 32, 0, s>>>(cm_, hp_, 2);
            launches_ += 1;
        }
    }

    void step(float* values, int nsteps, int sync) override {
        require_ready_for_step();
        REQUIRE(nsteps >= 0, FEAB_EINVAL, "step: nsteps < 0");
        REQUIRE(!lockstep_, FEAB_EINVAL, "step: this handle belongs to a lock-step group (use feab_mg_group_step)");
        CK(cudaSetDevice(device_));
        CK(cudaEventRecord(ev0_, stream_));
        // FEAB_SYNC_ROOT on a partitioned handle: only rank 0 touches PCIe; its array reaches the peers over NVLink.
        // Safe against the previous call: rank 0 gets here only after its wait for every peer's values_broadcast of
        // that call, i.e. after the peers' last read of their mirrors.
        const bool root_mode = world_ > 1 && sync == FEAB_SYNC_ROOT && values != nullptr;
        if (values) values_dirty_ = true;
        // instability stamps of this call: [call_first_stamp_, step_stamp_].  The stamp counter restarts (with a cleared flag)
        // BEFORE a call that would run past INT32_MAX, never inside one: atomicMax would keep a pre-wrap value for ever.
        call_first_stamp_ = 0;
        if ((int64_t)step_stamp_ + nsteps >= INT32_MAX - 1) {
            CK(cudaMemsetAsync(d_unstable_, 0, sizeof(int), stream_));
            step_stamp_ = 0;
        }
        mark_now_ = true;
        mark(0);
        mark_now_ = false;
        if (root_mode) {
            if (rank_ == 0) {
                CK(cudaMemcpyAsync(d_values_.p, values, V_ * sizeof(float), cudaMemcpyHostToDevice, stream_));
                root_signal<<<1, 32, 0, stream_>>>(cm_);
            } else {
                // the values of this rank's columns (own rows + ghosts) come from rank 0's mirror over NVLink
                root_wait<<<1, 32, 0, stream_>>>(cm_);
                pull_cols<<<std::max(1, std::min({cdiv64(n_cols_, kBlock), sms_ * 4, grid_cap_})), kBlock, 0, stream_>>>(
                    reinterpret_cast<const float*>(hp_.base[0] + hp_.off_vec[kVecValues][0]), node_of_col_p_, n_cols_, d_values_.p, cm_);
                launches_++;
            }
            launches_++;
        } else if (values) {
            CK(cudaMemcpyAsync(d_values_.p, values, V_ * sizeof(float), cudaMemcpyHostToDevice, stream_));
        }
        if (n_ > 0 && nsteps > 0) {
            // Fault injection for the tests (FEAB_FAULT_GSYNC=k at feab_create): before the k-th stepping call of this handle one
            // arrival is taken away from the one-GPU grid-sync counter (its base moves up by one), i.e. one CTA "never arrives":
            // the first grid sync of that call runs into its bounded wait and the call must end in FEAB_ECOMM.
            if (fault_gsync_call_ > 0 && ++step_calls_ == fault_gsync_call_) {
                CK(cudaStreamSynchronize(stream_));
                CommLocal l;
                CK(cudaMemcpy(&l, d_loc_, sizeof(l), cudaMemcpyDeviceToHost));
                l.gs_base += 1;
                CK(cudaMemcpy(d_loc_, &l, sizeof(l), cudaMemcpyHostToDevice));
            }
            ensure_operator();
            ph_push_state(stream_);
            // a device without cooperative launch runs BiCGSTAB through the WHILE graph
            const bool graph_mode = opt_.loop_mode == FEAB_LOOP_GRAPH_WHILE || (opt_.loop_mode == FEAB_LOOP_PERSISTENT && use_bicgstab_legacy());
            graph_mode_last_ = graph_mode;
            if (graph_mode) {
                dual_last_ = false;  // a graph replays separate solves
                last_stamp_ = 1;     // ... and one instability stamp: cleared here once per call, sticky over the call's graphs
                call_first_stamp_ = 1;
                CK(cudaMemsetAsync(d_unstable_, 0, sizeof(int), stream_));
            }
            if (graph_mode && (!graph_exec_ || graph_eq_ != eq_)) build_graph();
            cluster_steps_last_ = 0;
            if (!graph_mode && use_cluster()) {
                mark_now_ = true;
                mark(1);
                mark(2);
                launch_cluster_steps(nsteps);
                mark(3);
                mark(4);
            }
            for (int s = 0; s < nsteps && cluster_steps_last_ == 0; s++) {
                time_solves_ = !graph_mode && opt_.loop_mode == FEAB_LOOP_PERSISTENT && s == nsteps - 1;
                mark_now_ = !graph_mode && s == nsteps - 1;
                if (graph_mode) {
                    solve_kernel_ = "graph";
                    CK(cudaGraphLaunch(graph_exec_, stream_));
                    graph_steps_++;
                } else {
                    enqueue_step(false);
                }
            }
        }
        const bool stepped = mark_now_;
        mark_now_ = true;
        if (!stepped) {
            for (int i = 1; i <= 4; i++) mark(i);
        }
        if (values && sync) {
            ph_broadcast_values(stream_, root_mode);
            mark(5);
            if (!root_mode || rank_ == 0) CK(cudaMemcpyAsync(values, d_values_.p, V_ * sizeof(float), cudaMemcpyDeviceToHost, stream_));
        } else {
            mark(5);
        }
        mark(6);
        mark_now_ = false;
        CK(cudaEventRecord(ev1_, stream_));
        CK(cudaStreamSynchronize(stream_));
        if (step_prof_) {
            for (int i = 0; i < 6; i++) {
                float t = 0;
                phase_ms_[i] = cudaEventElapsedTime(&t, ev_ph_[i], ev_ph_[i + 1]) == cudaSuccess ? t : 0.0;
            }
            cudaGetLastError();
        }
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, ev0_, ev1_));
        last_ms_ = ms;
        last_steps_ = nsteps;
        if (n_ > 0 && nsteps > 0 && graph_mode_last_) account_graph_launches(nsteps);
        solve_ms_ = 0;
        if (cluster_steps_last_ > 0) {  // the cluster kernel holds whole steps: its time per step
            float t = 0;
            if (cudaEventElapsedTime(&t, ev_solve_[0], ev_solve_[1]) == cudaSuccess) solve_ms_ = t / cluster_steps_last_;
            cudaGetLastError();
        }
        if (time_solves_) {  // device time of the last step's persistent PCG kernels
            const int ns = eq_ == FEAB_EQ_REACTION_DIFFUSION && !dual_applies(false) ? 2 : 1;  // the dual path brackets both solves with one pair
            for (int sl = 0; sl < ns; sl++) {
                float t = 0;
                if (cudaEventElapsedTime(&t, ev_solve_[2 * sl], ev_solve_[2 * sl + 1]) == cudaSuccess) solve_ms_ += t;
            }
            cudaGetLastError();
            time_solves_ = false;
        }
        if (cluster_steps_last_ > 0 && h_flag_[8]) {  // pair mode of pcg_cluster: the other cluster never arrived
            const int who = h_flag_[8];
            h_flag_[8] = 0;
            CK(cudaMemset(&d_loc_->error, 0, sizeof(int)));
            // the launch was accounted for in full (xbar_count_ += 4 nsteps) but the device counter stopped short: start both
            // from zero again, or every later pair launch would wait for arrivals that never come
            CK(cudaMemset(d_xbar_, 0, sizeof(unsigned long long)));
            xbar_count_ = 0;
            throw Err{FEAB_ECOMM, "step: cluster " + std::to_string(who - 1) + " of the small-mesh kernel timed out waiting for the other cluster; "
                                  "the fields of this call are not valid (FEAB_CLUSTER_PAIR=0 runs the two solves in one cluster)"};
        }
        check_comm_error();
    }

    void read_states(CgState<R>* st2, BicgState<R>* bst) {
        CK(cudaMemcpy(st2, d_st_, 2 * sizeof(CgState<R>), cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(bst, d_bst_, sizeof(BicgState<R>), cudaMemcpyDeviceToHost));
    }
    // Kernel count of graph-launched steps: fixed nodes + 3 per PCG loop trip.  Only the last step's trip count is
    // known exactly (device state); it is used for all steps of the call -- an estimate, stated as such in DESIGN.md.
    void account_graph_launches(int nsteps) {
        CgState<R> st[2];
        BicgState<R> b;
        read_states(st, &b);
        int64_t per_step = world_ > 1 ? 5 : 3;  // gather, rhs, finish (+ state wait, collect)
        if (use_bicgstab()) {
            per_step += 1 + (int64_t)kBicgKernelsPerIter * b.trips;
        } else {
            per_step += 1 + kernels_per_iteration() * (int64_t)st[0].iter1;
            if (eq_ == FEAB_EQ_REACTION_DIFFUSION) per_step += 1 + kernels_per_iteration() * (int64_t)st[1].iter1;
        }
        launches_ += per_step * nsteps;
    }

    void stats(FeabStats* s) override {
        CK(cudaSetDevice(device_));
        std::memset(s, 0, sizeof(*s));
        CgState<R> st[2];
        BicgState<R> b;
        read_states(st, &b);
        if (use_bicgstab()) {
            s->iterations[0] = b.iters;
            s->spmv_passes[0] = 2 * (int64_t)b.trips;
            s->rel_residual[0] = b.rhs_norm > 0 ? (double)(b.r_norm / b.rhs_norm) : 0.0;
        } else {
            for (int i = 0; i < 2; i++) {
                s->iterations[i] = st[i].iters;
                s->spmv_passes[i] = st[i].iter1;
                s->rel_residual[i] = st[i].rhs2 > 0 ? std::sqrt((double)st[i].rn2 / (double)st[i].rhs2) : 0.0;
            }
            if (eq_ != FEAB_EQ_REACTION_DIFFUSION) {
                s->iterations[1] = 0;
                s->rel_residual[1] = 0;
                s->spmv_passes[1] = 0;
            }
        }
        s->last_call_ms = last_ms_;
        s->last_call_steps = last_steps_;
        s->kernel_launches = launches_;
        s->solve_ms = solve_ms_;
        s->shared_iterations = dual_last_ && !use_bicgstab() ? std::min(st[0].iter1, st[1].iter1) : 0;
    }

    // ---------------------------------------------------------------- measurement hook
    void time_kernel(int kernel, int reps, double* ms_out, double* bytes_out) override {
        REQUIRE(have_mesh_ && n_ > 0, FEAB_ENOMESH, "time_kernel: no mesh");
        REQUIRE(reps > 0 && kernel >= 0 && kernel <= 3, FEAB_EINVAL, "time_kernel: bad arguments");
        REQUIRE(world_ == 1, FEAB_EINVAL, "time_kernel: single-GPU handles only (partitioned kernels wait for their peers)");
        Comm one;            // a one-GPU communicator of its own ...
        one.loc = d_loc_;    // ... with the handle's flags behind it: every kernel reads CommLocal::error (fail closed)
        CK(cudaSetDevice(device_));
        ensure_operator();
        const double n = (double)n_, nnz = (double)nnz_, rb = sizeof(R);
        // make the loop state "running" with harmless scalars
        CgState<R> h;
        std::memset(&h, 0, sizeof(h));
        h.abs_new[0] = h.abs_new[1] = (R)1;
        h.thr = (R)0;
        h.iter1 = 1;
        h.maxit = INT32_MAX - 4;
        CK(cudaMemcpy(d_st_, &h, sizeof(h), cudaMemcpyHostToDevice));
        // p.Ap partials huge -> alpha ~ 0 (x, r unchanged); |r|^2 partials 1 -> not converged; r.z partials 0 -> beta = 0
        std::vector<R> init((size_t)g_part_ * 3, (R)0);
        for (int i = 0; i < g_part_; i++) {
            init[i] = Ops<R>::huge() / (R)(4 * g_part_);
            init[(size_t)g_part_ + i] = (R)1;
        }
        CK(cudaMemcpy(d_part_.p, init.data(), init.size() * sizeof(R), cudaMemcpyHostToDevice));
        const StepScalars sc = scalars();
        cudaGraphConditionalHandle none = 0;
        auto launch = [&]() {
            switch (kernel) {
                case 0:
                    launch_spmv(stream_, d_A0_.p, d_p_[0].p, d_y_.p, nullptr);
                    break;
                case 1:
                    cg_update<R><<<g_vec_, kBlock, 0, stream_>>>(d_w_[0].p, d_r_[0].p, d_p_[0].p, d_y_.p, d_dinv0_.p, n_, d_st_, part_pap(), spmv_parts(),
                                                                 part_rr(), part_rz(), one);
                    break;
                case 2:
                    cg_direction<R><<<g_vec_, kBlock, 0, stream_>>>(d_r_[0].p, d_p_[0].p, d_dinv0_.p, n_, d_st_, part_rr(), part_rz(), g_vec_, none, 0, one);
                    break;
                default:
                    launch_rhs(stream_, sc, part_rhs(), one, 0);
                    break;
            }
        };
        for (int i = 0; i < 3; i++) launch();
        CK(cudaGetLastError());
        CK(cudaEventRecord(ev0_, stream_));
        for (int i = 0; i < reps; i++) launch();
        CK(cudaEventRecord(ev1_, stream_));
        CK(cudaStreamSynchronize(stream_));
        CK(cudaGetLastError());
        launches_ += reps + 3;
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, ev0_, ev1_));
        *ms_out = ms / reps;
        // algorithmic bytes
        double bytes = 0;
        switch (kernel) {
            case 0: bytes = nnz * (rb + 4) + (n + 1) * 4 + 2 * n * rb; break;
            case 1: bytes = 7 * n * rb; break;
            case 2: bytes = 4 * n * rb; break;
            default:
                if (eq_ == FEAB_EQ_REACTION_DIFFUSION)
                    bytes = nnz * (rb + 4) + 4 * n + 4 * n * rb + 4 * n * rb;  // + r0,p0,r1,p1 written
                else if (eq_ == FEAB_EQ_WAVE)
                    bytes = nnz * (2 * rb + 4) + 4 * n + 3 * n * rb + 2 * n * rb;
                else
                    bytes = nnz * (rb + 4) + 4 * n + 2 * n * rb + 2 * n * rb;
                break;
        }
        *bytes_out = bytes;
        // leave a clean state behind
        CK(cudaMemset(d_st_, 0, 2 * sizeof(CgState<R>)));
    }

    // FEAB_PCG_PROF=3: when did the CTAs of pcg_dual enter and leave the three grid syncs of the last iteration?  (stderr)
    void report_arrivals() {
        if (!prof_arrivals_ || arrivals_grid_ <= 0) return;
        const int G = arrivals_grid_;
        std::vector<unsigned long long> h((size_t)6 * G);
        CK(cudaMemcpy(h.data(), d_arrivals_.p, h.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
        static const char* names[3] = {"sync1 (after SpMV)", "sync2 (after update)", "sync3 (after direction)"};
        for (int k = 0; k < 3; k++) {
            std::vector<unsigned long long> in(h.begin() + (size_t)(2 * k) * G, h.begin() + (size_t)(2 * k + 1) * G);
            std::vector<unsigned long long> out(h.begin() + (size_t)(2 * k + 1) * G, h.begin() + (size_t)(2 * k + 2) * G);
            std::sort(in.begin(), in.end());
            std::sort(out.begin(), out.end());
            const double first = (double)in.front();
            fprintf(stderr, "[feab] rank %d %s, %d CTAs: arrivals after the first one (us): p10 %.2f p50 %.2f p90 %.2f p99 %.2f last %.2f | first leaves %.2f, "
                            "last leaves %.2f after the last arrival\n",
                    rank_, names[k], G, (in[G / 10] - first) / 1e3, (in[G / 2] - first) / 1e3, (in[G * 9 / 10] - first) / 1e3, (in[G * 99 / 100] - first) / 1e3,
                    (in.back() - first) / 1e3, ((double)out.front() - (double)in.back()) / 1e3, ((double)out.back() - (double)in.back()) / 1e3);
        }
    }

    void pcg_profile(double* out16, int reset) override {
        for (int i = 0; i < 16; i++) out16[i] = 0;
        report_arrivals();
        if (step_prof_)
            for (int i = 0; i < 6; i++) out16[10 + i] = phase_ms_[i];
        if (!d_prof_) return;
        CK(cudaSetDevice(device_));
        unsigned long long h[16];
        CK(cudaMemcpy(h, d_prof_, sizeof(h), cudaMemcpyDeviceToHost));
        for (int i = 0; i < 10; i++) out16[i] = (double)h[i];   // [10, 16) are the step-phase times above
        if (reset) CK(cudaMemset(d_prof_, 0, sizeof(h)));
    }

    // ---------------------------------------------------------------- partitioned solve: configuration
    void mg_config(int rank, int world) override {
        REQUIRE(world >= 1 && world <= kMaxRanks && rank >= 0 && rank < world, FEAB_EINVAL, "mg_config: need 0 <= rank < world <= 8");
        REQUIRE(!have_mesh_, FEAB_EINVAL, "mg_config: must be called before feab_set_mesh");
        rank_ = rank;
        world_ = world;
    }
    void mg_set_partition(const int32_t* part, int64_t V) override {
        REQUIRE(part && V > 0, FEAB_EINVAL, "mg_set_partition: bad arguments");
        REQUIRE(!have_mesh_, FEAB_EINVAL, "mg_set_partition: must be called before feab_set_mesh");
        for (int64_t i = 0; i < V; i++) REQUIRE(part[i] >= 0 && part[i] < world_, FEAB_EINVAL, "mg_set_partition: part id out of range");
        h_part_.assign(part, part + V);
        part_given_ = true;
    }
    void mg_export(FeabMgExport* e) override {
        REQUIRE(have_mesh_ && e, FEAB_ENOMESH, "mg_export: no mesh");
        CK(cudaSetDevice(device_));
        std::memset(e, 0, sizeof(*e));
        cudaIpcMemHandle_t hd;
        if (cudaIpcGetMemHandle(&hd, d_arena_.p) == cudaSuccess)
            std::memcpy(e->ipc_handle, &hd, sizeof(hd));
        else
            cudaGetLastError();  // same-process connections do not need it
        e->base_ptr = (uint64_t)(uintptr_t)d_arena_.p;
        e->arena_bytes = d_arena_.count;
        for (int v = 0; v < 5; v++) e->off_vec[v] = off_vec_[v];
        e->n_own = n_;
        e->n_ghost = n_ghost_;
        e->ghost_base = ghost_base_;
        e->num_nodes = V_;
        e->rank = rank_;
        e->world = world_;
        e->device = device_;
        e->pid = (int32_t)getpid();
    }
    void mg_connect(int q, const FeabMgExport* e, int same_process) override {
        REQUIRE(have_mesh_ && e, FEAB_ENOMESH, "mg_connect: no mesh");
        REQUIRE(world_ > 1 && q >= 0 && q < world_ && q != rank_, FEAB_EINVAL, "mg_connect: bad peer rank");
        REQUIRE(e->rank == q && e->world == world_ && e->num_nodes == V_, FEAB_EINVAL, "mg_connect: export does not belong to that peer / mesh");
        CK(cudaSetDevice(device_));
        char* base = nullptr;
        if (same_process) {
            base = reinterpret_cast<char*>((uintptr_t)e->base_ptr);
            if (e->device != device_) {
                cudaError_t pe = cudaDeviceEnablePeerAccess(e->device, 0);
                if (pe != cudaSuccess && pe != cudaErrorPeerAccessAlreadyEnabled) CK(pe);
                cudaGetLastError();
            }
        } else {
            if (ipc_opened_[q]) {
                cudaIpcCloseMemHandle(ipc_opened_[q]);
                ipc_opened_[q] = nullptr;
            }
            cudaIpcMemHandle_t hd;
            std::memcpy(&hd, e->ipc_handle, sizeof(hd));
            void* ptr = nullptr;
            CK(cudaIpcOpenMemHandle(&ptr, hd, cudaIpcMemLazyEnablePeerAccess));
            ipc_opened_[q] = ptr;
            base = static_cast<char*>(ptr);
        }
        for (int b = 0; b < hp_.nnb; b++)
            if (hp_.nb_rank[b] == q) {
                const int64_t cnt = hp_.send_off[b + 1] - hp_.send_off[b];
                REQUIRE(hp_.dst_col[b] >= e->ghost_base && hp_.dst_col[b] + cnt <= e->ghost_base + e->n_ghost, FEAB_EINVAL,
                        "mg_connect: peer's ghost layout disagrees with this rank's send list (different mesh or partition?)");
            }
        cm_.ctrl[q] = reinterpret_cast<CommCtrl*>(base);
        hp_.base[q] = base;
        for (int v = 0; v < 5; v++) hp_.off_vec[v][q] = e->off_vec[v];
        connected_[q] = true;
        drop_graphs();  // the step graph bakes Comm / HaloPlan into its kernel nodes
    }
    void mg_get_layout(int64_t* n_own, int64_t* n_int, int64_t* n_ghost, int64_t* n_cols) override {
        REQUIRE(have_mesh_, FEAB_ENOMESH, "mg_get_layout: no mesh");
        if (n_own) *n_own = n_;
        if (n_int) *n_int = n_int_;
        if (n_ghost) *n_ghost = n_ghost_;
        if (n_cols) *n_cols = n_cols_;
    }
    void mg_get_partition(int32_t* part) override {
        REQUIRE(have_mesh_ && part, FEAB_ENOMESH, "mg_get_partition: no mesh");
        CK(cudaSetDevice(device_));
        if (world_ > 1)
            CK(cudaMemcpy(part, d_part_of_.p, V_ * sizeof(int), cudaMemcpyDeviceToHost));
        else
            std::memset(part, 0, V_ * sizeof(int));
    }
    void mg_get_maps(int32_t* row_node, int32_t* col_unknown) override {
        REQUIRE(have_mesh_, FEAB_ENOMESH, "mg_get_maps: no mesh");
        CK(cudaSetDevice(device_));
        if (row_node) CK(cudaMemcpy(row_node, node_of_col_p_, n_ * sizeof(int), cudaMemcpyDeviceToHost));
        if (col_unknown) {
            if (world_ > 1) {
                CK(cudaMemcpy(col_unknown, d_glob_of_col_.p, n_cols_ * sizeof(int), cudaMemcpyDeviceToHost));
            } else {
                for (int64_t i = 0; i < n_cols_; i++) col_unknown[i] = (int32_t)i;
            }
        }
    }

    // ---------------------------------------------------------------- lock-step driving of virtual ranks
    cudaStream_t ls_prepare(cudaStream_t shared) override {
        require_ready_for_step();
        REQUIRE(!use_bicgstab(), FEAB_EINVAL, "group_step: BiCGSTAB is not supported");
        CK(cudaSetDevice(device_));
        CK(cudaStreamSynchronize(stream_));
        ls_stream_ = shared ? shared : stream_;
        lockstep_ = true;
        call_first_stamp_ = 0;
        ensure_operator_on(ls_stream_);
        return ls_stream_;
    }
    void ls_phase(int phase, int slot) override {
        const R* A[2];
        const R* dinv[2];
        R* x[2];
        solves(A, dinv, x);
        cudaGraphConditionalHandle none = 0;
        cudaStream_t s = ls_stream_;
        switch (phase) {
            case LS_PUSH_STATE: ph_push_state(s); break;
            case LS_BEGIN: ph_step_begin(s, false); break;
            case LS_COLLECT: ph_step_collect(s, false); break;
            case LS_CG_BEGIN: ph_cg_begin(s, slot, x[slot], none, 0); launches_++; break;
            case LS_SPMV: ph_spmv(s, slot, A[slot]); break;
            case LS_UPDATE: ph_update(s, slot, dinv[slot], x[slot]); break;
            case LS_DIRECTION: ph_direction(s, slot, dinv[slot], none, 0); launches_ += kernels_per_iteration(); break;
            case LS_FINISH: ph_finish(s, false); break;
            default: throw Err{FEAB_EINVAL, "ls_phase: bad phase"};
        }
        CK(cudaGetLastError());
    }
    int ls_num_solves() override { return eq_ == FEAB_EQ_REACTION_DIFFUSION ? 2 : 1; }
    int ls_poll_done(int slot) override {
        CK(cudaMemcpyAsync(h_flag_, &(d_st_ + slot)->done, sizeof(int), cudaMemcpyDeviceToHost, ls_stream_));
        CK(cudaStreamSynchronize(ls_stream_));
        return *h_flag_;
    }
    int64_t ls_maxit() override { return maxit(); }
    void ls_finish_call(int nsteps) override {
        CK(cudaStreamSynchronize(ls_stream_));
        lockstep_ = false;
        last_steps_ = nsteps;
        last_ms_ = 0;
        check_comm_error();
    }
    int device() const override { return device_; }
    const char* last_solve_kernel() const override { return solve_kernel_; }

    int sms() const { return sms_; }
    const cudaDeviceProp& prop() const { return prop_; }

   private:
    int device_;
    cudaDeviceProp prop_;
    int sms_ = 0;
    cudaStream_t stream_ = nullptr, stream2_ = nullptr;
    cudaEvent_t ev0_ = nullptr, ev1_ = nullptr;
    int* h_flag_ = nullptr;

    bool have_mesh_ = false;
    int64_t V_ = 0, T_ = 0, n_ = 0, nnz_ = 0;
    int bc_ = FEAB_BC_DIRICHLET;  // include/FEM/FEMContext.hpp:32
    int eq_ = FEAB_EQ_WAVE;       // include/FEM/FEMContext.hpp:31
    int max_row_ = 0;
    float vel_[3] = {1.0f, 0.0f, 0.0f};  // include/FEM/EquationParameters.hpp:24
    FeabParams par_[4];
    FeabSolverOptions opt_;

    DevBuf<float> d_xyz_, d_values_;
    DevBuf<uint32_t> d_tri_;
    DevBuf<uint8_t> d_onb_, d_advp_;
    DevBuf<int> d_idx_, d_node_of_, d_rowptr_, d_col_, d_diag_, d_smap_, d_cptr_, d_corner_;
    DevBuf<long long> d_scan_bsum_, d_scan_total_, d_brush_;
    DevBuf<R> d_M_, d_K_, d_Adv_, d_AdvT_, d_Md_, d_A0_, d_A1_, d_dinv0_, d_dinv1_;
    DevBuf<R> d_u_, d_v_, d_y_, d_r_[2], d_p_[2], d_w_[5], d_part_, d_rhs_tot_;

    // ---- partition (world_ > 1; with world_ == 1 the maps alias idx_map / node_of) ----
    int rank_ = 0, world_ = 1;
    std::vector<int> h_part_;          // part of every mesh vertex (all ranks hold the same array)
    bool part_given_ = false;
    DevBuf<int> d_part_of_, d_rowmap_, d_colmap_, d_node_of_col_, d_glob_of_col_, d_send_rows_, d_send_nodes_, d_push_tab_, d_dir_nodes_;
    int64_t n_dir_ = 0;                  // Dirichlet nodes (no unknown) of the whole mesh
    const int* rowmap_p_ = nullptr;      // per node: local row or -1
    const int* colmap_p_ = nullptr;      // per node: local column (row or ghost) or -1
    const int* node_of_col_p_ = nullptr; // per local column: mesh node
    int64_t nglob_ = 0;                  // unknowns of the whole mesh
    int64_t n_int_ = 0, n_ghost_ = 0, ghost_base_ = 0, n_cols_ = 0;
    DevBuf<unsigned char> d_arena_;      // CommCtrl | p[0] | p[1] | u | v | values  (peers map this allocation)
    size_t off_vec_[5] = {0, 0, 0, 0, 0};
    CommLocal* d_loc_ = nullptr;
    Comm cm_;
    HaloPlan hp_;
    bool connected_[kMaxRanks] = {};
    void* ipc_opened_[kMaxRanks] = {};
    bool lockstep_ = false;
    cudaStream_t ls_stream_ = nullptr;   // shared stream while driven phase by phase by feab_mg_group_step
    CgState<R>* d_st_ = nullptr;
    BicgState<R>* d_bst_ = nullptr;
    int* d_unstable_ = nullptr;

    int num_tiles_ = 1, g_tile_ = 1, g_vec_ = 1, g_part_ = 1;
    int spmv_variant_ = 4, prefetch_x_ = 0;  // 4 = TMA pipeline, 512-row tiles, 2 stages, 2 CTAs/SM (profiles/r01_spmv_sweep.txt)
    bool op_valid_ = false;
    int op_eq_ = -1, op_transposed_ = 0;
    FeabParams op_par_{};

    cudaGraph_t graph_ = nullptr;
    cudaGraphExec_t graph_exec_ = nullptr;
    int graph_eq_ = -1;
    int64_t graph_steps_ = 0;

    // persistent whole-solve kernel
    int pcg_max_grid_ = 0;
    int dual_max_grid_ = 0;
    const char* solve_kernel_ = "";      // feab_last_solve_kernel
    bool prof_arrivals_ = false;         // FEAB_PCG_PROF=3
    DevBuf<unsigned long long> d_arrivals_;
    int arrivals_grid_ = 0;
    int bicg_max_grid_ = 0;
    int cluster_ctas_ = 0;      // largest cluster size pcg_cluster can be launched with (0: not available / switched off)
    int cluster_steps_last_ = 0;
    bool cluster_pair_ = false;  // two clusters of that size can be resident together (FEAB_CLUSTER_PAIR=0: never use the second one)
    unsigned long long* d_xbar_ = nullptr;  // arrivals at pcg_cluster's inter-cluster barriers
    unsigned long long xbar_count_ = 0;     // ... and how many the launches so far account for
    int cluster_threads_ = 512;   // 16 warps per CTA and 128 registers per thread measured faster than 768 x 80 and 1024 x 64 (FEAB_CLUSTER_THREADS=1024)
    int64_t cluster_ent_key_ = -1;
    int cluster_ent_max_ = 0;
    int grid_cap_ = INT32_MAX;
    int fault_gsync_call_ = 0, step_calls_ = 0;  // FEAB_FAULT_GSYNC (test hook)
    bool coop_launch_ = true;  // FEAB_COOP=0: plain launches of the persistent kernels (launch_resident)
    int zigzag_ = -1;    // FEAB_ZIGZAG: order of phase C of the persistent PCG kernels (PcgArgs::zigzag); -1 = phase_c_order() decides
    bool dual_ = true;   // FEAB_DUAL=0 runs the two Gray-Scott solves one after the other
    bool dual_last_ = false;
    DevBuf<R> d_dpart_;
    bool rhs_tma_ = true;
    bool prof_dual_ = false;
    unsigned long long* d_prof_ = nullptr;  // FEAB_PCG_PROF=1: phase timers of the persistent kernel (CTA 0)
    DevBuf<R> d_ppart_;
    cudaEvent_t ev_solve_[4] = {};
    bool time_solves_ = false, graph_mode_last_ = false;
    // FEAB_STEP_PROF=1: device time of the pieces of the LAST step of a feab_step call (events on the solver stream):
    // [0] upload (+ root push), [1] gather + RHS kernel, [2] solve kernels, [3] epilogue, [4] value broadcast, [5] download
    bool step_prof_ = false, mark_now_ = false;
    cudaEvent_t ev_ph_[8] = {};
    double phase_ms_[6] = {0, 0, 0, 0, 0, 0};
    void mark(int i) {
        if (step_prof_ && mark_now_) cudaEventRecord(ev_ph_[i], stream_);
    }
    double solve_ms_ = 0;

    double last_ms_ = 0;
    int64_t last_steps_ = 0;
    int64_t launches_ = 0;
    bool values_dirty_ = true;            // surface->values was written by somebody other than finish_step since the last step
    int step_stamp_ = 0, last_stamp_ = 0; // instability flag stamps (finish_step)
    int call_first_stamp_ = 0;            // stamp of the first step of the last feab_step call (0: no step yet)
};

// ---------------------------------------------------------------- BiCGSTAB driver
template <typename R>
void SolverT<R>::enqueue_bicgstab(const R* A, const R* dinv, R* x, bool capturing) {
    // vectors: r = d_r_[0] (holds b from rhs_init), r0 = d_w_[0], v = d_w_[1], p = d_p_[0] (reset), y = d_w_[2],
    //          s = d_r_[1], z = d_w_[3], t = d_y_
    BicgVectors<R> bv;
    bv.x = x;
    bv.r = d_r_[0].p;
    bv.r0 = d_w_[0].p;
    bv.v = d_w_[1].p;
    bv.p = d_p_[0].p;
    bv.y = d_w_[2].p;
    bv.s = d_r_[1].p;
    bv.z = d_w_[3].p;
    bv.t = d_y_.p;
    bv.b = d_w_[4].p;
    bv.dinv = dinv;
    BicgGrid bg;
    bg.g_tile = g_tile_;
    bg.g_vec = g_vec_;
    bg.g_part = g_part_;
    bg.num_tiles = num_tiles_;
    R* part = part_bi();
    const size_t sm1 = kTileCap * sizeof(R);
    CK(cudaMemsetAsync(x, 0, n_ * sizeof(R), stream_));
    if (capturing) {
        cudaStreamCaptureStatus status;
        cudaGraph_t graph;
        const cudaGraphNode_t* deps;
        size_t ndeps;
        CK(cudaStreamGetCaptureInfo(stream_, &status, nullptr, &graph, &deps, &ndeps));
        cudaGraphConditionalHandle cond;
        CK(cudaGraphConditionalHandleCreate(&cond, graph, 0, 0));
        bicg_begin<R><<<g_vec_, kBlock, 0, stream_>>>(bv, n_, part, g_part_);
        bicg_begin2<R><<<1, kBlock, 0, stream_>>>(d_bst_, part, g_vec_, tol(), maxit(), cond, 1);
        CK(cudaStreamGetCaptureInfo(stream_, &status, nullptr, &graph, &deps, &ndeps));
        cudaGraphNodeParams np = {cudaGraphNodeTypeConditional};
        np.type = cudaGraphNodeTypeConditional;
        np.conditional.handle = cond;
        np.conditional.type = cudaGraphCondTypeWhile;
        np.conditional.size = 1;
        cudaGraphNode_t node;
        CK(cudaGraphAddNode(&node, graph, deps, ndeps, &np));
        cudaGraph_t body = np.conditional.phGraph_out[0];
        CK(cudaStreamUpdateCaptureDependencies(stream_, &node, 1, cudaStreamSetCaptureDependencies));
        CK(cudaStreamBeginCaptureToGraph(stream2_, body, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal));
        bicg_iteration<R>(stream2_, d_rowptr_.p, d_col_.p, A, bv, n_, bg, d_bst_, part, sm1, cond, 1);
        CK(cudaStreamEndCapture(stream2_, nullptr));
    } else {
        cudaGraphConditionalHandle none = 0;
        DBG(stream_, "bicg: before begin n=%lld g_vec=%d g_tile=%d g_part=%d", (long long)n_, g_vec_, g_tile_, g_part_);
        bicg_begin<R><<<g_vec_, kBlock, 0, stream_>>>(bv, n_, part, g_part_);
        DBG(stream_, "bicg: begin");
        bicg_begin2<R><<<1, kBlock, 0, stream_>>>(d_bst_, part, g_vec_, tol(), maxit(), none, 0);
        DBG(stream_, "bicg: begin2");
        launches_ += 2;
        const int batch = 4;
        const int64_t mi = maxit();
        for (int64_t trips = 0; trips < 2 * mi + batch;) {
            for (int b = 0; b < batch; b++) {
                bicg_iteration<R>(stream_, d_rowptr_.p, d_col_.p, A, bv, n_, bg, d_bst_, part, sm1, none, 0);
                DBG(stream_, "bicg: iteration trips=%lld b=%d", (long long)trips, b);
            }
            launches_ += (int64_t)kBicgKernelsPerIter * batch;
            trips += batch;
            CK(cudaMemcpyAsync(h_flag_, &d_bst_->done, sizeof(int), cudaMemcpyDeviceToHost, stream_));
            CK(cudaStreamSynchronize(stream_));
            if (*h_flag_) break;
        }
    }
    CK(cudaGetLastError());
}

}  // namespace feab

// =================================================================================== C ABI
struct FeabHandle {
    int device = 0;
    int real_bytes = 8;
    std::string err;
    std::unique_ptr<feab::SolverBase> impl;
    cudaDeviceProp prop;
    void* reg_ptr = nullptr;   // host range page-locked by feab_register_host_values (undone on re-registration / destroy)
    void unregister_host() {
        if (reg_ptr) {
            cudaSetDevice(device);
            cudaHostUnregister(reg_ptr);
            cudaGetLastError();
            reg_ptr = nullptr;
        }
    }
    ~FeabHandle() {
        impl.reset();          // the solver drains its stream first: no copy may still read the range
        unregister_host();
    }
};

static thread_local std::string g_create_err;

#define FEAB_TRY(h, body)                                 \
    if (!(h)) return FEAB_EINVAL;                         \
    try {                                                 \
        body;                                             \
        return FEAB_OK;                                   \
    } catch (const feab::Err& e) {                        \
        (h)->err = e.msg;                                 \
        return e.code;                                    \
    } catch (const std::exception& e) {                   \
        (h)->err = e.what();                              \
        return FEAB_EINVAL;                               \
    } catch (...) {                                       \
        (h)->err = "unknown error";                       \
        return FEAB_EINVAL;                               \
    }

extern "C" {

int feab_create(int device, int real_bytes, FeabHandle** out) {
    if (!out) return FEAB_EINVAL;
    *out = nullptr;
    if (real_bytes != 4 && real_bytes != 8) {
        g_create_err = "feab_create: real_bytes must be 4 or 8";
        return FEAB_EINVAL;
    }
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0 || device < 0 || device >= count) {
        g_create_err = std::string("feab_create: no usable CUDA device (") + (e != cudaSuccess ? cudaGetErrorString(e) : "bad ordinal") +
                       "); libfeab has no CPU fallback";
        cudaGetLastError();
        return FEAB_ENOGPU;
    }
    auto* h = new FeabHandle();
    h->device = device;
    h->real_bytes = real_bytes;
    try {
        cudaGetDeviceProperties(&h->prop, device);
        if (real_bytes == 4)
            h->impl.reset(new feab::SolverT<float>(device));
        else
            h->impl.reset(new feab::SolverT<double>(device));
    } catch (const feab::Err& er) {
        g_create_err = er.msg;
        int code = er.code;
        delete h;
        return code;
    }
    *out = h;
    return FEAB_OK;
}

int feab_destroy(FeabHandle* h) {
    if (!h) return FEAB_EINVAL;
    delete h;
    return FEAB_OK;
}

const char* feab_last_error(const FeabHandle* h) { return h ? h->err.c_str() : g_create_err.c_str(); }

int feab_set_mesh(FeabHandle* h, int64_t V, const float* xyz, int64_t T, const uint32_t* tri, const uint8_t* onb) {
    FEAB_TRY(h, h->impl->set_mesh(V, xyz, T, tri, onb));
}
int feab_set_mesh_ex(FeabHandle* h, int64_t V, const float* xyz, int64_t T, const uint32_t* tri, const uint8_t* onb, int bc, const float* vel) {
    FEAB_TRY(h, h->impl->set_mesh_ex(V, xyz, T, tri, onb, bc, vel));
}
int feab_set_boundary_condition(FeabHandle* h, int bc) { FEAB_TRY(h, h->impl->set_bc(bc)); }
int feab_assemble(FeabHandle* h, const float* velocity3) { FEAB_TRY(h, h->impl->assemble(velocity3)); }
int feab_get_counts(const FeabHandle* h, int64_t* V, int64_t* n, int64_t* nnz, int* maxrow) {
    FeabHandle* hh = const_cast<FeabHandle*>(h);
    FEAB_TRY(hh, hh->impl->get_counts(V, n, nnz, maxrow));
}
int feab_get_idx_map(FeabHandle* h, int32_t* out) { FEAB_TRY(h, h->impl->get_idx_map(out)); }
int feab_get_csr(FeabHandle* h, int which, int64_t* nnz, int32_t* rowptr, int32_t* col, void* val) {
    FEAB_TRY(h, h->impl->get_csr(which, nnz, rowptr, col, val));
}
int feab_set_equation(FeabHandle* h, int eq) { FEAB_TRY(h, h->impl->set_equation(eq)); }
int feab_set_params(FeabHandle* h, int eq, const FeabParams* p) { FEAB_TRY(h, h->impl->set_params(eq, p)); }
int feab_get_params(const FeabHandle* h, int eq, FeabParams* p) {
    FeabHandle* hh = const_cast<FeabHandle*>(h);
    FEAB_TRY(hh, hh->impl->get_params(eq, p));
}
int feab_set_solver_options(FeabHandle* h, const FeabSolverOptions* o) { FEAB_TRY(h, h->impl->set_options(o)); }
int feab_get_solver_options(const FeabHandle* h, FeabSolverOptions* o) {
    FeabHandle* hh = const_cast<FeabHandle*>(h);
    FEAB_TRY(hh, hh->impl->get_options(o));
}
int feab_clear(FeabHandle* h) { FEAB_TRY(h, h->impl->clear()); }
int feab_set_values(FeabHandle* h, const float* v) { FEAB_TRY(h, h->impl->set_values(v)); }
int feab_get_values(FeabHandle* h, float* v) { FEAB_TRY(h, h->impl->get_values(v)); }
int feab_values_device(FeabHandle* h, const float** out) {
    if (!out) return FEAB_EINVAL;
    FEAB_TRY(h, *out = h->impl->values_device());
}
static void register_host_values(FeabHandle* h, float* values, int64_t count) {
    h->unregister_host();
    if (!values) return;
    if (count <= 0) throw feab::Err{FEAB_EINVAL, "register_host_values: count <= 0"};
    cudaError_t e = cudaSetDevice(h->device);
    if (e == cudaSuccess) e = cudaHostRegister(values, (size_t)count * sizeof(float), cudaHostRegisterPortable);
    if (e != cudaSuccess) {
        cudaGetLastError();
        throw feab::Err{(int)e, std::string("register_host_values: cudaHostRegister: ") + cudaGetErrorString(e)};
    }
    h->reg_ptr = values;
}
int feab_register_host_values(FeabHandle* h, float* values, int64_t count) { FEAB_TRY(h, register_host_values(h, values, count)); }
int feab_brush(FeabHandle* h, const int64_t* nodes, int64_t count, float strength) { FEAB_TRY(h, h->impl->brush(nodes, count, strength)); }
int feab_step(FeabHandle* h, float* values, int nsteps, int sync) { FEAB_TRY(h, h->impl->step(values, nsteps, sync)); }
int feab_unstable(FeabHandle* h, int* out) {
    if (!out) return FEAB_EINVAL;
    FEAB_TRY(h, *out = h->impl->unstable());
}
int feab_last_stats(FeabHandle* h, FeabStats* s) {
    if (!s) return FEAB_EINVAL;
    FEAB_TRY(h, h->impl->stats(s));
}
int feab_get_state(FeabHandle* h, int which, void* out) { FEAB_TRY(h, h->impl->get_state(which, out)); }
int feab_set_state(FeabHandle* h, int which, const void* in) { FEAB_TRY(h, h->impl->set_state(which, in)); }
int feab_time_kernel(FeabHandle* h, int kernel, int reps, double* ms, double* bytes) {
    if (!ms || !bytes) return FEAB_EINVAL;
    FEAB_TRY(h, h->impl->time_kernel(kernel, reps, ms, bytes));
}
int feab_pcg_profile(FeabHandle* h, double* out16, int reset) {
    if (!out16) return FEAB_EINVAL;
    FEAB_TRY(h, h->impl->pcg_profile(out16, reset));
}
int feab_mg_config(FeabHandle* h, int rank, int world) { FEAB_TRY(h, h->impl->mg_config(rank, world)); }
int feab_mg_set_partition(FeabHandle* h, const int32_t* part, int64_t V) { FEAB_TRY(h, h->impl->mg_set_partition(part, V)); }
int feab_partition_rcb(int64_t V, const float* xyz, int parts, int32_t* part_out) {
    if (V <= 0 || !xyz || parts < 1 || !part_out) return FEAB_EINVAL;
    feab::rcb_partition(V, xyz, parts, part_out);
    return FEAB_OK;
}
int feab_mg_export(FeabHandle* h, FeabMgExport* out) { FEAB_TRY(h, h->impl->mg_export(out)); }
int feab_mg_connect(FeabHandle* h, int peer, const FeabMgExport* e, int same_process) { FEAB_TRY(h, h->impl->mg_connect(peer, e, same_process)); }
int feab_mg_get_layout(FeabHandle* h, int64_t* n_own, int64_t* n_int, int64_t* n_ghost, int64_t* n_cols) {
    FEAB_TRY(h, h->impl->mg_get_layout(n_own, n_int, n_ghost, n_cols));
}
int feab_mg_get_maps(FeabHandle* h, int32_t* row_node, int32_t* col_unknown) { FEAB_TRY(h, h->impl->mg_get_maps(row_node, col_unknown)); }
int feab_mg_get_partition(FeabHandle* h, int32_t* part) { FEAB_TRY(h, h->impl->mg_get_partition(part)); }

static void group_step_impl(FeabHandle** hs, int count, int nsteps) {
    FeabHandle* h0 = hs[0];
    using B = feab::SolverBase;
        for (int i = 1; i < count; i++)
            if (hs[i]->impl->device() != h0->impl->device()) throw feab::Err{FEAB_EINVAL, "group_step: all handles must live on one device"};
        // one stream for every virtual rank: stream order is the only synchronisation, no kernel ever spins
        cudaStream_t s = h0->impl->ls_prepare(nullptr);
        for (int i = 1; i < count; i++) hs[i]->impl->ls_prepare(s);
        auto all = [&](int phase, int slot) {
            for (int i = 0; i < count; i++) hs[i]->impl->ls_phase(phase, slot);
        };
        if (nsteps > 0) all(B::LS_PUSH_STATE, 0);
        for (int st = 0; st < nsteps; st++) {
            all(B::LS_BEGIN, 0);
            all(B::LS_COLLECT, 0);
            const int ns = h0->impl->ls_num_solves();
            for (int slot = 0; slot < ns; slot++) {
                all(B::LS_CG_BEGIN, slot);
                const int64_t mi = h0->impl->ls_maxit();
                const int batch = 8;
                for (int64_t it = 0; it < mi + batch; it += batch) {
                    for (int b = 0; b < batch; b++) {
                        all(B::LS_SPMV, slot);
                        all(B::LS_UPDATE, slot);
                        all(B::LS_DIRECTION, slot);
                    }
                    if (h0->impl->ls_poll_done(slot)) break;
                }
            }
            all(B::LS_FINISH, 0);
        }
        for (int i = 0; i < count; i++) hs[i]->impl->ls_finish_call(nsteps);
}

int feab_mg_group_step(FeabHandle** hs, int count, int nsteps) {
    if (!hs || count < 1 || nsteps < 0) return FEAB_EINVAL;
    for (int i = 0; i < count; i++)
        if (!hs[i]) return FEAB_EINVAL;
    FEAB_TRY(hs[0], group_step_impl(hs, count, nsteps));
}

int feab_last_solve_kernel(FeabHandle* h, char* name64) {
    if (!h || !name64) return FEAB_EINVAL;
    std::strncpy(name64, h->impl->last_solve_kernel(), 63);
    name64[63] = 0;
    return FEAB_OK;
}

int feab_device_info(const FeabHandle* h, int* sm_count, int64_t* total_mem, char* name64) {
    if (!h) return FEAB_EINVAL;
    if (sm_count) *sm_count = h->prop.multiProcessorCount;
    if (total_mem) *total_mem = (int64_t)h->prop.totalGlobalMem;
    if (name64) {
        std::strncpy(name64, h->prop.name, 63);
        name64[63] = 0;
    }
    return FEAB_OK;
}

}  // extern "C"
