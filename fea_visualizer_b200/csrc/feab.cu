// libfeab.so -- host driver and C ABI (include/feab.h) around the sm_100a kernels in pattern.cuh,
// assembly.cuh and solver.cuh.  One handle = one mesh on one GPU; everything stays resident in HBM.
// There is no CPU fallback: without a CUDA device feab_create fails with FEAB_ENOGPU.
#include "../../include/feab.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "assembly.cuh"
#include "bicgstab.cuh"
#include "common.cuh"
#include "pattern.cuh"
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
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    ~DevBuf() { release(); }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        count = 0;
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
    virtual void brush(const int64_t* nodes, int64_t count, float strength) = 0;
    virtual void step(float* values, int nsteps, int sync) = 0;
    virtual int unstable() = 0;
    virtual void stats(FeabStats* s) = 0;
    virtual void get_state(int which, void* out) = 0;
    virtual void set_state(int which, const void* in) = 0;
    virtual void time_kernel(int kernel, int reps, double* ms, double* bytes) = 0;
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
        CK(cudaMalloc(&d_st_, 2 * sizeof(CgState<R>)));
        CK(cudaMemset(d_st_, 0, 2 * sizeof(CgState<R>)));
        CK(cudaMalloc(&d_bst_, sizeof(BicgState<R>)));
        CK(cudaMemset(d_bst_, 0, sizeof(BicgState<R>)));
        d_scan_total_.alloc(1);
        if (const char* v = getenv("FEAB_SPMV")) spmv_variant_ = atoi(v);
        if (const char* v = getenv("FEAB_PREFETCH_X")) prefetch_x_ = atoi(v);
        setup_spmv_variants();
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
        spmv_dot_tma<R, ROWS, CPR, STAGES, MINB><<<grid, ROWS, Cfg::kSmemBytes, s>>>(d_rowptr_.p, d_col_.p, A, x, y, n_, tiles, st, part_pap(), prefetch_x_);
    }
    template <int ROWS, int MINB>
    int grid_tma() const {
        return std::min(std::max(1, cdiv64(n_, ROWS)), sms_ * MINB);
    }
    // number of p.Ap partials the SpMV of the current variant writes
    int spmv_parts() const {
        switch (spmv_variant_) {
            case 1: return grid_tma<256, 2>();
            case 2: return grid_tma<512, 1>();
            case 3: return grid_tma<256, 3>();
            case 4: return grid_tma<512, 2>();
            case 5: return grid_tma<1024, 1>();
            default: return g_tile_;
        }
    }
    void launch_spmv(cudaStream_t s, const R* A, const R* x, R* y, CgState<R>* st) {
        switch (spmv_variant_) {
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
        if (d_st_) cudaFree(d_st_);
        if (d_bst_) cudaFree(d_bst_);
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
        V_ = V;
        T_ = T;
        d_xyz_.alloc(3 * V);
        d_tri_.alloc(3 * std::max<int64_t>(T, 1));
        d_onb_.alloc(V);
        CK(cudaMemcpyAsync(d_xyz_.p, xyz, 3 * V * sizeof(float), cudaMemcpyHostToDevice, stream_));
        if (T) CK(cudaMemcpyAsync(d_tri_.p, tri, 3 * T * sizeof(uint32_t), cudaMemcpyHostToDevice, stream_));
        CK(cudaMemcpyAsync(d_onb_.p, onb, V, cudaMemcpyHostToDevice, stream_));
        d_values_.alloc(V);
        CK(cudaMemsetAsync(d_values_.p, 0, V * sizeof(float), stream_));
        CK(cudaStreamSynchronize(stream_));
        // validate triangle indices on the host copy (the reference would index out of bounds)
        for (int64_t i = 0; i < 3 * T; i++) REQUIRE(tri[i] < (uint64_t)V, FEAB_EINVAL, "set_mesh: triangle index out of range");
        have_mesh_ = true;
        symbolic();
        assemble(nullptr);
        clear();
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

    void symbolic() {
        drop_graphs();
        op_valid_ = false;
        const int64_t V = V_, T = T_;
        const int gV = cdiv64(V, kBlock), gT = std::max(1, cdiv64(T, kBlock));
        // --- idx_map (exclusive scan of the unknown flags) ---
        DevBuf<int> flag, excl;
        flag.alloc(V);
        excl.alloc(V + 1);
        unknown_flags<<<gV, kBlock, 0, stream_>>>(d_onb_.p, V, bc_, flag.p);
        long long n_ll = 0;
        exclusive_scan(flag.p, V, excl.p, &n_ll);
        n_ = n_ll;
        d_idx_.alloc(V);
        d_node_of_.alloc(std::max<int64_t>(n_, 1));
        make_idx_map<<<gV, kBlock, 0, stream_>>>(flag.p, excl.p, V, d_idx_.p, d_node_of_.p);
        launches_ += 2;
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
        if (T) count_row_candidates<<<gT, kBlock, 0, stream_>>>(d_tri_.p, T, d_idx_.p, cnt.p);
        long long ncand = 0;
        exclusive_scan(cnt.p, n, cptr.p, &ncand);
        REQUIRE(ncand < (long long)INT32_MAX, FEAB_EINVAL, "symbolic: candidate list exceeds int32");
        cand.alloc(std::max<long long>(ncand, 1));
        if (T) fill_row_candidates<<<gT, kBlock, 0, stream_>>>(d_tri_.p, T, d_idx_.p, cptr.p, fill.p, cand.p);
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
        d_tpos_.alloc(std::max<int64_t>(nnz_, 1));
        d_advp_.alloc(std::max<int64_t>(nnz_, 1));
        build_transpose_map<<<gN, kBlock, 0, stream_>>>(d_rowptr_.p, d_col_.p, n, d_node_of_.p, d_onb_.p, d_tpos_.p, d_advp_.p);
        launches_ += 5;
        CK(cudaMemcpyAsync(&max_row_, maxlen.p, sizeof(int), cudaMemcpyDeviceToHost, stream_));
        // --- scatter map (scatter assembly) ---
        d_smap_.alloc(std::max<int64_t>(9 * T, 1));
        if (T) build_scatter_map<<<gT, kBlock, 0, stream_>>>(d_tri_.p, T, d_idx_.p, d_rowptr_.p, d_col_.p, d_smap_.p);
        // --- ordered corner adjacency (ordered assembly) ---
        d_cptr_.alloc(n + 1);
        CK(cudaMemsetAsync(cnt.p, 0, (n + 1) * sizeof(int), stream_));
        CK(cudaMemsetAsync(fill.p, 0, (n + 1) * sizeof(int), stream_));
        if (T) count_row_corners<<<gT, kBlock, 0, stream_>>>(d_tri_.p, T, d_idx_.p, cnt.p);
        long long ncorner = 0;
        exclusive_scan(cnt.p, n, d_cptr_.p, &ncorner);
        d_corner_.alloc(std::max<long long>(ncorner, 1));
        if (T) fill_row_corners<<<gT, kBlock, 0, stream_>>>(d_tri_.p, T, d_idx_.p, d_cptr_.p, fill.p, d_corner_.p);
        sort_row_corners<<<gN, kBlock, 0, stream_>>>(d_cptr_.p, d_corner_.p, n);
        launches_ += 4;
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(stream_));
        // --- numeric storage ---
        const int64_t nz = std::max<int64_t>(nnz_, 1), nn = std::max<int64_t>(n, 1);
        d_M_.alloc(nz);
        d_K_.alloc(nz);
        d_Adv_.alloc(nz);
        d_Md_.alloc(nz);
        d_A0_.alloc(nz);
        d_A1_.alloc(nz);
        d_dinv0_.alloc(nn);
        d_dinv1_.alloc(nn);
        d_y_.alloc(nn);
        for (int s = 0; s < 2; s++) {
            d_r_[s].alloc(nn);
            d_p_[s].alloc(nn);
        }
        for (int s = 0; s < 5; s++) d_w_[s].alloc(nn);
        num_tiles_ = std::max(1, cdiv64(n, kTileRows));
        g_tile_ = std::min(num_tiles_, sms_ * 8);
        g_vec_ = std::max(1, std::min(cdiv64(std::max<int64_t>(n / Vec<R>::N, 1), kBlock), sms_ * 8));
        g_part_ = std::max(std::max(g_tile_, g_vec_), sms_ * 8);
        d_part_.alloc((size_t)16 * g_part_);
        resize_state();
    }

    void resize_state() {
        const int64_t nn = std::max<int64_t>(n_, 1);
        if (d_u_.count != (size_t)nn || !d_u_.p) {
            d_u_.alloc(nn);
            d_v_.alloc(nn);
            drop_graphs();
        }
        CK(cudaMemsetAsync(d_u_.p, 0, nn * sizeof(R), stream_));
        CK(cudaMemsetAsync(d_v_.p, 0, nn * sizeof(R), stream_));
        CK(cudaStreamSynchronize(stream_));
    }

    void assemble(const float* vel) override {
        REQUIRE(have_mesh_, FEAB_ENOMESH, "assemble: no mesh");
        CK(cudaSetDevice(device_));
        if (vel) std::memcpy(vel_, vel, 3 * sizeof(float));
        op_valid_ = false;
        const int64_t nz = std::max<int64_t>(nnz_, 1);
        if (opt_.assembly == FEAB_ASSEMBLY_ORDERED) {
            if (n_ > 0) {
                assemble_ordered<R><<<cdiv64(n_, kBlock), kBlock, 0, stream_>>>(d_xyz_.p, d_tri_.p, n_, d_onb_.p, d_idx_.p, d_cptr_.p,
                                                                                  d_corner_.p, d_rowptr_.p, d_col_.p, (R)vel_[0], (R)vel_[1],
                                                                                  (R)vel_[2], d_M_.p, d_K_.p, d_Adv_.p);
                launches_++;
            }
        } else {
            CK(cudaMemsetAsync(d_M_.p, 0, nz * sizeof(R), stream_));
            CK(cudaMemsetAsync(d_K_.p, 0, nz * sizeof(R), stream_));
            CK(cudaMemsetAsync(d_Adv_.p, 0, nz * sizeof(R), stream_));
            if (T_ > 0) {
                assemble_scatter<R><<<cdiv64(T_, kBlock), kBlock, 0, stream_>>>(d_xyz_.p, d_tri_.p, T_, d_onb_.p, d_smap_.p, (R)vel_[0],
                                                                                  (R)vel_[1], (R)vel_[2], d_M_.p, d_K_.p, d_Adv_.p);
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
        REQUIRE(o->loop_mode == 0 || o->loop_mode == 1, FEAB_EINVAL, "set_solver_options: loop_mode");
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
    }
    void set_values(const float* v) override {
        REQUIRE(have_mesh_ && v, FEAB_ENOMESH, "set_values: no mesh");
        CK(cudaSetDevice(device_));
        CK(cudaMemcpyAsync(d_values_.p, v, V_ * sizeof(float), cudaMemcpyHostToDevice, stream_));
        CK(cudaStreamSynchronize(stream_));
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
    }
    int unstable() override {
        REQUIRE(have_mesh_, FEAB_ENOMESH, "unstable: no mesh");
        CK(cudaSetDevice(device_));
        int f = 0;
        CK(cudaMemcpy(&f, d_unstable_, sizeof(int), cudaMemcpyDeviceToHost));
        return f;
    }

    // ---------------------------------------------------------------- operator cache (K2)
    StepScalars scalars() const {
        const FeabParams& p = par_[eq_];
        // promote the reference's floats exactly
        return StepScalars{(double)p.time_step, (double)p.c[0], (double)p.c[1], (double)p.c[2], (double)p.c[3]};
    }
    bool use_bicgstab() const { return eq_ == FEAB_EQ_ADVECTION_DIFFUSION && opt_.solver == FEAB_SOLVER_BICGSTAB; }

    void ensure_operator() {
        const FeabParams& p = par_[eq_];
        const int transposed = (eq_ == FEAB_EQ_ADVECTION_DIFFUSION && !use_bicgstab()) ? 1 : 0;
        if (op_valid_ && op_eq_ == eq_ && op_transposed_ == transposed && std::memcmp(&op_par_, &p, sizeof(p)) == 0) return;
        REQUIRE(p.time_step != 0.0f, FEAB_EINVAL, "time_step must be non-zero");
        if (nnz_ > 0) {
            int g = std::min(cdiv64(nnz_, kBlock), sms_ * 16);
            build_operator<R><<<g, kBlock, 0, stream_>>>(eq_, transposed, nnz_, d_M_.p, d_K_.p, d_Adv_.p, d_tpos_.p, scalars(), d_Md_.p, d_A0_.p,
                                                         d_A1_.p);
            inv_diag<R><<<cdiv64(n_, kBlock), kBlock, 0, stream_>>>(d_A0_.p, d_diag_.p, n_, d_dinv0_.p);
            launches_ += 2;
            if (eq_ == FEAB_EQ_REACTION_DIFFUSION) {
                inv_diag<R><<<cdiv64(n_, kBlock), kBlock, 0, stream_>>>(d_A1_.p, d_diag_.p, n_, d_dinv1_.p);
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

    void launch_iteration(cudaStream_t s, int slot, const R* A, const R* dinv, R* x, cudaGraphConditionalHandle cond, int use_cond) {
        CgState<R>* st = d_st_ + slot;
        launch_spmv(s, A, d_p_[slot].p, d_y_.p, st);
        cg_update<R><<<g_vec_, kBlock, 0, s>>>(x, d_r_[slot].p, d_p_[slot].p, d_y_.p, dinv, n_, st, part_pap(), spmv_parts(), part_rr(), part_rz());
        cg_direction<R><<<g_vec_, kBlock, 0, s>>>(d_r_[slot].p, d_p_[slot].p, dinv, n_, st, part_rr(), part_rz(), g_vec_, cond, use_cond);
    }

    // PCG solve of slot `slot` into x.  capturing: build WHILE node into the graph being captured on stream_.
    void enqueue_pcg(int slot, const R* A, const R* dinv, R* x, bool capturing) {
        CgState<R>* st = d_st_ + slot;
        const R* rhs = part_rhs() + (size_t)(slot * 2) * g_tile_;
        CK(cudaMemsetAsync(x, 0, n_ * sizeof(R), stream_));
        if (capturing) {
            cudaStreamCaptureStatus status;
            cudaGraph_t graph;
            const cudaGraphNode_t* deps;
            size_t ndeps;
            CK(cudaStreamGetCaptureInfo(stream_, &status, nullptr, &graph, &deps, &ndeps));
            cudaGraphConditionalHandle cond;
            CK(cudaGraphConditionalHandleCreate(&cond, graph, 0, 0));
            cg_begin<R><<<1, kBlock, 0, stream_>>>(st, rhs, rhs + g_tile_, g_tile_, tol(), maxit(), cond, 1);
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
            cg_begin<R><<<1, kBlock, 0, stream_>>>(st, rhs, rhs + g_tile_, g_tile_, tol(), maxit(), none, 0);
            launches_ += 1;
            const int batch = 8;
            const int64_t mi = maxit();
            for (int64_t done_iters = 0; done_iters < mi + batch;) {
                for (int b = 0; b < batch; b++) launch_iteration(stream_, slot, A, dinv, x, none, 0);
                launches_ += 3 * batch;
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
        const int gV = cdiv64(V_, kBlock);
        const StepScalars sc = scalars();
        const size_t sm2 = 2 * kTileCap * sizeof(R);
        CK(cudaMemsetAsync(d_unstable_, 0, sizeof(int), stream_));
        R* target = eq_ == FEAB_EQ_REACTION_DIFFUSION ? d_v_.p : d_u_.p;
        gather_values<R><<<gV, kBlock, 0, stream_>>>(d_values_.p, d_idx_.p, V_, target);
        R* prt = part_rhs();
        switch (eq_) {
            case FEAB_EQ_HEAT:
                rhs_init<R, 0><<<g_tile_, kBlock, sm2, stream_>>>(d_rowptr_.p, d_col_.p, d_Md_.p, d_K_.p, d_u_.p, d_v_.p, d_dinv0_.p, d_dinv1_.p, n_,
                                                                  num_tiles_, sc, d_r_[0].p, d_p_[0].p, d_r_[1].p, d_p_[1].p, prt);
                break;
            case FEAB_EQ_WAVE:
                rhs_init<R, 1><<<g_tile_, kBlock, sm2, stream_>>>(d_rowptr_.p, d_col_.p, d_Md_.p, d_K_.p, d_u_.p, d_v_.p, d_dinv0_.p, d_dinv1_.p, n_,
                                                                  num_tiles_, sc, d_r_[0].p, d_p_[0].p, d_r_[1].p, d_p_[1].p, prt);
                break;
            case FEAB_EQ_ADVECTION_DIFFUSION:
                rhs_init<R, 2><<<g_tile_, kBlock, sm2, stream_>>>(d_rowptr_.p, d_col_.p, d_Md_.p, d_K_.p, d_u_.p, d_v_.p, d_dinv0_.p, d_dinv1_.p, n_,
                                                                  num_tiles_, sc, d_r_[0].p, d_p_[0].p, d_r_[1].p, d_p_[1].p, prt);
                break;
            default:
                rhs_init<R, 3><<<g_tile_, kBlock, sm2, stream_>>>(d_rowptr_.p, d_col_.p, d_Md_.p, d_K_.p, d_u_.p, d_v_.p, d_dinv0_.p, d_dinv1_.p, n_,
                                                                  num_tiles_, sc, d_r_[0].p, d_p_[0].p, d_r_[1].p, d_p_[1].p, prt);
                break;
        }
        if (!capturing) launches_ += 2;
        CK(cudaGetLastError());
        switch (eq_) {
            case FEAB_EQ_HEAT: enqueue_pcg(0, d_A0_.p, d_dinv0_.p, d_u_.p, capturing); break;
            case FEAB_EQ_WAVE: enqueue_pcg(0, d_A0_.p, d_dinv0_.p, d_v_.p, capturing); break;
            case FEAB_EQ_ADVECTION_DIFFUSION:
                if (use_bicgstab())
                    enqueue_bicgstab(d_A0_.p, d_dinv0_.p, d_u_.p, capturing);
                else
                    enqueue_pcg(0, d_A0_.p, d_dinv0_.p, d_u_.p, capturing);
                break;
            default:
                enqueue_pcg(0, d_A0_.p, d_dinv0_.p, d_u_.p, capturing);
                enqueue_pcg(1, d_A1_.p, d_dinv1_.p, d_v_.p, capturing);
                break;
        }
        finish_step<R><<<gV, kBlock, 0, stream_>>>(d_values_.p, d_idx_.p, V_, d_u_.p, d_v_.p, eq_ == FEAB_EQ_WAVE ? 1 : 0, (R)sc.dt,
                                                   eq_ == FEAB_EQ_REACTION_DIFFUSION ? 1 : 0, d_unstable_);
        if (!capturing) launches_ += 1;
        CK(cudaGetLastError());
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

    void step(float* values, int nsteps, int sync) override {
        REQUIRE(have_mesh_, FEAB_ENOMESH, "step: no mesh");
        REQUIRE(nsteps >= 0, FEAB_EINVAL, "step: nsteps < 0");
        CK(cudaSetDevice(device_));
        CK(cudaEventRecord(ev0_, stream_));
        if (values) CK(cudaMemcpyAsync(d_values_.p, values, V_ * sizeof(float), cudaMemcpyHostToDevice, stream_));
        if (n_ > 0 && nsteps > 0) {
            ensure_operator();
            const bool graph_mode = opt_.loop_mode == FEAB_LOOP_GRAPH_WHILE;
            if (graph_mode && (!graph_exec_ || graph_eq_ != eq_)) build_graph();
            for (int s = 0; s < nsteps; s++) {
                if (graph_mode) {
                    CK(cudaGraphLaunch(graph_exec_, stream_));
                    graph_steps_++;
                } else {
                    enqueue_step(false);
                }
            }
        }
        if (values && sync) CK(cudaMemcpyAsync(values, d_values_.p, V_ * sizeof(float), cudaMemcpyDeviceToHost, stream_));
        CK(cudaEventRecord(ev1_, stream_));
        CK(cudaStreamSynchronize(stream_));
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, ev0_, ev1_));
        last_ms_ = ms;
        last_steps_ = nsteps;
        if (n_ > 0 && nsteps > 0 && opt_.loop_mode == FEAB_LOOP_GRAPH_WHILE) account_graph_launches(nsteps);
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
        int64_t per_step = 3;  // gather, rhs, finish
        if (use_bicgstab()) {
            per_step += 1 + (int64_t)kBicgKernelsPerIter * b.trips;
        } else {
            per_step += 1 + 3 * (int64_t)st[0].iter1;
            if (eq_ == FEAB_EQ_REACTION_DIFFUSION) per_step += 1 + 3 * (int64_t)st[1].iter1;
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
            s->rel_residual[0] = b.rhs_norm > 0 ? (double)(b.r_norm / b.rhs_norm) : 0.0;
        } else {
            for (int i = 0; i < 2; i++) {
                s->iterations[i] = st[i].iters;
                s->rel_residual[i] = st[i].rhs2 > 0 ? std::sqrt((double)st[i].rn2 / (double)st[i].rhs2) : 0.0;
            }
            if (eq_ != FEAB_EQ_REACTION_DIFFUSION) {
                s->iterations[1] = 0;
                s->rel_residual[1] = 0;
            }
        }
        s->last_call_ms = last_ms_;
        s->last_call_steps = last_steps_;
        s->kernel_launches = launches_;
    }

    // ---------------------------------------------------------------- measurement hook
    void time_kernel(int kernel, int reps, double* ms_out, double* bytes_out) override {
        REQUIRE(have_mesh_ && n_ > 0, FEAB_ENOMESH, "time_kernel: no mesh");
        REQUIRE(reps > 0 && kernel >= 0 && kernel <= 3, FEAB_EINVAL, "time_kernel: bad arguments");
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
                                                                 part_rr(), part_rz());
                    break;
                case 2:
                    cg_direction<R><<<g_vec_, kBlock, 0, stream_>>>(d_r_[0].p, d_p_[0].p, d_dinv0_.p, n_, d_st_, part_rr(), part_rz(), g_vec_, none, 0);
                    break;
                default:
                    if (eq_ == FEAB_EQ_REACTION_DIFFUSION)
                        rhs_init<R, 3><<<g_tile_, kBlock, 2 * kTileCap * sizeof(R), stream_>>>(d_rowptr_.p, d_col_.p, d_Md_.p, d_K_.p, d_u_.p, d_v_.p,
                                                                                               d_dinv0_.p, d_dinv1_.p, n_, num_tiles_, sc, d_r_[0].p,
                                                                                               d_p_[0].p, d_r_[1].p, d_p_[1].p, part_rhs());
                    else if (eq_ == FEAB_EQ_WAVE)
                        rhs_init<R, 1><<<g_tile_, kBlock, 2 * kTileCap * sizeof(R), stream_>>>(d_rowptr_.p, d_col_.p, d_Md_.p, d_K_.p, d_u_.p, d_v_.p,
                                                                                               d_dinv0_.p, d_dinv1_.p, n_, num_tiles_, sc, d_r_[0].p,
                                                                                               d_p_[0].p, d_r_[1].p, d_p_[1].p, part_rhs());
                    else
                        rhs_init<R, 0><<<g_tile_, kBlock, 2 * kTileCap * sizeof(R), stream_>>>(d_rowptr_.p, d_col_.p, d_Md_.p, d_K_.p, d_u_.p, d_v_.p,
                                                                                               d_dinv0_.p, d_dinv1_.p, n_, num_tiles_, sc, d_r_[0].p,
                                                                                               d_p_[0].p, d_r_[1].p, d_p_[1].p, part_rhs());
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
    DevBuf<int> d_idx_, d_node_of_, d_rowptr_, d_col_, d_diag_, d_tpos_, d_smap_, d_cptr_, d_corner_;
    DevBuf<long long> d_scan_bsum_, d_scan_total_, d_brush_;
    DevBuf<R> d_M_, d_K_, d_Adv_, d_Md_, d_A0_, d_A1_, d_dinv0_, d_dinv1_;
    DevBuf<R> d_u_, d_v_, d_y_, d_r_[2], d_p_[2], d_w_[5], d_part_;
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

    double last_ms_ = 0;
    int64_t last_steps_ = 0;
    int64_t launches_ = 0;
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
