// K3 (TMA variant): y = A p fused with the partial of p.y, as a persistent, TMA-fed pipeline.
//
// One CTA per slot of a persistent grid walks tiles of ROWS consecutive CSR rows.  A tile's matrix data is
// three contiguous byte ranges (row pointers, values, column indices); thread 0 streams them into a ring of
// shared-memory stages with 1-D bulk TMA copies (cp.async.bulk ... mbarrier::complete_tx, SASS UBLKCP) that
// carry an L2 evict-first policy, STAGES tiles ahead of the consumers.  The consumers (all threads, one row
// each) wait on the stage's mbarrier, read their row's values / columns from shared memory, gather p from
// global memory (L1/L2), accumulate in Eigen's two-accumulator order and write y.  HBM stays busy while the
// consumers sit in gather latency, which the LDG-based version (solver.cuh: spmv_dot) cannot do because its
// loads, products and row sums are serialised by CTA barriers.
#pragma once
#include "common.cuh"
#include "solver.cuh"

namespace feab {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra DONE;\n"
        "bra LAB_WAIT;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void tma_load_1d(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar, uint64_t policy) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
                 : "memory");
}
__device__ __forceinline__ void l2_prefetch(const void* src_gmem, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src_gmem), "r"(bytes) : "memory");
}

// Geometry of one pipeline configuration.  CPR = staged entries per row of capacity (tiles with more entries
// fall back to direct global loads for that tile).
template <typename R, int ROWS, int CPR, int STAGES>
struct TmaCfg {
    static constexpr int kRows = ROWS;
    static constexpr int kCap = ROWS * CPR;
    static constexpr int kStages = STAGES;
    static constexpr size_t kValBytes = ((size_t)(kCap + 4) * sizeof(R) + 127) / 128 * 128;
    static constexpr size_t kColBytes = ((size_t)(kCap + 4) * 4 + 127) / 128 * 128;
    static constexpr size_t kRpBytes = ((size_t)(ROWS + 8) * 4 + 127) / 128 * 128;
    static constexpr size_t kStageBytes = kValBytes + kColBytes + kRpBytes;
    static constexpr size_t kSmemBytes = kStageBytes * STAGES + 128 /* barriers */ + 128 /* alignment slack */;
};

// Arrays must be over-allocated by kTmaPad elements: bulk copies round sizes up to 16 bytes.
constexpr int kTmaPad = 16;

template <typename R, int ROWS, int CPR, int STAGES, int MINB>
__global__ void __launch_bounds__(ROWS, MINB)
    spmv_dot_tma(const int* __restrict__ rowptr, const int* __restrict__ col, const R* __restrict__ val, const R* __restrict__ x,
                 R* __restrict__ y, int64_t n, int num_tiles, CgState<R>* __restrict__ st, R* __restrict__ part_pap, int prefetch_x,
                 Comm cm, HaloPlan hp, int first_halo_tile) {
    using Cfg = TmaCfg<R, ROWS, CPR, STAGES>;
    extern __shared__ unsigned char smraw_tma[];
    unsigned char* base = reinterpret_cast<unsigned char*>(((uintptr_t)smraw_tma + 127) & ~(uintptr_t)127);
    uint64_t* full = reinterpret_cast<uint64_t*>(base + Cfg::kStageBytes * STAGES);
    __shared__ R red[ROWS / 32 + 1];
    if (st != nullptr && st->done) return;

    auto stage_val = [&](int s) { return reinterpret_cast<R*>(base + (size_t)s * Cfg::kStageBytes); };
    auto stage_col = [&](int s) { return reinterpret_cast<int*>(base + (size_t)s * Cfg::kStageBytes + Cfg::kValBytes); };
    auto stage_rp = [&](int s) { return reinterpret_cast<int*>(base + (size_t)s * Cfg::kStageBytes + Cfg::kValBytes + Cfg::kColBytes); };

    uint64_t policy = 0;
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; s++) mbar_init(&full[s], 1);
        mbar_fence_init();
        policy = l2_evict_first_policy();
    }
    __syncthreads();

    auto issue = [&](int tile, int s) {  // thread 0 only
        const int64_t r0 = (int64_t)tile * ROWS;
        const int rows = (int)min((int64_t)ROWS, n - r0);
        const int tb = rowptr[r0], te = rowptr[r0 + rows];
        const int tba = tb & ~3;
        const int cnt_al = ((te - tba) + 3) & ~3;
        const uint32_t rp_b = (uint32_t)(((rows + 1 + 3) & ~3) * 4);
        const bool staged = (te - tba) <= Cfg::kCap;
        const uint32_t bytes = rp_b + (staged ? (uint32_t)cnt_al * (uint32_t)(sizeof(R) + 4) : 0u);
        mbar_expect_tx(&full[s], bytes);
        tma_load_1d(stage_rp(s), rowptr + r0, rp_b, &full[s], policy);
        if (staged) {
            tma_load_1d(stage_val(s), val + tba, (uint32_t)cnt_al * (uint32_t)sizeof(R), &full[s], policy);
            tma_load_1d(stage_col(s), col + tba, (uint32_t)cnt_al * 4u, &full[s], policy);
        }
        if (prefetch_x) l2_prefetch(x + r0, (uint32_t)(((rows * sizeof(R)) + 15) & ~15));
    };

    if (threadIdx.x == 0) {
        int t = blockIdx.x;
        for (int s = 0; s < STAGES && t < num_tiles; s++, t += gridDim.x) issue(t, s);
    }

    R acc = (R)0;
    int i = 0;
    bool halo_ok = cm.world == 1;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, i++) {
        if (!halo_ok && tile >= first_halo_tile) {
            // partitioned solve: rows from here on read ghost columns -- wait for the neighbours' pushes (comm.cuh)
            if (threadIdx.x == 0) wait_flags(cm, hp, 0, *(volatile unsigned long long*)&cm.loc->halo_seq);
            __syncthreads();
            halo_ok = true;
        }
        const int s = i % STAGES;
        const uint32_t parity = (uint32_t)((i / STAGES) & 1);
        mbar_wait(&full[s], parity);
        const int* rp = stage_rp(s);
        const int64_t r0 = (int64_t)tile * ROWS;
        const int rows = (int)min((int64_t)ROWS, n - r0);
        const int tb = rp[0], te = rp[rows];
        const int tba = tb & ~3;
        const bool staged = (te - tba) <= Cfg::kCap;
        if ((int)threadIdx.x < rows) {
            const int64_t r = r0 + threadIdx.x;
            const int kb = rp[threadIdx.x], ke = rp[threadIdx.x + 1];
            R ta = (R)0, tb2 = (R)0, xr = (R)0;
            if (staged) {
                const R* sv = stage_val(s) + (kb - tba);
                const int* sc = stage_col(s) + (kb - tba);
                const int len = ke - kb;
                R av[8], xv[8];
                int cv[8];
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    if (j < len) {
                        cv[j] = sc[j];
                        av[j] = sv[j];
                    } else {
                        cv[j] = (int)r;
                        av[j] = (R)0;
                    }
                }
#pragma unroll
                for (int j = 0; j < 8; j++) xv[j] = x[cv[j]];
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    if (j < len) {
                        if (j & 1)
                            tb2 += av[j] * xv[j];
                        else
                            ta += av[j] * xv[j];
                    }
                }
                xr = xv[7];  // slot 7 is a padding slot (col = r) for rows shorter than 8 ...
                if (len >= 8) {  // ... otherwise fetch it and finish the long row
                    xr = x[r];
                    for (int j = 8; j < len; j++) {
                        R p = sv[j] * x[sc[j]];
                        if (j & 1)
                            tb2 += p;
                        else
                            ta += p;
                    }
                }
            } else {
                int k = kb;
                for (; k + 1 < ke; k += 2) {
                    ta += val[k] * x[col[k]];
                    tb2 += val[k + 1] * x[col[k + 1]];
                }
                if (k < ke) ta += val[k] * x[col[k]];
                xr = x[r];
            }
            const R sum = ta + tb2;
            y[r] = sum;
            acc += xr * sum;
        }
        __syncthreads();  // every consumer is done with stage s
        if (threadIdx.x == 0) {
            const int next = tile + STAGES * gridDim.x;
            if (next < num_tiles) issue(next, s);
        }
    }
    acc = block_sum_n<R, ROWS / 32>(acc, red);
    if (threadIdx.x == 0) {
        part_pap[blockIdx.x] = acc;
        if (blockIdx.x == 0 && st != nullptr) st->iter1 += 1;
    }
    if (cm.world > 1) {
        __syncthreads();
        ar_publish<R, 1, ROWS / 32>(cm, part_pap, 0, red);
    }
}

}  // namespace feab
