// K6 (TMA variant): RHS build + PCG start for the equations whose RHS is one product with M/dt -- heat, advection-
// diffusion (CPUSolver.cpp:78, :96) and Gray-Scott (both species + the reaction terms in one pass, :134, :137) -- fused
// with r = b, p = D^-1 b and the partials of |b|^2, b.p (ConjugateGradient.h:43-60).  Same arithmetic as rhs_init
// (solver.cuh): per row ONE accumulator in ascending column order (Eigen's column-major sparse * dense), every product
// rounded before it is added; only the data movement differs: the matrix streams through the 2-stage bulk-TMA ring of
// spmv_tma.cuh instead of LDG + CTA barriers, which is what lifted the SpMV from 4.0 to 6.8 TB/s.
// The wave equation (two value arrays per entry) keeps rhs_init.
#pragma once
#include "spmv_tma.cuh"

namespace feab {

template <typename R, int EQ, int ROWS, int CPR, int STAGES, int MINB>
__global__ void __launch_bounds__(ROWS, MINB)
    rhs_init_tma(const int* __restrict__ rowptr, const int* __restrict__ col, const R* __restrict__ Md, const R* __restrict__ u,
                 const R* __restrict__ v, const R* __restrict__ dinv0, const R* __restrict__ dinv1, int64_t n, int num_tiles,
                 StepScalars sc, R* __restrict__ r0o, R* __restrict__ p0o, R* __restrict__ r1o, R* __restrict__ p1o,
                 R* __restrict__ part /* 4 x pstride: bb0, rz0, bb1, rz1 */, int pstride, Comm cm, HaloPlan hp, int publish) {
    static_assert(EQ == 0 || EQ == 2 || EQ == 3, "wave keeps rhs_init");
    using O = Ops<R>;
    using Cfg = TmaCfg<R, ROWS, CPR, STAGES>;
    constexpr int NW = ROWS / 32;
    extern __shared__ unsigned char smraw_tma[];
    unsigned char* base = reinterpret_cast<unsigned char*>(((uintptr_t)smraw_tma + 127) & ~(uintptr_t)127);
    uint64_t* full = reinterpret_cast<uint64_t*>(base + Cfg::kStageBytes * STAGES);
    __shared__ R red[NW + 1];
    auto stage_val = [&](int s) { return reinterpret_cast<R*>(base + (size_t)s * Cfg::kStageBytes); };
    auto stage_col = [&](int s) { return reinterpret_cast<int*>(base + (size_t)s * Cfg::kStageBytes + Cfg::kValBytes); };
    auto stage_rp = [&](int s) { return reinterpret_cast<int*>(base + (size_t)s * Cfg::kStageBytes + Cfg::kValBytes + Cfg::kColBytes); };

    if (comm_failed(cm)) return;  // fail closed (comm.cuh), one GPU included: a timed-out grid sync sets the same flag
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
        mbar_expect_tx(&full[s], rp_b + (staged ? (uint32_t)cnt_al * (uint32_t)(sizeof(R) + 4) : 0u));
        tma_load_1d(stage_rp(s), rowptr + r0, rp_b, &full[s], policy);
        if (staged) {
            tma_load_1d(stage_val(s), Md + tba, (uint32_t)cnt_al * (uint32_t)sizeof(R), &full[s], policy);
            tma_load_1d(stage_col(s), col + tba, (uint32_t)cnt_al * 4u, &full[s], policy);
        }
    };
    if (threadIdx.x == 0) {
        int t = blockIdx.x;
        for (int s = 0; s < STAGES && t < num_tiles; s++, t += gridDim.x) issue(t, s);
    }
    // the matrix stream is on its way; the ghost entries of u / v come from the neighbours' end-of-step pushes
    bool pushed = false;
    if (cm.world > 1) {
        if (threadIdx.x == 0) wait_flags(cm, hp, 1, *(volatile unsigned long long*)&cm.loc->state_seq);
        __syncthreads();
        if (comm_failed(cm)) {
            // the neighbours' u / v never arrived: build nothing, push nothing (comm.cuh contract).  The bulk copies already
            // issued still target this CTA's shared memory: wait them out before the CTA may exit.
            if (threadIdx.x == 0) {
                int t = blockIdx.x;
                for (int s = 0; s < STAGES && t < num_tiles; s++, t += gridDim.x) mbar_wait(&full[s], 0u);
            }
            return;
        }
    }
    // the vector surface->values feeds (u; Gray-Scott: v) is read through float (solver.cuh through_float)
    auto ldu = [&](int64_t c) { return EQ == 3 ? u[c] : through_float(u[c]); };
    auto ldv = [&](int64_t c) { return through_float(v[c]); };

    R bb0 = (R)0, rz0 = (R)0, bb1 = (R)0, rz1 = (R)0;
    int i = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, i++) {
        const int s = i % STAGES;
        mbar_wait(&full[s], (uint32_t)((i / STAGES) & 1));
        const int* rp = stage_rp(s);
        const int64_t r0 = (int64_t)tile * ROWS;
        const int rows = (int)min((int64_t)ROWS, n - r0);
        const int tb = rp[0], te = rp[rows];
        const int tba = tb & ~3;
        const bool staged = (te - tba) <= Cfg::kCap;
        if ((int)threadIdx.x < rows) {
            const int64_t r = r0 + threadIdx.x;
            const int kb = rp[threadIdx.x], ke = rp[threadIdx.x + 1];
            R s0 = (R)0, s1 = (R)0, ui, vi = (R)0;
            if (staged) {
                const R* sv = stage_val(s) + (kb - tba);
                const int* scl = stage_col(s) + (kb - tba);
                const int len = ke - kb;
                R av[8], uv[8], vv[8];
                int cv[8];
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    const bool in = j < len;
                    cv[j] = in ? scl[j] : (int)r;   // padding slots gather the row's own entry (slot 7 doubles as u[r] / v[r])
                    av[j] = in ? sv[j] : (R)0;
                }
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    uv[j] = ldu(cv[j]);
                    if (EQ == 3) vv[j] = ldv(cv[j]);
                }
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    if (j < len) {
                        s0 = O::add(s0, O::mul(av[j], uv[j]));
                        if (EQ == 3) s1 = O::add(s1, O::mul(av[j], vv[j]));
                    }
                }
                ui = uv[7];
                if (EQ == 3) vi = vv[7];
                if (len >= 8) {
                    ui = ldu(r);
                    if (EQ == 3) vi = ldv(r);
                    for (int j = 8; j < len; j++) {
                        const int c = scl[j];
                        s0 = O::add(s0, O::mul(sv[j], ldu(c)));
                        if (EQ == 3) s1 = O::add(s1, O::mul(sv[j], ldv(c)));
                    }
                }
            } else {
                for (int k = kb; k < ke; k++) {
                    const int c = col[k];
                    const R m = Md[k];
                    s0 = O::add(s0, O::mul(m, ldu(c)));
                    if (EQ == 3) s1 = O::add(s1, O::mul(m, ldv(c)));
                }
                ui = ldu(r);
                if (EQ == 3) vi = ldv(r);
            }
            if (EQ == 3) {
                const R f = (R)sc.c2, fk = O::add((R)sc.c2, (R)sc.c3);
                const R uvv = O::mul(ui, O::mul(vi, vi));
                const R bu = O::add(O::sub(s0, uvv), O::mul(f, O::sub((R)1, ui)));
                const R bv = O::sub(O::add(s1, uvv), O::mul(fk, vi));
                const R pu = O::mul(dinv0[r], bu), pv = O::mul(dinv1[r], bv);
                r0o[r] = bu;
                p0o[r] = pu;
                r1o[r] = bv;
                p1o[r] = pv;
                bb0 += bu * bu;
                rz0 += bu * pu;
                bb1 += bv * bv;
                rz1 += bv * pv;
                if (cm.world > 1 && r >= hp.n_int) {  // first search directions of the boundary rows -> neighbours' ghost tails
                    push_boundary_row<R>(hp, kVecP0, r, pu);
                    push_boundary_row<R>(hp, kVecP1, r, pv);
                    pushed = true;
                }
            } else {
                const R pz = O::mul(dinv0[r], s0);
                r0o[r] = s0;
                p0o[r] = pz;
                bb0 += s0 * s0;
                rz0 += s0 * pz;
                if (cm.world > 1 && r >= hp.n_int) {
                    push_boundary_row<R>(hp, kVecP0, r, pz);
                    pushed = true;
                }
            }
        }
        __syncthreads();  // every consumer is done with stage s
        if (threadIdx.x == 0) {
            const int next = tile + STAGES * gridDim.x;
            if (next < num_tiles) issue(next, s);
        }
    }
    bb0 = block_sum_n<R, NW>(bb0, red);
    __syncthreads();
    rz0 = block_sum_n<R, NW>(rz0, red);
    __syncthreads();
    if (EQ == 3) {
        bb1 = block_sum_n<R, NW>(bb1, red);
        __syncthreads();
        rz1 = block_sum_n<R, NW>(rz1, red);
        __syncthreads();
    }
    constexpr int K = EQ == 3 ? 4 : 2;
    if (threadIdx.x == 0) {
        part[blockIdx.x] = bb0;
        part[(size_t)pstride + blockIdx.x] = rz0;
        if (EQ == 3) {
            part[(size_t)2 * pstride + blockIdx.x] = bb1;
            part[(size_t)3 * pstride + blockIdx.x] = rz1;
        }
    }
    // cg_begin sums `pstride` partials per quantity (the grid of rhs_init): the slots this grid does not own are zero
    if (blockIdx.x == 0)
        for (int k = 0; k < K; k++)
            for (int q = (int)gridDim.x + (int)threadIdx.x; q < pstride; q += ROWS) part[(size_t)k * pstride + q] = (R)0;
    if (publish) {  // totals -> every rank's mailbox (persistent solve kernels and partitioned handles start from there)
        if (cm.world > 1 && __syncthreads_or(pushed ? 1 : 0)) __threadfence_system();
        __syncthreads();
        ar_publish<R, K, NW>(cm, part, pstride, red, &hp, cm.world > 1);
    }
}

}  // namespace feab
