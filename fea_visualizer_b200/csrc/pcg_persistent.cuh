// K3+K4+K5 as ONE persistent kernel: the whole Jacobi-PCG solve (ConjugateGradient.h:30-85) runs inside a single
// cooperative launch.  A grid of MINB CTAs per SM stays resident and walks the three phases of every iteration
//   A  y = A p (TMA-fed tile pipeline of spmv_tma.cuh) + partial of p.y
//   B  alpha; r -= alpha y; partials of |r|^2 and r.(D^-1 r)
//   C  convergence test (every CTA takes the same decision from the same fixed-order sums), beta,
//      x += alpha p (deferred from B: C reads p anyway), p = D^-1 r + beta p
// separated by grid barriers instead of kernel boundaries.  What this buys:
//   * one GPU: no launch gaps, no graph WHILE node, and the TMA prologue of the next SpMV is issued before phases B/C
//     so the matrix stream never drains;
//   * partitioned solve: per iteration 2 mailbox all-reduces + 1 halo push ride on the barriers (CTA 0 publishes /
//     pushes, every CTA consumes), instead of 4 kernel launches whose heads and tails wait for the peers.
// The arithmetic (operation order inside a row, per-CTA partials re-summed in CTA order) is the same as in the
// three-kernel version.  Vectors that change during the kernel (p, r, x, y) are never read through the non-coherent
// path; the barrier follows the fence / atomic / acquire pattern of cooperative groups, which also drops stale L1
// lines.
#pragma once
#include "comm.cuh"
#include "common.cuh"
#include "solver.cuh"
#include "spmv_tma.cuh"

namespace feab {

__device__ __forceinline__ unsigned long long ld_acquire_gpu_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// Monotonic 64-bit arrival counter, zeroed by the host before every launch.  `target` is this thread's running count.
__device__ __forceinline__ void grid_barrier(unsigned long long* counter, unsigned long long& target) {
    __syncthreads();
    target += gridDim.x;
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(counter, 1ull);
        while (ld_acquire_gpu_u64(counter) < target) {
        }
        __threadfence();
    }
    __syncthreads();
}

template <typename R>
struct PcgArgs {
    const int* rowptr;
    const int* col;
    const R* A;
    const R* dinv;
    R* x;
    R* r;
    R* p;
    R* y;
    int64_t n;
    int num_tiles;
    int first_halo_tile;
    CgState<R>* st;          // prepared by cg_begin (threshold, r.z of the first direction, early-outs)
    R* part;                 // 3 x gridDim.x: p.Ap | |r|^2 | r.z
    unsigned long long* bar; // grid barrier counter (zero at launch)
    int vec_p;               // which arena vector p is (kVecP0 / kVecP1): where the neighbours keep its ghosts
    unsigned long long* prof; // optional (FEAB_PCG_PROF=1): ns spent by CTA 0 in [A, bar1+sum, ar1, B, bar2+sum, ar2, C, bar3, push], iterations
};

// Sum of gridDim.x per-CTA partials written before the last grid barrier, in CTA order; identical in every CTA.
template <typename R, int NW>
__device__ __forceinline__ R sum_cta_partials(const R* part, R* red) {
    R s = (R)0;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x) s += __ldcg(part + i);
    s = block_sum_n<R, NW>(s, red);
    __syncthreads();
    return s;
}

// K such sums at once (partials of sum k at part + k * stride): one pass over L2, one block reduction; same bits.
template <typename R, int K, int NW>
__device__ __forceinline__ void sum_cta_partials_vec(const R* part, size_t stride, R (&out)[K], R* red /* (NW + 1) * K */) {
#pragma unroll
    for (int k = 0; k < K; k++) out[k] = (R)0;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x) {
#pragma unroll
        for (int k = 0; k < K; k++) out[k] += __ldcg(part + k * stride + i);
    }
    block_sum_vec<R, K, NW>(out, red);
}

// Mailbox all-reduce riding on a grid barrier: CTA 0 stores the rank totals into every peer's mailbox (one thread per
// peer, LL words: no fence), every CTA waits for the peers' totals in its own mailbox (one thread per peer) and adds
// them in rank order -- bit-identical on all CTAs of all ranks.
template <typename R, int K>
__device__ __forceinline__ void ar_exchange(const Comm& cm, unsigned long long& seq, R (&v)[K], double* s_pk /* kMaxRanks * 4 */) {
    seq += 1;
    const int q = threadIdx.x;
    if (q < cm.world) {
        if (q == cm.rank) {
            if (blockIdx.x == 0) cm.loc->ar_seq = seq;
#pragma unroll
            for (int k = 0; k < K; k++) s_pk[q * 4 + k] = (double)v[k];
        } else {
            if (blockIdx.x == 0) {
                ArSlot* sl = &cm.ctrl[q]->slot[seq & 1][cm.rank];
#pragma unroll
                for (int k = 0; k < K; k++) ll_store(sl, k, (double)v[k], seq);
            }
            const ArSlot* mine = &cm.ctrl[cm.rank]->slot[seq & 1][q];
#pragma unroll
            for (int k = 0; k < K; k++) s_pk[q * 4 + k] = ll_wait(mine, k, seq, cm.loc);
        }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < K; k++) {
        R acc = (R)0;
        for (int w = 0; w < cm.world; w++) acc += (R)s_pk[w * 4 + k];
        v[k] = acc;
    }
    __syncthreads();
}

template <typename R, int ROWS, int CPR, int STAGES, int MINB>
__global__ void __launch_bounds__(ROWS, MINB) pcg_persistent(PcgArgs<R> a, Comm cm, HaloPlan hp) {
    using Cfg = TmaCfg<R, ROWS, CPR, STAGES>;
    using O = Ops<R>;
    constexpr int NW = ROWS / 32;
    extern __shared__ unsigned char smraw_pcg[];
    unsigned char* base = reinterpret_cast<unsigned char*>(((uintptr_t)smraw_pcg + 127) & ~(uintptr_t)127);
    uint64_t* full = reinterpret_cast<uint64_t*>(base + Cfg::kStageBytes * STAGES);
    __shared__ R red[(NW + 1) * 2];
    __shared__ double s_v[kMaxRanks * 4];

    CgState<R>* st = a.st;
    if (st->done) return;  // rhs == 0 or already below the threshold (cg_begin); uniform over the grid
    const R thr = st->thr;
    const int maxit = st->maxit;
    R abs_old = st->abs_new[0];
    const int64_t n = a.n;
    const int num_tiles = a.num_tiles;
    R* const px = a.p;  // the SpMV gathers from p: plain (coherent) loads only

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
        const int tb = a.rowptr[r0], te = a.rowptr[r0 + rows];
        const int tba = tb & ~3;
        const int cnt_al = ((te - tba) + 3) & ~3;
        const uint32_t rp_b = (uint32_t)(((rows + 1 + 3) & ~3) * 4);
        const bool staged = (te - tba) <= Cfg::kCap;
        const uint32_t bytes = rp_b + (staged ? (uint32_t)cnt_al * (uint32_t)(sizeof(R) + 4) : 0u);
        mbar_expect_tx(&full[s], bytes);
        tma_load_1d(stage_rp(s), a.rowptr + r0, rp_b, &full[s], policy);
        if (staged) {
            tma_load_1d(stage_val(s), a.A + tba, (uint32_t)cnt_al * (uint32_t)sizeof(R), &full[s], policy);
            tma_load_1d(stage_col(s), a.col + tba, (uint32_t)cnt_al * 4u, &full[s], policy);
        }
    };
    // prologue of an SpMV phase: the first STAGES tiles of this CTA.  The matrix never changes, so it is issued
    // ahead of phases B / C of the previous iteration and lands while they run.
    unsigned fills = 0;  // tiles consumed so far by this CTA over the whole solve (stage / parity bookkeeping)
    auto prologue = [&]() {
        if (threadIdx.x == 0) {
            int t = blockIdx.x;
            for (int s = 0; s < STAGES && t < num_tiles; s++, t += gridDim.x) issue(t, (int)((fills + s) % STAGES));
        }
    };
    prologue();

    unsigned long long bar_target = 0;
    unsigned long long ar_seq = 0, halo_seq = 0;
    if (cm.world > 1) {
        ar_seq = *(volatile unsigned long long*)&cm.loc->ar_seq;
        halo_seq = *(volatile unsigned long long*)&cm.loc->halo_seq;
    }
    R* part_pap = a.part;
    R* part_rr = a.part + gridDim.x;
    R* part_rz = a.part + 2 * (size_t)gridDim.x;
    constexpr int N = Vec<R>::N;
    const int64_t nvec = n / N;
    const int64_t gthreads = (int64_t)gridDim.x * blockDim.x;
    const int64_t gtid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;

    // partitioned solve: mirror a boundary row's new p on every neighbour that keeps it as a ghost
    auto push_row = [&](int64_t row, R val) {
        if (row < hp.n_int) return;
        const int* t = hp.push_tab + (size_t)(row - hp.n_int) * hp.nnb;
        for (int b = 0; b < hp.nnb; b++) {
            const int d = t[b];
            if (d >= 0) reinterpret_cast<R*>(hp.base[hp.nb_rank[b]] + hp.off_vec[a.vec_p][hp.nb_rank[b]])[d] = val;
        }
    };
    int it = st->iter1, iters = 0;  // 0 after cg_begin; pcg_dual.cuh hands a running solve over with its count
    const int it0 = it;
    R rr = (R)0;
    const bool prof = a.prof != nullptr && blockIdx.x == 0 && threadIdx.x == 0;
    unsigned long long t_last = prof ? globaltimer_ns() : 0ull;
#define PCG_STAMP(slot)                               \
    if (prof) {                                       \
        const unsigned long long t_now = globaltimer_ns(); \
        a.prof[slot] += t_now - t_last;               \
        t_last = t_now;                               \
    }
    while (true) {
        // ---------------- phase A: y = A p, partial of p.y  (SparseDenseProduct.h:66-88, ConjugateGradient.h:66) ----
        R acc = (R)0;
        bool halo_ok = cm.world == 1;
        int k = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, k++) {
            if (!halo_ok && tile >= a.first_halo_tile) {
                if (threadIdx.x == 0) wait_flags(cm, hp, 0, halo_seq);
                __syncthreads();
                halo_ok = true;
            }
            const unsigned f = fills + (unsigned)k;
            const int s = (int)(f % STAGES);
            mbar_wait(&full[s], (f / STAGES) & 1u);
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
                    for (int j = 0; j < 8; j++) xv[j] = px[cv[j]];
#pragma unroll
                    for (int j = 0; j < 8; j++) {
                        if (j < len) {
                            if (j & 1)
                                tb2 += av[j] * xv[j];
                            else
                                ta += av[j] * xv[j];
                        }
                    }
                    xr = xv[7];
                    if (len >= 8) {
                        xr = px[r];
                        for (int j = 8; j < len; j++) {
                            R pr = sv[j] * px[sc[j]];
                            if (j & 1)
                                tb2 += pr;
                            else
                                ta += pr;
                        }
                    }
                } else {
                    int q = kb;
                    for (; q + 1 < ke; q += 2) {
                        ta += a.A[q] * px[a.col[q]];
                        tb2 += a.A[q + 1] * px[a.col[q + 1]];
                    }
                    if (q < ke) ta += a.A[q] * px[a.col[q]];
                    xr = px[r];
                }
                const R sum = ta + tb2;
                a.y[r] = sum;
                acc += xr * sum;
            }
            __syncthreads();  // every consumer is done with stage s
            if (threadIdx.x == 0) {
                const int next = tile + STAGES * gridDim.x;
                if (next < num_tiles) issue(next, s);
            }
        }
        fills += (unsigned)k;
        it++;
        acc = block_sum_n<R, NW>(acc, red);
        if (threadIdx.x == 0) part_pap[blockIdx.x] = acc;
        __syncthreads();
        prologue();  // next iteration's first tiles stream in while phases B / C run
        PCG_STAMP(0)
        grid_barrier(a.bar, bar_target);
        R pap[1] = {sum_cta_partials<R, NW>(part_pap, red)};
        PCG_STAMP(1)
        if (cm.world > 1) ar_exchange<R, 1>(cm, ar_seq, pap, s_v);
        PCG_STAMP(2)

        // ---------------- phase B: r -= alpha y, |r|^2, r.z  (ConjugateGradient.h:67-76) --------------------------
        // x += alpha p is deferred to phase C, which reads p anyway (one vector read less per iteration; same
        // operations on the same operands, so the same bits).
        const R alpha = abs_old / pap[0];
        R prr = (R)0, prz = (R)0;
        for (int64_t i = gtid; i < nvec; i += gthreads) {
            R rv[N], yv[N], dv[N];
            vload<R>(a.r + i * N, rv);
            vload<R>(a.y + i * N, yv);
            vload<R>(a.dinv + i * N, dv);
#pragma unroll
            for (int q = 0; q < N; q++) {
                rv[q] = O::sub(rv[q], O::mul(alpha, yv[q]));
                R z = O::mul(dv[q], rv[q]);
                prr += rv[q] * rv[q];
                prz += rv[q] * z;
            }
            vstore<R>(a.r + i * N, rv);
        }
        if (blockIdx.x == 0) {
            for (int64_t i = nvec * N + threadIdx.x; i < n; i += blockDim.x) {
                R rn = O::sub(a.r[i], O::mul(alpha, a.y[i]));
                a.r[i] = rn;
                R z = O::mul(a.dinv[i], rn);
                prr += rn * rn;
                prz += rn * z;
            }
        }
        {
            R pr2[2] = {prr, prz};
            block_sum_vec<R, 2, NW>(pr2, red);
            if (threadIdx.x == 0) {
                part_rr[blockIdx.x] = pr2[0];
                part_rz[blockIdx.x] = pr2[1];
            }
        }
        PCG_STAMP(3)
        grid_barrier(a.bar, bar_target);
        R t2[2];
        sum_cta_partials_vec<R, 2, NW>(part_rr, (size_t)gridDim.x, t2, red);  // part_rz = part_rr + gridDim.x
        PCG_STAMP(4)
        if (cm.world > 1) ar_exchange<R, 2>(cm, ar_seq, t2, s_v);
        PCG_STAMP(5)
        rr = t2[0];
        const R rz = t2[1];

        // ---------------- convergence test: same sums, same decision in every CTA (ConjugateGradient.h:72-73) -----
        const bool conv = rr < thr;
        if (conv || it >= maxit || !(rr <= O::huge())) {  // also stop on NaN / Inf instead of spinning to 2n
            iters = conv ? it - 1 : it;
            if (!conv) abs_old = rz;
            // the deferred x += alpha p of this last iteration
            for (int64_t i = gtid; i < nvec; i += gthreads) {
                R xv[N], pv[N];
                vload<R>(a.x + i * N, xv);
                vload<R>(a.p + i * N, pv);
#pragma unroll
                for (int q = 0; q < N; q++) xv[q] = O::add(xv[q], O::mul(alpha, pv[q]));
                vstore<R>(a.x + i * N, xv);
            }
            if (blockIdx.x == 0)
                for (int64_t i = nvec * N + threadIdx.x; i < n; i += blockDim.x) a.x[i] = O::add(a.x[i], O::mul(alpha, a.p[i]));
            break;
        }
        // ---------------- phase C: x += alpha p; p = D^-1 r + beta p  (ConjugateGradient.h:68, :77-80) -----------
        const R beta = rz / abs_old;
        abs_old = rz;
        bool pushed = false;
        for (int64_t i = gtid; i < nvec; i += gthreads) {
            R xv[N], rv[N], pv[N], dv[N];
            vload<R>(a.x + i * N, xv);
            vload<R>(a.r + i * N, rv);
            vload<R>(a.p + i * N, pv);
            vload<R>(a.dinv + i * N, dv);
#pragma unroll
            for (int q = 0; q < N; q++) {
                xv[q] = O::add(xv[q], O::mul(alpha, pv[q]));
                pv[q] = O::add(O::mul(dv[q], rv[q]), O::mul(beta, pv[q]));
            }
            vstore<R>(a.x + i * N, xv);
            vstore<R>(a.p + i * N, pv);
            if (cm.world > 1 && i * N + N > hp.n_int) {  // boundary rows: straight into the neighbours' ghost tails
#pragma unroll
                for (int q = 0; q < N; q++) push_row(i * N + q, pv[q]);
                pushed = true;
            }
        }
        if (blockIdx.x == 0)
            for (int64_t i = nvec * N + threadIdx.x; i < n; i += blockDim.x) {
                const R pi = a.p[i];
                a.x[i] = O::add(a.x[i], O::mul(alpha, pi));
                const R pn = O::add(O::mul(a.dinv[i], a.r[i]), O::mul(beta, pi));
                a.p[i] = pn;
                if (cm.world > 1) {
                    push_row(i, pn);
                    pushed = true;
                }
            }
        if (cm.world > 1 && __syncthreads_or(pushed ? 1 : 0)) __threadfence_system();  // remote stores done before the barrier
        PCG_STAMP(6)
        grid_barrier(a.bar, bar_target);
        PCG_STAMP(7)
        if (cm.world > 1) {
            // every boundary row of the new direction already sits in its neighbours' ghost tails: the CTAs that
            // pushed ran a system-scope fence before arriving at the barrier, so the data is visible system-wide
            // and a relaxed flag store is enough (a release here would cost another system fence on the critical path).  Every CTA counts the push so that its next halo wait
            // asks for the right flag value.
            halo_seq += 1;
            if (blockIdx.x == 0) {  // one thread per neighbour: the flag stores leave in parallel
                if (threadIdx.x == 0) cm.loc->halo_seq = halo_seq;
                if ((int)threadIdx.x < hp.nnb) st_relaxed_sys(&cm.ctrl[hp.nb_rank[threadIdx.x]]->halo_flag[cm.rank], halo_seq);
            }
        }
        PCG_STAMP(8)
    }
#undef PCG_STAMP
    if (prof) a.prof[9] += (unsigned long long)(it - it0);
    // drain the prologue issued for an iteration that never runs, so no bulk copy is in flight at exit
    {
        int k = 0;
        for (int tile = blockIdx.x; tile < num_tiles && k < STAGES; tile += gridDim.x, k++) {
            const unsigned f = fills + (unsigned)k;
            mbar_wait(&full[f % STAGES], (f / STAGES) & 1u);
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        st->rn2 = rr;
        st->abs_new[it & 1] = abs_old;
        st->iter1 = it;
        st->iters = iters;
        st->done = 1;
    }
}

}  // namespace feab
