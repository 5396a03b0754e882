// K3+K4+K5 as ONE persistent kernel: the whole Jacobi-PCG solve (ConjugateGradient.h:30-85) runs inside a single
// cooperative launch.  A grid of MINB CTAs per SM stays resident and walks the three phases of every iteration
//   A  y = A p (TMA-fed tile pipeline of spmv_tma.cuh) + partial of p.y
//   B  alpha; r -= alpha y; partials of |r|^2 and r.(D^-1 r)
//   C  convergence test (every CTA takes the same decision from the same fixed-order sums), beta,
//      x += alpha p (deferred from B: C reads p anyway), p = D^-1 r + beta p
// separated by grid-wide syncs instead of kernel boundaries.  What this buys:
//   * one GPU: no launch gaps, no graph WHILE node, and the TMA prologue of the next SpMV is issued before phases B/C
//     so the matrix stream never drains;
//   * partitioned solve: per iteration the 2 all-reduces ARE the two grid syncs (comm.cuh gsync_allreduce: the last CTA
//     to arrive publishes the rank total to every rank's mailbox, every CTA waits on the mailbox) and the halo flags of
//     the new direction are raised by the last CTA to arrive at the third.
// The start of the solve (threshold, early-outs: ConjugateGradient.h:43-60) is taken inside the kernel from the sums the
// RHS kernel published, and x = 0 is never written: the first iteration stores 0 + alpha p.  So a time step is
// RHS kernel -> solve kernel(s) -> epilogue kernel, nothing else.
// The arithmetic (operation order inside a row, per-CTA partials re-summed in CTA order) is the same as in the
// three-kernel version.  Vectors that change during the kernel (p, r, x, y) are never read through the non-coherent
// path; the syncs follow the fence / atomic / fence pattern of cooperative groups, which also drops stale L1 lines.
#pragma once
#include "comm.cuh"
#include "common.cuh"
#include "solver.cuh"
#include "spmv_tma.cuh"

namespace feab {

template <typename R>
struct PcgArgs {
    const int* rowptr;
    const int* col;
    const R* A[2];    // per system (Gray-Scott has two; the others use [0])
    const R* dinv[2];
    R* x[2];
    R* r[2];
    R* p[2];
    R* y;
    int64_t n;
    int num_tiles;
    int first_halo_tile;
    CgState<R>* st;    // [2]
    R* part;           // 3 x gridDim.x: p.Ap | |r|^2 | r.z
    int slot;          // >= 0: solve this system; -1: whichever of the two pcg_dual left running (the tail of a Gray-Scott step)
    int begin;         // > 0: this is the first solve kernel of the step: number of systems (1 or 2) whose start is taken here
                       // from the RHS kernel's published sums (mailbox all-reduce number ar_seq); 0: CgState is already set
    R tol;
    int maxit;
    int zigzag;        // order of phase C.  0: ascending.  2: ascending, but starting at the boundary rows (numbered last) and wrapping
                       // round, so that their new p is on its way to the neighbours over NVLink while the interior is updated.
                       // 1: DESCENDING (boundary rows first, and the rows phase B touched last) -- 13 % slower in that phase
                       // (208 vs 240 us at 16 M rows, profiles/r02_v17_*).  Default 0: also at N = 8 it beats 2 (342 vs 325 steps/s); FEAB_ZIGZAG
    unsigned long long* prof; // optional (FEAB_PCG_PROF=1): ns spent by CTA 0 in [A, sync1 (barrier + all-reduce of p.Ap), -, B, sync2, -, C, sync3, -], iterations
};

// Start of the step's solves, ConjugateGradient.h:43-60 with x0 = 0 (IterativeSolverBase.h:357-360): residual = b, so
// |b|^2 and b.D^-1 b are all that is needed; the RHS kernel's last CTA all-reduced them into the mailbox (all-reduce
// number *seq, [2 s] = |b_s|^2, [2 s + 1] = b_s.D^-1 b_s).  Every CTA computes the same state; thread 0 of CTA 0 records it.
template <typename R, int NS>
__device__ __forceinline__ bool pcg_begin(const Comm& cm, unsigned long long seq, R tol, int maxit, CgState<R>* st_global, CgState<R> (&out)[NS],
                                          double* s_pk, int* s_ctl) {
    if ((int)threadIdx.x < cm.world) {
        const ArSlot* mine = &cm.ctrl[cm.rank]->slot[seq & 1][threadIdx.x];
        bool ok = true;
#pragma unroll
        for (int k = 0; k < 2 * NS; k++) s_pk[threadIdx.x * 4 + k] = ll_wait(mine, k, seq, cm.loc, &ok, (int)threadIdx.x);
        if (!ok) s_ctl[1] = 1;
        __threadfence();
    }
    __syncthreads();
#pragma unroll
    for (int s = 0; s < NS; s++) {
        R bb = (R)0, rz = (R)0;
        for (int w = 0; w < cm.world; w++) {
            bb += (R)s_pk[w * 4 + 2 * s];
            rz += (R)s_pk[w * 4 + 2 * s + 1];
        }
        R thr = tol * tol * bb;
        if (thr < Ops<R>::tiny()) thr = Ops<R>::tiny();
        CgState<R>& c = out[s];
        c.rhs2 = bb;
        c.thr = thr;
        c.rn2 = bb == (R)0 ? (R)0 : bb;
        c.tol = tol;
        c.abs_new[0] = rz;
        c.abs_new[1] = (R)0;
        c.iter1 = 0;
        c.iters = 0;
        c.maxit = maxit;
        c.done = (bb == (R)0 || bb < thr || maxit <= 0) ? 1 : 0;
        if (blockIdx.x == 0 && threadIdx.x == 0) st_global[s] = c;
    }
    const bool failed = s_ctl[1] != 0;
    __syncthreads();
    return !failed;
}

// x = 0 for a system whose solve ended before its first iteration (rhs == 0: ConjugateGradient.h:45-49; residual already
// below the threshold: x keeps the zero guess)
template <typename R>
__device__ __forceinline__ void zero_solution(R* x, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) x[i] = (R)0;
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
    __shared__ double s_pk[kMaxRanks * 4];
    __shared__ int s_ctl[2];

    if (comm_failed(cm)) return;  // an earlier kernel of this call lost a peer: fail closed
    if (threadIdx.x == 0) s_ctl[1] = 0;
    unsigned long long ar_seq = *(volatile unsigned long long*)&cm.loc->ar_seq;
    unsigned long long halo_seq = *(volatile unsigned long long*)&cm.loc->halo_seq;
    unsigned long long bar_seq = ld_volatile_u64(&cm.loc->gs_release);
    unsigned long long gs_target = ld_volatile_u64(&cm.loc->gs_base);  // one GPU: arrival count that completes the next grid sync (comm.cuh)
    __syncthreads();

    // ---- which system, and its state ----
    int slot = a.slot;
    CgState<R> c;
    if (a.begin > 0) {
        CgState<R> b2[2];
        bool ok;
        if (a.begin == 2) {
            ok = pcg_begin<R, 2>(cm, ar_seq, a.tol, a.maxit, a.st, b2, s_pk, s_ctl);
        } else {
            CgState<R> b1[1];
            ok = pcg_begin<R, 1>(cm, ar_seq, a.tol, a.maxit, a.st, b1, s_pk, s_ctl);
            b2[0] = b1[0];
            b2[1] = b1[0];
        }
        if (!ok) return;
        for (int s = 0; s < a.begin; s++)
            if (b2[s].done) zero_solution<R>(a.x[s], a.n);
        c = b2[slot];
    } else {
        if (slot < 0) {
            const int d0 = a.st[0].done, d1 = a.st[1].done;
            if (d0 && d1) return;
            slot = d0 ? 1 : 0;
        }
        c = a.st[slot];
    }
    if (c.done) return;  // uniform over the grid
    CgState<R>* st = a.st + slot;
    const R* const Aval = a.A[slot];
    const R* const dinv = a.dinv[slot];
    R* const xs = a.x[slot];
    R* const rs = a.r[slot];
    R* const px = a.p[slot];  // the SpMV gathers from p: plain (coherent) loads only
    const int vec_p = slot == 0 ? kVecP0 : kVecP1;
    const R thr = c.thr;
    const int maxit = c.maxit;
    R abs_old = c.abs_new[0];
    const int64_t n = a.n;
    const int num_tiles = a.num_tiles;

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
            tma_load_1d(stage_val(s), Aval + tba, (uint32_t)cnt_al * (uint32_t)sizeof(R), &full[s], policy);
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

    R* part_pap = a.part;
    R* part_rr = a.part + gridDim.x;
    constexpr int N = Vec<R>::N;
    const int64_t nvec = n / N;
    const int64_t gthreads = (int64_t)gridDim.x * blockDim.x;
    const int64_t gtid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t c_start = (a.zigzag == 2 && cm.world > 1) ? (min(nvec, (int64_t)hp.n_int / N) & ~(int64_t)63) : 0;  // first 16-byte vector of phase C

    // partitioned solve: mirror a boundary row's new p on every neighbour that keeps it as a ghost
    auto push_row = [&](int64_t row, R val) {
        if (row < hp.n_int) return;
        const int* t = hp.push_tab + (size_t)(row - hp.n_int) * hp.nnb;
        for (int b = 0; b < hp.nnb; b++) {
            const int d = t[b];
            if (d >= 0) reinterpret_cast<R*>(hp.base[hp.nb_rank[b]] + hp.off_vec[vec_p][hp.nb_rank[b]])[d] = val;
        }
    };
    int it = c.iter1, iters = 0;  // 0 at the start of a solve; pcg_dual.cuh hands a running solve over with its count
    const int it0 = it;
    R rr = c.rn2;
    bool failed = false;
    const bool prof = a.prof != nullptr && blockIdx.x == 0 && threadIdx.x == 0;
    unsigned long long t_last = prof ? globaltimer_ns() : 0ull;
#define PCG_STAMP(slot_)                              \
    if (prof) {                                       \
        const unsigned long long t_now = globaltimer_ns(); \
        a.prof[slot_] += t_now - t_last;              \
        t_last = t_now;                               \
    }
    while (true) {
        // ---------------- phase A: y = A p, partial of p.y  (SparseDenseProduct.h:66-88, ConjugateGradient.h:66) ----
        R acc = (R)0;
        bool halo_ok = cm.world == 1;
        int k = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, k++) {
            if (!halo_ok && tile >= a.first_halo_tile) {
                if (threadIdx.x == 0 && !wait_flags(cm, hp, 0, halo_seq)) s_ctl[1] = 1;
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
                        ta += Aval[q] * px[a.col[q]];
                        tb2 += Aval[q + 1] * px[a.col[q + 1]];
                    }
                    if (q < ke) ta += Aval[q] * px[a.col[q]];
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
        R pap[1] = {block_sum_n<R, NW>(acc, red)};
        __syncthreads();
        prologue();  // next iteration's first tiles stream in while phases B / C run
        PCG_STAMP(0)
        if (!gsync_allreduce<R, 1, NW>(cm, ar_seq, part_pap, pap, red, s_pk, s_ctl, gs_target)) {
            failed = true;
            break;
        }
        PCG_STAMP(1)

        // ---------------- phase B: r -= alpha y, |r|^2, r.z  (ConjugateGradient.h:67-76) --------------------------
        // x += alpha p is deferred to phase C, which reads p anyway (one vector read less per iteration; same
        // operations on the same operands, so the same bits).
        const R alpha = abs_old / pap[0];
        R prr = (R)0, prz = (R)0;
        for (int64_t i = gtid; i < nvec; i += gthreads) {
            R rv[N], yv[N], dv[N];
            vload<R>(rs + i * N, rv);
            vload<R>(a.y + i * N, yv);
            vload<R>(dinv + i * N, dv);
#pragma unroll
            for (int q = 0; q < N; q++) {
                rv[q] = O::sub(rv[q], O::mul(alpha, yv[q]));
                R z = O::mul(dv[q], rv[q]);
                prr += rv[q] * rv[q];
                prz += rv[q] * z;
            }
            vstore<R>(rs + i * N, rv);
        }
        if (blockIdx.x == 0) {
            for (int64_t i = nvec * N + threadIdx.x; i < n; i += blockDim.x) {
                R rn = O::sub(rs[i], O::mul(alpha, a.y[i]));
                rs[i] = rn;
                R z = O::mul(dinv[i], rn);
                prr += rn * rn;
                prz += rn * z;
            }
        }
        R t2[2] = {prr, prz};
        block_sum_vec<R, 2, NW>(t2, red);
        PCG_STAMP(3)
        if (!gsync_allreduce<R, 2, NW>(cm, ar_seq, part_rr, t2, red, s_pk, s_ctl, gs_target)) {  // part_rz = part_rr + gridDim.x
            failed = true;
            break;
        }
        PCG_STAMP(4)
        rr = t2[0];
        const R rz = t2[1];
        const bool first = it == 1;  // x0 = 0 is never stored: the first update writes 0 + alpha p (same bits as x += alpha p on zeros)

        // ---------------- convergence test: same sums, same decision in every CTA (ConjugateGradient.h:72-73) -----
        const bool conv = rr < thr;
        if (conv || it >= maxit || !(rr <= O::huge())) {  // also stop on NaN / Inf instead of spinning to 2n
            iters = conv ? it - 1 : it;
            if (!conv) abs_old = rz;
            // the deferred x += alpha p of this last iteration
            for (int64_t i = gtid; i < nvec; i += gthreads) {
                R xv[N], pv[N];
                if (first) {
#pragma unroll
                    for (int q = 0; q < N; q++) xv[q] = (R)0;
                } else {
                    vload<R>(xs + i * N, xv);
                }
                vload<R>(px + i * N, pv);
#pragma unroll
                for (int q = 0; q < N; q++) xv[q] = O::add(xv[q], O::mul(alpha, pv[q]));
                vstore<R>(xs + i * N, xv);
            }
            if (blockIdx.x == 0)
                for (int64_t i = nvec * N + threadIdx.x; i < n; i += blockDim.x) xs[i] = O::add(first ? (R)0 : xs[i], O::mul(alpha, px[i]));
            break;
        }
        // ---------------- phase C: x += alpha p; p = D^-1 r + beta p  (ConjugateGradient.h:68, :77-80) -----------
        const R beta = rz / abs_old;
        abs_old = rz;
        bool pushed = false;
        for (int64_t j = gtid; j < nvec; j += gthreads) {
            int64_t i = j + c_start;  // ascending, boundary rows (the last ones) first on a partition
            if (i >= nvec) i -= nvec;
            if (a.zigzag == 1) i = nvec - 1 - j;
            R xv[N], rv[N], pv[N], dv[N];
            if (first) {
#pragma unroll
                for (int q = 0; q < N; q++) xv[q] = (R)0;
            } else {
                vload<R>(xs + i * N, xv);
            }
            vload<R>(rs + i * N, rv);
            vload<R>(px + i * N, pv);
            vload<R>(dinv + i * N, dv);
#pragma unroll
            for (int q = 0; q < N; q++) {
                xv[q] = O::add(xv[q], O::mul(alpha, pv[q]));
                pv[q] = O::add(O::mul(dv[q], rv[q]), O::mul(beta, pv[q]));
            }
            vstore<R>(xs + i * N, xv);
            vstore<R>(px + i * N, pv);
            if (cm.world > 1 && i * N + N > hp.n_int) {  // boundary rows: straight into the neighbours' ghost tails
#pragma unroll
                for (int q = 0; q < N; q++) push_row(i * N + q, pv[q]);
                pushed = true;
            }
        }
        if (blockIdx.x == 0)
            for (int64_t i = nvec * N + threadIdx.x; i < n; i += blockDim.x) {
                const R pi = px[i];
                xs[i] = O::add(first ? (R)0 : xs[i], O::mul(alpha, pi));
                const R pn = O::add(O::mul(dinv[i], rs[i]), O::mul(beta, pi));
                px[i] = pn;
                if (cm.world > 1) {
                    push_row(i, pn);
                    pushed = true;
                }
            }
        // remote stores are complete system-wide before this CTA arrives at the sync; the last CTA to arrive then raises the
        // neighbours' halo flags with release stores (comm.cuh gsync_barrier)
        if (cm.world > 1 && __syncthreads_or(pushed ? 1 : 0)) __threadfence_system();
        PCG_STAMP(6)
        halo_seq += 1;
        if (!gsync_barrier(cm, hp, bar_seq, cm.world > 1 ? halo_seq : 0ull, s_ctl, gs_target)) {
            failed = true;
            break;
        }
        PCG_STAMP(7)
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
        st->iters = failed ? it : iters;
        st->done = 1;
        if (failed) cm.loc->error = 1;
    }
}

}  // namespace feab
