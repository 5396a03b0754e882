// Gray-Scott: the u-solve and the v-solve of one time step (CPUSolver.cpp:139-140) as ONE persistent kernel, for as long as
// both are iterating.  Both right-hand sides are built from the old fields (:134-137) and A_u = M/dt + Du K, A_v = M/dt + Dv K
// share rowptr / columns, so the two Jacobi-PCG recurrences (ConjugateGradient.h:30-85, each with its own alpha, beta and
// stopping test) can advance in lock-step:
//   * one index stream (rowptr + columns, 32 B/row) feeds both SpMVs: a 256-row tile is staged once together with BOTH value
//     streams; threads 0..255 of the CTA multiply system 0, threads 256..511 system 1;
//   * the three grid barriers of an iteration are shared.
// Per shared iteration this moves 336 instead of 384 B/row and pays 3 instead of 6 barriers.  As soon as one system stops
// (converged / maxit / non-finite), the kernel finishes the other one's iteration, stores its state and exits; the
// single-system kernel (pcg_persistent.cuh) resumes it from CgState (abs_new[0], iter1).  Every operation on a system is
// the one pcg_persistent performs; only the grouping of the per-CTA partial sums differs (256-row tiles), so results agree
// with the other loop drivers to summation order, not bit for bit.  On a partitioned handle the halo exchange and the
// all-reduces are shared too: one flag value per iteration covers both directions, one mailbox round carries both p.Ap
// (2 values) and one both |r|^2 and r.z (4 values).
// x (phase C only) and the phase-B read of y are streaming accesses (ld/st.cs): touched once per iteration, they should
// not push r, p, dinv of two systems out of L2.  The 2 + 4 + 4 sums of an iteration are reduced as vectors (common.cuh).
#pragma once
#include "pcg_persistent.cuh"

namespace feab {

template <typename R>
struct PcgDualArgs {
    const int* rowptr;
    const int* col;
    const R* A[2];
    const R* dinv[2];
    R* x[2];
    R* r[2];
    R* p[2];
    R* y[2];
    int64_t n;
    int num_tiles;
    CgState<R>* st;           // [2]: written here (the start of both solves is taken from the RHS kernel's published sums)
    R* part;                  // 6 x gridDim.x: pap0 | pap1 | rr0 | rr1 | rz0 | rz1
    int first_halo_tile;      // partitioned: first tile with a row that reads a ghost column
    R tol;
    int maxit;
    int zigzag;  // see PcgArgs::zigzag
    unsigned long long* prof; // optional (FEAB_PCG_PROF=2): ns spent by CTA 0, same slots as PcgArgs::prof
    unsigned long long* arrivals;  // optional (FEAB_PCG_PROF=3): globaltimer of every CTA when it enters / leaves each of the three grid
                                   // syncs of the LAST iteration, [sync][enter, leave][gridDim.x] -- how much of a sync is waiting for the slowest CTA
};

template <typename R, int ROWS, int CPR, int STAGES>
struct DualCfg {
    static constexpr int kCap = ROWS * CPR;
    static constexpr size_t kValBytes = ((size_t)(kCap + 4) * sizeof(R) + 127) / 128 * 128;
    static constexpr size_t kColBytes = ((size_t)(kCap + 4) * 4 + 127) / 128 * 128;
    static constexpr size_t kRpBytes = ((size_t)(ROWS + 8) * 4 + 127) / 128 * 128;
    static constexpr size_t kStageBytes = 2 * kValBytes + kColBytes + kRpBytes;
    static constexpr size_t kSmemBytes = kStageBytes * STAGES + 128 + 128;
};

template <typename R, int ROWS, int CPR, int STAGES, int MINB>
__global__ void __launch_bounds__(2 * ROWS, MINB) pcg_dual(PcgDualArgs<R> a, Comm cm, HaloPlan hp) {
    using Cfg = DualCfg<R, ROWS, CPR, STAGES>;
    using O = Ops<R>;
    constexpr int NWT = 2 * ROWS / 32;  // warps per CTA
    constexpr int NWH = ROWS / 32;      // warps per half
    extern __shared__ unsigned char smraw_dual[];
    unsigned char* base = reinterpret_cast<unsigned char*>(((uintptr_t)smraw_dual + 127) & ~(uintptr_t)127);
    uint64_t* full = reinterpret_cast<uint64_t*>(base + Cfg::kStageBytes * STAGES);
    __shared__ R red[(NWT + 1) * 4];
    __shared__ double s_v[kMaxRanks * 4];
    __shared__ int s_ctl[2];

    if (comm_failed(cm)) return;  // an earlier kernel of this call lost a peer: fail closed
    if (threadIdx.x == 0) s_ctl[1] = 0;
    unsigned long long ar_seq = *(volatile unsigned long long*)&cm.loc->ar_seq;
    unsigned long long halo_seq = *(volatile unsigned long long*)&cm.loc->halo_seq;
    unsigned long long bar_seq = ld_volatile_u64(&cm.loc->gs_release);
    unsigned long long gs_target = ld_volatile_u64(&cm.loc->gs_base);  // one GPU: arrival count that completes the next grid sync (comm.cuh)
    __syncthreads();
    const int64_t n = a.n;
    CgState<R> c2[2];
    if (!pcg_begin<R, 2>(cm, ar_seq, a.tol, a.maxit, a.st, c2, s_v, s_ctl)) return;
    if (c2[0].done || c2[1].done) {  // one solve ended before it began: the single-system kernel takes the other one (uniform)
        for (int s = 0; s < 2; s++)
            if (c2[s].done) zero_solution<R>(a.x[s], n);
        return;
    }
    R thr[2] = {c2[0].thr, c2[1].thr};
    const int maxit = a.maxit;
    R abs_old[2] = {c2[0].abs_new[0], c2[1].abs_new[0]};
    const int num_tiles = a.num_tiles;
    const int h = (int)threadIdx.x / ROWS;   // which system this thread multiplies
    const int t = (int)threadIdx.x % ROWS;   // its row inside the tile

    auto stage_val = [&](int s, int sys) { return reinterpret_cast<R*>(base + (size_t)s * Cfg::kStageBytes + (size_t)sys * Cfg::kValBytes); };
    auto stage_col = [&](int s) { return reinterpret_cast<int*>(base + (size_t)s * Cfg::kStageBytes + 2 * Cfg::kValBytes); };
    auto stage_rp = [&](int s) { return reinterpret_cast<int*>(base + (size_t)s * Cfg::kStageBytes + 2 * Cfg::kValBytes + Cfg::kColBytes); };

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
        mbar_expect_tx(&full[s], rp_b + (staged ? (uint32_t)cnt_al * (uint32_t)(2 * sizeof(R) + 4) : 0u));
        tma_load_1d(stage_rp(s), a.rowptr + r0, rp_b, &full[s], policy);
        if (staged) {
            tma_load_1d(stage_val(s, 0), a.A[0] + tba, (uint32_t)cnt_al * (uint32_t)sizeof(R), &full[s], policy);
            tma_load_1d(stage_val(s, 1), a.A[1] + tba, (uint32_t)cnt_al * (uint32_t)sizeof(R), &full[s], policy);
            tma_load_1d(stage_col(s), a.col + tba, (uint32_t)cnt_al * 4u, &full[s], policy);
        }
    };
    unsigned fills = 0;
    auto prologue = [&]() {
        if (threadIdx.x == 0) {
            int tl = blockIdx.x;
            for (int s = 0; s < STAGES && tl < num_tiles; s++, tl += gridDim.x) issue(tl, (int)((fills + s) % STAGES));
        }
    };
    prologue();

    // partitioned solve: mirror a boundary row's new p_s on every neighbour that keeps it as a ghost
    auto push_row = [&](int sys, int64_t row, R val) {
        if (row < hp.n_int) return;
        const int* tb = hp.push_tab + (size_t)(row - hp.n_int) * hp.nnb;
        const int vec = sys == 0 ? kVecP0 : kVecP1;
        for (int b = 0; b < hp.nnb; b++) {
            const int d = tb[b];
            if (d >= 0) reinterpret_cast<R*>(hp.base[hp.nb_rank[b]] + hp.off_vec[vec][hp.nb_rank[b]])[d] = val;
        }
    };
    const int G = (int)gridDim.x;
    R* const part = a.part;
    constexpr int N = Vec<R>::N;
    const int64_t nvec = n / N;
    const int64_t gthreads = (int64_t)gridDim.x * blockDim.x;
    const int64_t gtid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t c_start = (a.zigzag == 2 && cm.world > 1) ? (min(nvec, (int64_t)hp.n_int / N) & ~(int64_t)63) : 0;  // see PcgArgs::zigzag
    const R* const Ah = a.A[h];
    R* const ph = a.p[h];
    R* const yh = a.y[h];

    int it = 0;
    const int it0 = it;
    bool failed = false;
    int iters[2] = {0, 0};
    bool fin[2] = {false, false};
    R rr[2] = {(R)0, (R)0};
    const bool prof = a.prof != nullptr && blockIdx.x == 0 && threadIdx.x == 0;
    unsigned long long t_last = prof ? globaltimer_ns() : 0ull;
#define DUAL_ARRIVAL(idx)                                                                                   \
    if (a.arrivals != nullptr && threadIdx.x == 0) a.arrivals[(size_t)(idx)*gridDim.x + blockIdx.x] = globaltimer_ns();
#define DUAL_STAMP(slot)                                   \
    if (prof) {                                            \
        const unsigned long long t_now = globaltimer_ns(); \
        a.prof[slot] += t_now - t_last;                    \
        t_last = t_now;                                    \
    }
    while (true) {
        // ---------------- phase A: y_s = A_s p_s for both systems from one staged index stream ----------------
        R acc = (R)0;
        bool halo_ok = cm.world == 1;
        int k = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, k++) {
            if (!halo_ok && tile >= a.first_halo_tile) {  // both directions' ghosts arrive under one flag value
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
            if (t < rows) {
                const int64_t r = r0 + t;
                const int kb = rp[t], ke = rp[t + 1];
                R ta = (R)0, tb2 = (R)0, xr = (R)0;
                if (staged) {
                    const R* sv = stage_val(s, h) + (kb - tba);
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
                    for (int j = 0; j < 8; j++) xv[j] = ph[cv[j]];
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
                        xr = ph[r];
                        for (int j = 8; j < len; j++) {
                            R pr = sv[j] * ph[sc[j]];
                            if (j & 1)
                                tb2 += pr;
                            else
                                ta += pr;
                        }
                    }
                } else {
                    int q = kb;
                    for (; q + 1 < ke; q += 2) {
                        ta += Ah[q] * ph[a.col[q]];
                        tb2 += Ah[q + 1] * ph[a.col[q + 1]];
                    }
                    if (q < ke) ta += Ah[q] * ph[a.col[q]];
                    xr = ph[r];
                }
                const R sum = ta + tb2;
                yh[r] = sum;
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
        // per-half sums in a fixed order: warps [0, NWH) belong to system 0, [NWH, 2 NWH) to system 1
        acc = warp_sum(acc);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
        __syncthreads();
        R pap[2] = {(R)0, (R)0};
        if (threadIdx.x == 0) {
#pragma unroll
            for (int s = 0; s < 2; s++)
                for (int w = 0; w < NWH; w++) pap[s] += red[s * NWH + w];
        }
        __syncthreads();
        prologue();
        DUAL_STAMP(0)
        DUAL_ARRIVAL(0)
        if (!gsync_allreduce<R, 2, NWT>(cm, ar_seq, part, pap, red, s_v, s_ctl, gs_target)) {
            failed = true;
            break;
        }
        DUAL_ARRIVAL(1)
        DUAL_STAMP(1)

        // ---------------- phase B (both systems): r -= alpha y, |r|^2, r.z ----------------
        R alpha[2];
        R pb[4];  // |r_0|^2, |r_1|^2, r_0.z_0, r_1.z_1 of this CTA
#pragma unroll
        for (int s = 0; s < 2; s++) {
            alpha[s] = abs_old[s] / pap[s];
            R prr = (R)0, prz = (R)0;
            R* const rs = a.r[s];
            const R* const ys = a.y[s];
            const R* const ds = a.dinv[s];
            for (int64_t i = gtid; i < nvec; i += gthreads) {
                R rv[N], yv[N], dv[N];
                vload<R>(rs + i * N, rv);
                vload_cs<R>(ys + i * N, yv);
                vload<R>(ds + i * N, dv);
#pragma unroll
                for (int q = 0; q < N; q++) {
                    rv[q] = O::sub(rv[q], O::mul(alpha[s], yv[q]));
                    R z = O::mul(dv[q], rv[q]);
                    prr += rv[q] * rv[q];
                    prz += rv[q] * z;
                }
                vstore<R>(rs + i * N, rv);
            }
            if (blockIdx.x == 0) {
                for (int64_t i = nvec * N + threadIdx.x; i < n; i += blockDim.x) {
                    R rn = O::sub(rs[i], O::mul(alpha[s], ys[i]));
                    rs[i] = rn;
                    R z = O::mul(ds[i], rn);
                    prr += rn * rn;
                    prz += rn * z;
                }
            }
            pb[s] = prr;
            pb[2 + s] = prz;
        }
        block_sum_vec<R, 4, NWT>(pb, red);
        DUAL_STAMP(3)
        DUAL_ARRIVAL(2)
        if (!gsync_allreduce<R, 4, NWT>(cm, ar_seq, part + (size_t)2 * G, pb, red, s_v, s_ctl, gs_target)) {
            failed = true;
            break;
        }
        DUAL_ARRIVAL(3)
        DUAL_STAMP(4)
        R rz[2];
        rr[0] = pb[0];
        rr[1] = pb[1];
        rz[0] = pb[2];
        rz[1] = pb[3];
        const bool first = it == 1;  // x0 = 0 is never stored: the first update writes 0 + alpha p
        bool pushed = false;
        // ---------------- stopping tests (ConjugateGradient.h:72-73), per system; direction update for the ones that go on ----
#pragma unroll
        for (int s = 0; s < 2; s++) {
            const bool conv = rr[s] < thr[s];
            fin[s] = conv || it >= maxit || !(rr[s] <= O::huge());
            R* const xs = a.x[s];
            R* const ps = a.p[s];
            if (fin[s]) {
                iters[s] = conv ? it - 1 : it;
                if (!conv) abs_old[s] = rz[s];
                for (int64_t i = gtid; i < nvec; i += gthreads) {  // the deferred x += alpha p of the last iteration
                    R xv[N], pv[N];
                    if (first) {
#pragma unroll
                        for (int q = 0; q < N; q++) xv[q] = (R)0;
                    } else {
                        vload_cs<R>(xs + i * N, xv);
                    }
                    vload<R>(ps + i * N, pv);
#pragma unroll
                    for (int q = 0; q < N; q++) xv[q] = O::add(xv[q], O::mul(alpha[s], pv[q]));
                    vstore_cs<R>(xs + i * N, xv);
                }
                if (blockIdx.x == 0)
                    for (int64_t i = nvec * N + threadIdx.x; i < n; i += blockDim.x) xs[i] = O::add(first ? (R)0 : xs[i], O::mul(alpha[s], ps[i]));
            } else {
                const R beta = rz[s] / abs_old[s];
                abs_old[s] = rz[s];
                const R* const rs = a.r[s];
                const R* const ds = a.dinv[s];
                for (int64_t j = gtid; j < nvec; j += gthreads) {
                    int64_t i = j + c_start;  // ascending, boundary rows (the last ones) first on a partition
                    if (i >= nvec) i -= nvec;
                    if (a.zigzag == 1) i = nvec - 1 - j;
                    R xv[N], rv[N], pv[N], dv[N];
                    if (first) {
#pragma unroll
                        for (int q = 0; q < N; q++) xv[q] = (R)0;
                    } else {
                        vload_cs<R>(xs + i * N, xv);
                    }
                    vload<R>(rs + i * N, rv);
                    vload<R>(ps + i * N, pv);
                    vload<R>(ds + i * N, dv);
#pragma unroll
                    for (int q = 0; q < N; q++) {
                        xv[q] = O::add(xv[q], O::mul(alpha[s], pv[q]));
                        pv[q] = O::add(O::mul(dv[q], rv[q]), O::mul(beta, pv[q]));
                    }
                    vstore_cs<R>(xs + i * N, xv);
                    vstore<R>(ps + i * N, pv);
                    if (cm.world > 1 && i * N + N > hp.n_int) {  // boundary rows: straight into the neighbours' ghost tails
#pragma unroll
                        for (int q = 0; q < N; q++) push_row(s, i * N + q, pv[q]);
                        pushed = true;
                    }
                }
                if (blockIdx.x == 0)
                    for (int64_t i = nvec * N + threadIdx.x; i < n; i += blockDim.x) {
                        const R pi = ps[i];
                        xs[i] = O::add(first ? (R)0 : xs[i], O::mul(alpha[s], pi));
                        const R pn = O::add(O::mul(ds[i], rs[i]), O::mul(beta, pi));
                        ps[i] = pn;
                        if (cm.world > 1) {
                            push_row(s, i, pn);
                            pushed = true;
                        }
                    }
            }
        }
        DUAL_STAMP(6)
        if (fin[0] && fin[1]) break;
        if (cm.world == 1) {
            if (fin[0] || fin[1]) break;  // the kernel boundary is the barrier the survivor's next SpMV needs
            DUAL_ARRIVAL(4)
            if (!gsync_barrier(cm, hp, bar_seq, 0ull, s_ctl, gs_target)) {
                failed = true;
                break;
            }
            DUAL_ARRIVAL(5)
            DUAL_STAMP(7)
        } else {
            // same hand-shake as pcg_persistent.cuh: pushers fence, grid sync, the last CTA to arrive raises the flags; one
            // flag value covers both directions.  A survivor's last push here is what its first SpMV in the single-system
            // kernel waits for.
            if (__syncthreads_or(pushed ? 1 : 0)) __threadfence_system();
            halo_seq += 1;
            DUAL_ARRIVAL(4)
            if (!gsync_barrier(cm, hp, bar_seq, halo_seq, s_ctl, gs_target)) {
                failed = true;
                break;
            }
            DUAL_ARRIVAL(5)
            DUAL_STAMP(7)
            if (fin[0] || fin[1]) break;
        }
    }
#undef DUAL_STAMP
#undef DUAL_ARRIVAL
    if (prof) a.prof[9] += (unsigned long long)(it - it0);
    // drain the prologue issued for an iteration that does not run here
    {
        int k = 0;
        for (int tile = blockIdx.x; tile < num_tiles && k < STAGES; tile += gridDim.x, k++) {
            const unsigned f = fills + (unsigned)k;
            mbar_wait(&full[f % STAGES], (f / STAGES) & 1u);
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        if (failed) cm.loc->error = 1;
        for (int s = 0; s < 2; s++) {
            CgState<R>* st = a.st + s;
            st->rn2 = rr[s];
            st->iter1 = it;
            if (fin[s] || failed) {
                st->abs_new[it & 1] = abs_old[s];
                st->iters = iters[s];
                st->done = 1;
            } else {
                st->abs_new[0] = abs_old[s];  // pcg_persistent resumes from here
            }
        }
    }
}

}  // namespace feab
