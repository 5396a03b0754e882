// K7 as ONE persistent kernel: the whole Jacobi right-preconditioned BiCGSTAB solve of the non-symmetric advection-diffusion
// operator (lib/eigen/Eigen/src/IterativeLinearSolvers/BiCGSTAB.h:32-119, with the two stated deviations of bicgstab.cuh /
// oracle/fea_oracle_impl.h: relative stopping test, deterministic restart) inside a single cooperative launch, on the
// skeleton of pcg_persistent.cuh.  One loop trip walks five phases separated by grid-wide syncs:
//   P   beta;  p = r + beta (p - w v);  y = D^-1 p                      (:79-82)   -> barrier (+ halo flags of y)
//   S1  v = A y (TMA tile pipeline), partials r0.v, v.v                   (:84-87)   -> all-reduce
//   S   alpha = rho / r0.v;  s = r - alpha v;  z = D^-1 s                 (:99-102)  -> barrier (+ halo flags of z)
//   S2  t = A z, partials t.s, t.t                                        (:103-105) -> all-reduce
//   X   w = t.s / t.t;  x += alpha y + w z;  r = s - w t;  |r|^2, r0.r   (:106-113) -> all-reduce
// and the two breakdown restarts (:68-77, :89-97) run r = b - A x through the same tile pipeline.  Every CTA of every rank
// takes every decision from the same bit-identical sums (comm.cuh gsync_allreduce), so the control flow is uniform across
// the grid and across the ranks; no host round trip, no WHILE graph.  On a partition the boundary rows of y and z are stored
// straight into the neighbours' ghost tails by the phase that produces them and the flags are raised by the last CTA to
// arrive at the barrier; the SpMV waits for them only before its first boundary tile.
// The start (|b|, early-outs: :44-49) is taken from the sums the RHS kernel published; r0 = r and the saved b are written by
// the first P phase, p = v = 0 and x = 0 are never stored (the first trip treats them as zeros).
// Row sums use ONE accumulator in ascending column order with separately rounded products, like spmv_plain / the oracle's
// spmv_n (the column walk of a column-major matrix adds into y[row] in that order).
#pragma once
#include "bicgstab.cuh"
#include "comm.cuh"
#include "common.cuh"
#include "pcg_persistent.cuh"
#include "solver.cuh"
#include "spmv_tma.cuh"

namespace feab {

template <typename R>
struct BicgArgs {
    const int* rowptr;
    const int* col;
    const R* A;
    const R* dinv;
    R *x, *r, *r0, *v, *p, *y, *s, *z, *t, *b;   // y, z and x live in the comm arena (ghost tails): kVecP0, kVecP1, kVecU
    int64_t n;
    int num_tiles;
    int first_halo_tile;
    BicgState<R>* st;
    R* part;          // 4 x gridDim.x
    R tol;
    int maxit;
    unsigned long long* prof;  // optional: ns spent by CTA 0 in [S1+S2, all-reduces, -, P+S+X, barriers, ...], trips in [9]
};

template <typename R, int ROWS, int CPR, int STAGES, int MINB>
__global__ void __launch_bounds__(ROWS, MINB) bicg_persistent(BicgArgs<R> a, Comm cm, HaloPlan hp) {
    using Cfg = TmaCfg<R, ROWS, CPR, STAGES>;
    using O = Ops<R>;
    constexpr int NW = ROWS / 32;
    constexpr int N = Vec<R>::N;
    extern __shared__ unsigned char smraw_bicg[];
    unsigned char* base = reinterpret_cast<unsigned char*>(((uintptr_t)smraw_bicg + 127) & ~(uintptr_t)127);
    uint64_t* full = reinterpret_cast<uint64_t*>(base + Cfg::kStageBytes * STAGES);
    __shared__ R red[(NW + 1) * 2];
    __shared__ double s_pk[kMaxRanks * 4];
    __shared__ int s_ctl[2];

    if (comm_failed(cm)) return;
    if (threadIdx.x == 0) s_ctl[1] = 0;
    unsigned long long ar_seq = *(volatile unsigned long long*)&cm.loc->ar_seq;
    unsigned long long halo_seq = *(volatile unsigned long long*)&cm.loc->halo_seq;
    unsigned long long bar_seq = ld_volatile_u64(&cm.loc->gs_release);
    unsigned long long gs_target = ld_volatile_u64(&cm.loc->gs_base);  // one GPU: arrival count that completes the next grid sync (comm.cuh)
    __syncthreads();

    const int64_t n = a.n;
    const int num_tiles = a.num_tiles;
    const int64_t nvec = n / N;
    const int64_t gthreads = (int64_t)gridDim.x * blockDim.x;
    const int64_t gtid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool lead = blockIdx.x == 0 && threadIdx.x == 0;

    // ---- start of the solve: |b|^2 from the RHS kernel's all-reduce (BiCGSTAB.h:41-49; x0 = 0 so r = b) ----
    R bb = (R)0;
    {
        if ((int)threadIdx.x < cm.world) {
            const ArSlot* mine = &cm.ctrl[cm.rank]->slot[ar_seq & 1][threadIdx.x];
            bool ok = true;
            s_pk[threadIdx.x * 4] = ll_wait(mine, 0, ar_seq, cm.loc, &ok, (int)threadIdx.x);
            if (!ok) s_ctl[1] = 1;
            __threadfence();
        }
        __syncthreads();
        if (s_ctl[1]) return;
        for (int w = 0; w < cm.world; w++) bb += (R)s_pk[w * 4];
        __syncthreads();
    }
    const R rhs_norm = O::sqrt(bb);
    R r0_norm = rhs_norm, r_norm = rhs_norm;
    if (rhs_norm == (R)0 || !(rhs_norm > a.tol * rhs_norm) || a.maxit <= 0) {
        zero_solution<R>(a.x, n);
        if (lead) {
            BicgState<R> c;
            c.rho[0] = c.rho[1] = (R)1;
            c.alpha = (R)0;
            c.w = (R)1;
            c.r0_norm = c.r_norm = c.rhs_norm = rhs_norm;
            c.tol = a.tol;
            c.iters = c.trips = c.restarts = c.restart_req = c.restart_now = c.skip = 0;
            c.done = 1;
            c.maxit = a.maxit;
            *a.st = c;
        }
        return;
    }

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
    unsigned fills = 0;
    auto prologue = [&]() {  // first STAGES tiles of the NEXT product: the matrix never changes, so they stream in during the vector phases
        if (threadIdx.x == 0) {
            int t = blockIdx.x;
            for (int s = 0; s < STAGES && t < num_tiles; s++, t += gridDim.x) issue(t, (int)((fills + s) % STAGES));
        }
    };
    prologue();

    bool failed = false;
    const bool prof = a.prof != nullptr && lead;
    unsigned long long t_last = prof ? globaltimer_ns() : 0ull;
#define BICG_STAMP(slot_)                                   \
    if (prof) {                                             \
        const unsigned long long t_now = globaltimer_ns();  \
        a.prof[slot_] += t_now - t_last;                    \
        t_last = t_now;                                     \
    }

    // dst = A src (+ partials dotv.dst and dst.dst); src is gathered with plain (coherent) loads: it changes inside this kernel
    auto spmv = [&](const R* src, R* dst, const R* dotv, R (&d)[2]) {
        R d0 = (R)0, d1 = (R)0;
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
                R sum = (R)0;
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
                    for (int j = 0; j < 8; j++) xv[j] = src[cv[j]];
#pragma unroll
                    for (int j = 0; j < 8; j++)
                        if (j < len) sum = O::add(sum, O::mul(av[j], xv[j]));
                    for (int j = 8; j < len; j++) sum = O::add(sum, O::mul(sv[j], src[sc[j]]));
                } else {
                    for (int q = kb; q < ke; q++) sum = O::add(sum, O::mul(a.A[q], src[a.col[q]]));
                }
                dst[r] = sum;
                if (dotv != nullptr) {
                    d0 += dotv[r] * sum;
                    d1 += sum * sum;
                }
            }
            __syncthreads();  // every consumer is done with stage s
            if (threadIdx.x == 0) {
                const int next = tile + STAGES * gridDim.x;
                if (next < num_tiles) issue(next, s);
            }
        }
        fills += (unsigned)k;
        d[0] = d0;
        d[1] = d1;
        __syncthreads();
        prologue();
    };
    // boundary row `row` of arena vector `vec` -> every neighbour that keeps it as a ghost
    auto push_row = [&](int vec, int64_t row, R val) {
        if (row < hp.n_int) return;
        push_boundary_row<R>(hp, vec, row, val);
    };
    auto barrier_with_halo = [&](bool pushed) -> bool {
        if (cm.world > 1 && __syncthreads_or(pushed ? 1 : 0)) __threadfence_system();
        halo_seq += 1;
        return gsync_barrier(cm, hp, bar_seq, cm.world > 1 ? halo_seq : 0ull, s_ctl, gs_target);
    };
    // r = b - A x, r0 = r, returns |r|^2 (both breakdown restarts, BiCGSTAB.h:71-72 and :92; deterministic r0 = r)
    bool x_unset = true;  // x = 0 has not been stored yet
    auto restart_residual = [&](R& rr_out) -> bool {
        if (x_unset) {
            zero_solution<R>(a.x, n);
            x_unset = false;
        }
        bool pushed = false;
        if (cm.world > 1) {
            for (int64_t i = hp.n_int + gtid; i < n; i += gthreads) {
                push_row(kVecU, i, a.x[i]);
                pushed = true;
            }
        }
        if (!barrier_with_halo(pushed)) return false;
        R d[2];
        spmv(a.x, a.t, nullptr, d);
        if (!gsync_barrier(cm, hp, bar_seq, 0ull, s_ctl, gs_target)) return false;   // t complete (rows of a tile are not the rows of the vector loop)
        R rr = (R)0;
        for (int64_t i = gtid; i < n; i += gthreads) {
            const R rn = O::sub(a.b[i], a.t[i]);
            a.r[i] = rn;
            a.r0[i] = rn;
            rr += rn * rn;
        }
        R t1[1] = {block_sum_n<R, NW>(rr, red)};
        __syncthreads();
        if (!gsync_allreduce<R, 1, NW>(cm, ar_seq, a.part, t1, red, s_pk, s_ctl, gs_target)) return false;
        rr_out = t1[0];
        return true;
    };

    R rho = (R)1, alpha = (R)0, w = (R)1;
    R r0r = bb;           // r0.r of the current residual (r0 = r = b at the start)
    int it = 0, trips = 0, restarts = 0;
    bool first = true;    // p = v = 0, r0 / b not stored yet
    while (r_norm > a.tol * rhs_norm && it < a.maxit) {
        trips++;
        const R rho_old = rho;
        rho = r0r;
        {
            const R mx = r0_norm > r_norm ? r0_norm : r_norm, mn = r0_norm < r_norm ? r0_norm : r_norm;
            if (!first && fabs((double)rho) / (double)mx < (double)(O::eps() * mn)) {  // r too orthogonal to r0 (:68-77)
                R rr;
                if (!restart_residual(rr)) {
                    failed = true;
                    break;
                }
                rho = rr;
                r0_norm = O::sqrt(rr);
                alpha = (R)0;
                w = (R)1;
                if (restarts++ == 0) it = 0;
            }
        }
        const R beta = (rho / rho_old) * (alpha / w);
        // ---------------- phase P: p = r + beta (p - w v); y = D^-1 p ------------------------------------------------
        bool pushed = false;
        for (int64_t i = gtid; i < nvec; i += gthreads) {
            R rv[N], pv[N], vv[N], dv[N], yv[N];
            vload<R>(a.r + i * N, rv);
            vload<R>(a.dinv + i * N, dv);
            if (first) {
                vstore<R>(a.r0 + i * N, rv);
                vstore<R>(a.b + i * N, rv);
#pragma unroll
                for (int q = 0; q < N; q++) pv[q] = rv[q];
            } else {
                vload<R>(a.p + i * N, pv);
                vload<R>(a.v + i * N, vv);
#pragma unroll
                for (int q = 0; q < N; q++) pv[q] = O::add(rv[q], O::mul(beta, O::sub(pv[q], O::mul(w, vv[q]))));
            }
#pragma unroll
            for (int q = 0; q < N; q++) yv[q] = O::mul(dv[q], pv[q]);
            vstore<R>(a.p + i * N, pv);
            vstore<R>(a.y + i * N, yv);
            if (cm.world > 1 && i * N + N > hp.n_int) {
#pragma unroll
                for (int q = 0; q < N; q++) push_row(kVecP0, i * N + q, yv[q]);
                pushed = true;
            }
        }
        if (blockIdx.x == 0)
            for (int64_t i = nvec * N + threadIdx.x; i < n; i += blockDim.x) {
                const R ri = a.r[i];
                R pn;
                if (first) {
                    a.r0[i] = ri;
                    a.b[i] = ri;
                    pn = ri;
                } else {
                    pn = O::add(ri, O::mul(beta, O::sub(a.p[i], O::mul(w, a.v[i]))));
                }
                const R yn = O::mul(a.dinv[i], pn);
                a.p[i] = pn;
                a.y[i] = yn;
                if (cm.world > 1) {
                    push_row(kVecP0, i, yn);
                    pushed = true;
                }
            }
        first = false;
        BICG_STAMP(3)
        if (!barrier_with_halo(pushed)) {
            failed = true;
            break;
        }
        BICG_STAMP(4)
        // ---------------- S1: v = A y, theta = r0.v, |v|^2 -----------------------------------------------------------
        R d[2];
        spmv(a.y, a.v, a.r0, d);
        block_sum_vec<R, 2, NW>(d, red);
        BICG_STAMP(0)
        if (!gsync_allreduce<R, 2, NW>(cm, ar_seq, a.part, d, red, s_pk, s_ctl, gs_target)) {
            failed = true;
            break;
        }
        BICG_STAMP(1)
        const R theta = d[0], v_norm = O::sqrt(d[1]);
        {
            const R mx = r0_norm > v_norm ? r0_norm : v_norm, mn = r0_norm < v_norm ? r0_norm : v_norm;
            if (fabs((double)theta) / (double)mx < (double)(O::eps() * mn)) {  // small angle between r0 and v (:89-97)
                R rr;
                if (!restart_residual(rr)) {
                    failed = true;
                    break;
                }
                r0_norm = r_norm = O::sqrt(rr);
                r0r = rr;
                rho = (R)1;
                alpha = (R)0;
                w = (R)1;
                if (restarts++ == 0) it = 0;
                if (restarts > 8) break;
                continue;
            }
        }
        alpha = rho / theta;
        // ---------------- phase S: s = r - alpha v; z = D^-1 s ---------------------------------------------------------
        pushed = false;
        for (int64_t i = gtid; i < nvec; i += gthreads) {
            R rv[N], vv[N], dv[N], sv[N], zv[N];
            vload<R>(a.r + i * N, rv);
            vload<R>(a.v + i * N, vv);
            vload<R>(a.dinv + i * N, dv);
#pragma unroll
            for (int q = 0; q < N; q++) {
                sv[q] = O::sub(rv[q], O::mul(alpha, vv[q]));
                zv[q] = O::mul(dv[q], sv[q]);
            }
            vstore<R>(a.s + i * N, sv);
            vstore<R>(a.z + i * N, zv);
            if (cm.world > 1 && i * N + N > hp.n_int) {
#pragma unroll
                for (int q = 0; q < N; q++) push_row(kVecP1, i * N + q, zv[q]);
                pushed = true;
            }
        }
        if (blockIdx.x == 0)
            for (int64_t i = nvec * N + threadIdx.x; i < n; i += blockDim.x) {
                const R sn = O::sub(a.r[i], O::mul(alpha, a.v[i]));
                const R zn = O::mul(a.dinv[i], sn);
                a.s[i] = sn;
                a.z[i] = zn;
                if (cm.world > 1) {
                    push_row(kVecP1, i, zn);
                    pushed = true;
                }
            }
        BICG_STAMP(3)
        if (!barrier_with_halo(pushed)) {
            failed = true;
            break;
        }
        BICG_STAMP(4)
        // ---------------- S2: t = A z, t.s, t.t ------------------------------------------------------------------------
        spmv(a.z, a.t, a.s, d);
        block_sum_vec<R, 2, NW>(d, red);
        BICG_STAMP(0)
        if (!gsync_allreduce<R, 2, NW>(cm, ar_seq, a.part, d, red, s_pk, s_ctl, gs_target)) {
            failed = true;
            break;
        }
        BICG_STAMP(1)
        w = d[1] > (R)0 ? d[0] / d[1] : (R)0;
        // ---------------- phase X: x += alpha y + w z; r = s - w t; |r|^2, r0.r ----------------------------------------
        R prr = (R)0, prho = (R)0;
        for (int64_t i = gtid; i < nvec; i += gthreads) {
            R xv[N], yv[N], zv[N], sv[N], tv[N], r0v[N], rv[N];
            if (x_unset) {
#pragma unroll
                for (int q = 0; q < N; q++) xv[q] = (R)0;
            } else {
                vload<R>(a.x + i * N, xv);
            }
            vload<R>(a.y + i * N, yv);
            vload<R>(a.z + i * N, zv);
            vload<R>(a.s + i * N, sv);
            vload<R>(a.t + i * N, tv);
            vload<R>(a.r0 + i * N, r0v);
#pragma unroll
            for (int q = 0; q < N; q++) {
                xv[q] = O::add(xv[q], O::add(O::mul(alpha, yv[q]), O::mul(w, zv[q])));
                rv[q] = O::sub(sv[q], O::mul(w, tv[q]));
                prr += rv[q] * rv[q];
                prho += r0v[q] * rv[q];
            }
            vstore<R>(a.x + i * N, xv);
            vstore<R>(a.r + i * N, rv);
        }
        if (blockIdx.x == 0)
            for (int64_t i = nvec * N + threadIdx.x; i < n; i += blockDim.x) {
                a.x[i] = O::add(x_unset ? (R)0 : a.x[i], O::add(O::mul(alpha, a.y[i]), O::mul(w, a.z[i])));
                const R rn = O::sub(a.s[i], O::mul(w, a.t[i]));
                a.r[i] = rn;
                prr += rn * rn;
                prho += a.r0[i] * rn;
            }
        x_unset = false;
        R t2[2] = {prr, prho};
        block_sum_vec<R, 2, NW>(t2, red);
        BICG_STAMP(3)
        if (!gsync_allreduce<R, 2, NW>(cm, ar_seq, a.part + (size_t)2 * gridDim.x, t2, red, s_pk, s_ctl, gs_target)) {  // not the array of the sync before it (comm.cuh)
            failed = true;
            break;
        }
        BICG_STAMP(1)
        r_norm = O::sqrt(t2[0]);
        r0r = t2[1];
        ++it;
        if (!(t2[0] <= O::huge())) break;  // NaN / Inf: stop instead of spinning to maxit
    }
#undef BICG_STAMP
    if (x_unset && !failed) zero_solution<R>(a.x, n);
    if (prof) a.prof[9] += (unsigned long long)trips;
    // drain the prologue issued for a product that never runs
    {
        int k = 0;
        for (int tile = blockIdx.x; tile < num_tiles && k < STAGES; tile += gridDim.x, k++) {
            const unsigned f = fills + (unsigned)k;
            mbar_wait(&full[f % STAGES], (f / STAGES) & 1u);
        }
    }
    if (lead) {
        BicgState<R> c;
        c.rho[0] = c.rho[1] = rho;
        c.alpha = alpha;
        c.w = w;
        c.r0_norm = r0_norm;
        c.r_norm = r_norm;
        c.rhs_norm = rhs_norm;
        c.tol = a.tol;
        c.iters = it;
        c.trips = trips;
        c.restarts = restarts;
        c.restart_req = c.restart_now = c.skip = 0;
        c.done = 1;
        c.maxit = a.maxit;
        *a.st = c;
        if (failed) cm.loc->error = 1;
    }
}

}  // namespace feab
