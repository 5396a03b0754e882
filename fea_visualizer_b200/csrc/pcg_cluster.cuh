// Small meshes (the reference's interactive sizes: 10 k - 65 k rows): whole time steps -- right-hand side, the Jacobi-PCG
// solve(s) and the epilogue of CPUSolver::advance_time (src/FEM/CPUSolver.cpp:64-147), for any number of consecutive steps --
// in ONE launch of ONE thread-block cluster (up to 16 CTAs, one per SM).  At these sizes a PCG iteration moves well under a
// megabyte, so the time of the grid-wide persistent kernel (pcg_persistent.cuh) is three global-memory grid barriers and
// three kernel-wide sweeps per iteration (~12 us) plus three launches per step; here
//   * the dot products are all-gathered through distributed shared memory in ONE hop: every CTA sends its partial to every
//     CTA's slot with st.async (the store signals the receiver's mbarrier with complete_tx), every CTA waits on its own
//     mbarrier for the C x K x 8 bytes and adds the partials in CTA order -> bit-identical sums everywhere, uniform control
//     flow, no atomics, and no barrier.cluster round trip (arrive + wait) on top of the data; the one sync per iteration that
//     must order GLOBAL stores (the new direction, gathered by the neighbours) is a hardware cluster barrier (~0.5 us);
//   * a CTA owns a contiguous block of rows for the whole launch and keeps their r, A p, D^-1, x and p in shared memory,
//     together with its rows of the matrix as far as the 227 KB allow (row pointers, then columns, then the values of the system
//     being solved): a cluster barrier drops the L1, so whatever stays in global memory is re-read through L2 every
//     iteration.  Only p (gathered by the neighbours' SpMV) and the state u, v cross CTAs (plain loads after the barrier);
//   * the entries of a row are loaded in batches of 8 with all gathers of a batch in flight together (the loop is
//     latency-bound: three dependent L2 round trips per entry otherwise);
//   * nothing returns to the host between the steps of a device-resident call (feab_step(NULL, K, 0)).
// Same recurrence (ConjugateGradient.h:30-85: x0 = 0, threshold, rhs == 0 early-out, 2n cap), same operation order inside a
// row (two accumulators for A p, one for the right-hand side) as the other drivers; the partial sums are grouped by CTA, so
// results agree with them to summation order (<= 1e-12) and runs are bit-reproducible.  Gray-Scott's two solves run one
// after the other (the reference's order), each with its own values of the shared pattern -- or, when the device can keep two
// clusters resident (ClusterArgs::pair), SIDE BY SIDE: both right-hand sides are built from the old fields
// (CPUSolver.cpp:134-137) and the two solves are independent (:139-140), so cluster 0 takes the u system and cluster 1 the v
// system of every step.  The clusters meet twice per step through one global counter (after both have read the old fields,
// after both have stored the new ones); everything else, including the arithmetic of each solve, is unchanged.
#pragma once
#include <cooperative_groups.h>

#include <type_traits>

#include "comm.cuh"
#include "common.cuh"
#include "solver.cuh"
#include "spmv_tma.cuh"

namespace feab {

namespace cg = cooperative_groups;

constexpr int kClusterMaxCtas = 16;
constexpr int kClusterMaxRowsPerCta = 4096;          // 5 vectors of that many reals in shared memory: 160 KB in fp64
constexpr int kClusterSmemBudget = 227 * 1024 - 4096;  // dynamic shared memory the launch may ask for (the static part is ~2 KB)

template <typename R>
struct ClusterArgs {
    const int* rowptr;
    const int* col;
    const R* Md;       // M / dt
    const R* Kv;       // K (wave right-hand side)
    const R* A[2];
    const R* dinv[2];
    R* u;
    R* v;
    R* r[2];           // right-hand sides of the step (global: system 1's waits for solve 0 to finish)
    R* p[2];           // the direction vector the SpMV gathers from (pair mode: one per cluster)
    float* values;
    const int* node_of;    // row -> mesh node
    const int* dir_nodes;  // mesh nodes without a row (Dirichlet), zeroed on exit like map_vector_to_surface
    int64_t n_dir;
    int64_t n;
    int eq;
    StepScalars sc;
    R tol;
    int maxit;
    int nsteps;
    CgState<R>* st;    // [2]: state of the LAST step's solves (statistics)
    int* unstable;
    int stamp0;        // instability stamp of the first step of this launch
    int col_cap;       // entries per CTA whose column indices have room in shared memory (0: read them from global)
    int val_cap;       // the same for the values of the system being solved
    int pair;                  // Gray-Scott on two clusters: cluster s solves system s
    unsigned long long* xbar;  // pair mode: arrivals of cluster leaders at the inter-cluster barriers, monotone over the handle's life
    unsigned long long xbar_base;  // its value when this launch starts (kept by the host)
    CommLocal* loc;            // a barrier that never completed makes the later ones give up at once ...
    int* xerr;                 // ... and is reported to the host here (pinned host memory; FEAB_ECOMM)
    unsigned long long* prof;  // optional (FEAB_PCG_PROF=1): ns of CTA 0 in [SpMV, all-reduce 1, -, update, all-reduce 2, -, direction, barrier, -], iterations
};

// Vectors written by other CTAs of the cluster (p; u, v between steps) are read with PLAIN loads: every cluster barrier is a
// release / acquire at cluster scope and drops the L1, so a line read after it is never stale, and within a phase the L1 serves
// the neighbouring rows that share a 32-byte sector (with ld.cg every gathered value costs its own sector from L2: 4 x the
// bytes -- measured, 1.1 us per row pass against 0.3).  The pointers are volatile-free on purpose; nothing here is __ldg'd.
template <typename R>
__device__ __forceinline__ R ld_l2(const R* p) {
    return *p;
}

// shared-memory address `addr` of this CTA's window -> the same offset in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, int rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
// 8 bytes into another CTA's shared memory; their arrival is counted (complete_tx, release at cluster scope) on that CTA's mbarrier
__device__ __forceinline__ void st_async_f64(uint32_t remote_addr, double v, uint32_t remote_bar) {
    asm volatile("st.async.shared::cluster.mbarrier::complete_tx::bytes.b64 [%0], %1, [%2];" ::"r"(remote_addr), "l"(__double_as_longlong(v)),
                 "r"(remote_bar)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra DONE;\n"
        "bra LAB_WAIT;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

template <typename R, int THREADS>
__global__ void __launch_bounds__(THREADS, 1) pcg_cluster(ClusterArgs<R> a) {
    using O = Ops<R>;
    constexpr int kClusterThreads = THREADS;
    constexpr int NW = kClusterThreads / 32;
    cg::cluster_group cluster = cg::this_cluster();
    const int C = (int)cluster.num_blocks();
    const int c = (int)cluster.block_rank();
    const bool pair = a.pair != 0;
    const int myc = pair ? (int)blockIdx.x / C : 0;   // which cluster this CTA belongs to
    const int t = (int)threadIdx.x;
    const int n = (int)a.n;
    const int rpc = (n + C - 1) / C;
    const int lo = min(n, c * rpc), hi = min(n, lo + rpc);
    const int rows = hi - lo;

    extern __shared__ __align__(16) unsigned char smraw_cl[];
    R* __restrict__ s_r = reinterpret_cast<R*>(smraw_cl);   // own rows: residual
    R* __restrict__ s_y = s_r + rpc;                        // A p
    R* __restrict__ s_d = s_y + rpc;                        // D^-1
    R* __restrict__ s_x = s_d + rpc;                        // solution
    R* __restrict__ s_p = s_x + rpc;                        // own rows of the direction (the global copy is for the neighbours)
    R* s_val = s_p + rpc;                      // values of this CTA's rows (val_cap entries)
    int* s_col = reinterpret_cast<int*>(s_val + a.val_cap);   // their columns (col_cap entries)
    int* s_rp = s_col + a.col_cap;                            // row pointers relative to the CTA's first entry (rpc + 1)
    __shared__ double s_slot[2][kClusterMaxCtas][4];
    __shared__ __align__(8) uint64_t s_bar[2];   // arrival of the C x K partials of an all-reduce, one per slot set
    __shared__ R red[(NW + 1) * 4];
    int par = 0;
    unsigned phbits = 0;                         // bit q: the phase parity s_bar[q] completes next
    if (t == 0) {
        mbar_init(&s_bar[0], 1);
        mbar_init(&s_bar[1], 1);
        mbar_fence_init();
    }

    const int e0 = a.rowptr[lo];
    const int ne = a.rowptr[hi] - e0;          // entries of this CTA's rows
    const bool cols_sm = a.col_cap > 0 && ne <= a.col_cap;
    const bool vals_sm = a.val_cap > 0 && ne <= a.val_cap;
    for (int i = t; i <= rows; i += kClusterThreads) s_rp[i] = a.rowptr[lo + i] - e0;
    if (cols_sm)
        for (int k = t; k < ne; k += kClusterThreads) s_col[k] = a.col[e0 + k];
    const int* const colp = cols_sm ? s_col : a.col + e0;   // indexed by (entry - e0)
    cluster.sync();   // every CTA's mbarriers exist before anybody sends to them (also the CTA barrier for the copies above)

    // all-reduce of K sums over the cluster.  Every CTA's partial travels to every CTA (its own included) as st.async words
    // that complete a transaction count on the receiver's mbarrier; a CTA has the sums as soon as the C x K words have landed.
    // It is an execution barrier too (nobody gets past it before everybody has sent, i.e. has finished the phase before it),
    // but it does NOT order global memory: the phases around it exchange nothing through global memory (A p, r, x, D^-1 live
    // in the owner's shared memory; p is published before a real cluster barrier).  Slot sets and mbarriers alternate by
    // parity: a CTA can be at most one all-reduce ahead of the slowest one, because the next one needs everybody's word.
    auto allreduce = [&](auto& v, auto kc) {
        constexpr int K = decltype(kc)::value;
        // CTA sum in the order of block_sum_vec (warp shuffles, then warp 0 over the per-warp results); only warp 0 needs it
        const int lane = t & 31, w = t >> 5;
#pragma unroll
        for (int k = 0; k < K; k++) {
            const R ws = warp_sum(v[k]);
            if (lane == 0) red[w * K + k] = ws;
        }
        __syncthreads();
        if (w == 0) {
            uint32_t dst = 0, rbar = 0;
            if (lane < C) {
                dst = mapa_u32(smem_u32(&s_slot[par][c][0]), lane);
                rbar = mapa_u32(smem_u32(&s_bar[par]), lane);
            }
#pragma unroll
            for (int k = 0; k < K; k++) {
                R x = lane < NW ? red[lane * K + k] : (R)0;
                x = warp_sum(x);
                x = __shfl_sync(0xffffffffu, x, 0);
                if (lane < C) st_async_f64(dst + 8u * k, (double)x, rbar);
            }
            if (lane == 0) mbar_expect_tx(&s_bar[par], (uint32_t)(C * K * 8));
        }
        mbar_wait_cluster(&s_bar[par], (phbits >> par) & 1u);
        phbits ^= 1u << par;
#pragma unroll
        for (int k = 0; k < K; k++) {
            R acc = (R)0;
            for (int q = 0; q < C; q++) acc += (R)s_slot[par][q][k];
            v[k] = acc;
        }
        par ^= 1;
    };
    // pair mode: both clusters.  cluster.sync() orders the global stores of this cluster's CTAs before the leader's gpu-scope
    // fence + arrival; the leader's acquire and the second cluster.sync() (which drops every CTA's L1) order the other
    // cluster's stores before everything after it.  The spin is bounded; a timeout is reported, not survived.
    unsigned long long xcount = a.xbar_base;
    auto xsync = [&]() {
        xcount += 2;
        cluster.sync();
        if (c == 0 && t == 0) {
            __threadfence();
            atomicAdd(a.xbar, 1ull);
            if (ld_volatile_u64(a.xbar) < xcount) {
                const unsigned long long t0 = globaltimer_ns();
                while (ld_volatile_u64(a.xbar) < xcount) {
                    if (globaltimer_ns() - t0 > kSpinTimeoutNs || *(volatile int*)&a.loc->error) {
                        comm_fail(a.loc, 7, xcount, 1 - myc);
                        *(volatile int*)a.xerr = 1 + myc;
                        __threadfence_system();
                        break;
                    }
                }
            }
            __threadfence();
        }
        cluster.sync();
    };
    // sum_k val[k] * f(x[col[k]]) over the entries [kb, ke) of one row (relative to e0), ONE accumulator, ascending order,
    // separately rounded products (Eigen's column-major sparse * dense adds into y[row] in this order); batches of 8
    auto row_dot_seq = [&](R acc, int kb, int ke, const R* val /* indexed by entry - e0 */, const R* x, bool thru_float, R scale, bool scaled,
                           int self) -> R {
        for (int k0 = kb; k0 < ke; k0 += 8) {
            const int len = min(8, ke - k0);
            int cv[8];
            R mv[8], xv[8];
#pragma unroll
            for (int j = 0; j < 8; j++) {
                cv[j] = j < len ? colp[k0 + j] : self;
                mv[j] = j < len ? val[k0 + j] : (R)0;
            }
#pragma unroll
            for (int j = 0; j < 8; j++) xv[j] = ld_l2(x + cv[j]);
#pragma unroll
            for (int j = 0; j < 8; j++) {
                if (j < len) {
                    R xj = thru_float ? through_float(xv[j]) : xv[j];
                    R m = mv[j];
                    if (scaled) {  // wave: (cc K_ij) * (-1 u_j)
                        m = O::mul(scale, m);
                        xj = O::mul((R)-1, xj);
                    }
                    acc = O::add(acc, O::mul(m, xj));
                }
            }
        }
        return acc;
    };

    const bool prof = a.prof != nullptr && c == 0 && t == 0 && myc == 0;
    unsigned long long t_last = 0;
    if (prof) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_last));
#define CL_STAMP(slot_)                                                   \
    if (prof) {                                                           \
        unsigned long long t_now;                                         \
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_now));        \
        a.prof[slot_] += t_now - t_last;                                  \
        t_last = t_now;                                                   \
    }
    const int eq = a.eq;
    const int ns = eq == 3 ? 2 : 1;
    const R cc = O::mul((R)a.sc.c0, (R)a.sc.c0);
    const R dt = (R)a.sc.dt;
    const R* const Mdp = a.Md + e0;
    for (int step = 0; step < a.nsteps; step++) {
        // ---------------- right-hand side(s) + PCG start: r = b, |b|^2, b.D^-1 b  (solver.cuh rhs_init) ----------------
        //   eq 0 / 2: b = (M/dt) u;   eq 1: b = (M/dt) v - ((c c) K) u, the K.u terms accumulate onto the M.v sum;
        //   eq 3: b_u = ((M/dt) u - u v v) + f (1 - u), b_v = ((M/dt) v + u v v) - (f + k) v, both from the old fields
        R s4[4] = {(R)0, (R)0, (R)0, (R)0};  // |b_0|^2, b_0.p_0, |b_1|^2, b_1.p_1
        for (int i = t; i < rows; i += kClusterThreads) {
            const int r = lo + i;
            const int kb = s_rp[i], ke = s_rp[i + 1];
            if (eq == 3) {
                const R f = (R)a.sc.c2, fk = O::add((R)a.sc.c2, (R)a.sc.c3);
                const R ui = ld_l2(a.u + r), vi = through_float(ld_l2(a.v + r));
                const R uvv = O::mul(ui, O::mul(vi, vi));
                if (!pair || myc == 0) {
                    const R s0 = row_dot_seq((R)0, kb, ke, Mdp, a.u, false, (R)0, false, r);
                    const R bu = O::add(O::sub(s0, uvv), O::mul(f, O::sub((R)1, ui)));
                    a.r[0][r] = bu;
                    const R pu = O::mul(a.dinv[0][r], bu);
                    s4[0] += bu * bu;
                    s4[1] += bu * pu;
                }
                if (!pair || myc == 1) {
                    const R s1 = row_dot_seq((R)0, kb, ke, Mdp, a.v, true, (R)0, false, r);
                    const R bv = O::sub(O::add(s1, uvv), O::mul(fk, vi));
                    a.r[1][r] = bv;
                    const R pv = O::mul(a.dinv[1][r], bv);
                    s4[2] += bv * bv;
                    s4[3] += bv * pv;
                }
            } else {
                R s0;
                if (eq == 1) {
                    s0 = row_dot_seq((R)0, kb, ke, Mdp, a.v, false, (R)0, false, r);
                    s0 = row_dot_seq(s0, kb, ke, a.Kv + e0, a.u, true, cc, true, r);
                } else {
                    s0 = row_dot_seq((R)0, kb, ke, Mdp, a.u, true, (R)0, false, r);
                }
                a.r[0][r] = s0;
                const R pz = O::mul(a.dinv[0][r], s0);
                s4[0] += s0 * s0;
                s4[1] += s0 * pz;
            }
        }
        allreduce(s4, std::integral_constant<int, 4>());  // every CTA has read the old u, v after this barrier
        if (pair) xsync();                                // ... and so has every CTA of the other cluster

        bool bad = false;
        for (int s = 0; s < ns; s++) {
            if (pair && s != myc) continue;
            R* const pg = a.p[pair ? s : 0];
            R* const xg = eq == 1 ? a.v : (s == 0 ? a.u : a.v);
            const R bb = s4[2 * s];
            R abs_old = s4[2 * s + 1];
            R thr = a.tol * a.tol * bb;
            if (thr < O::tiny()) thr = O::tiny();
            int it = 0, iters = 0;
            R rr = bb;
            const bool skip = bb == (R)0 || bb < thr || a.maxit <= 0;
            // own rows into shared memory: r = b, D^-1, x = 0, p = D^-1 b (also published for the neighbours' gathers)
            for (int i = t; i < rows; i += kClusterThreads) {
                const int r = lo + i;
                const R b = a.r[s][r], d = a.dinv[s][r];
                const R pz = O::mul(d, b);
                s_r[i] = b;
                s_d[i] = d;
                s_x[i] = (R)0;
                s_p[i] = pz;
                pg[r] = pz;
            }
            if (vals_sm && !skip)
                for (int k = t; k < ne; k += kClusterThreads) s_val[k] = a.A[s][e0 + k];
            const R* const valp = vals_sm ? s_val : a.A[s] + e0;
            cluster.sync();  // p complete (and the shared copies, CTA-wide)
            CL_STAMP(8)
            while (!skip) {
                // ---- y = A p, p.y   (SparseDenseProduct.h:66-88 two-accumulator order, ConjugateGradient.h:66) ----
                R pap[1] = {(R)0};
                for (int i = t; i < rows; i += kClusterThreads) {
                    const int kb = s_rp[i], ke = s_rp[i + 1];
                    R ta = (R)0, tb = (R)0;
                    for (int k0 = kb; k0 < ke; k0 += 8) {   // k0 - kb stays even: the even / odd split is the row's own
                        const int len = min(8, ke - k0);
                        int cv[8];
                        R av[8], xv[8];
#pragma unroll
                        for (int j = 0; j < 8; j++) {
                            cv[j] = j < len ? colp[k0 + j] : lo + i;
                            av[j] = j < len ? valp[k0 + j] : (R)0;
                        }
#pragma unroll
                        for (int j = 0; j < 8; j++) xv[j] = ld_l2(pg + cv[j]);
#pragma unroll
                        for (int j = 0; j < 8; j++) {
                            if (j < len) {
                                if (j & 1)
                                    tb += av[j] * xv[j];
                                else
                                    ta += av[j] * xv[j];
                            }
                        }
                    }
                    const R sum = ta + tb;
                    s_y[i] = sum;
                    pap[0] += s_p[i] * sum;
                }
                it++;
                CL_STAMP(0)
                allreduce(pap, std::integral_constant<int, 1>());
                CL_STAMP(1)
                // ---- x += alpha p; r -= alpha y; |r|^2, r.z   (:67-76) ----
                const R alpha = abs_old / pap[0];
                R t2[2] = {(R)0, (R)0};
#pragma unroll 4
                for (int i = t; i < rows; i += kClusterThreads) {
                    s_x[i] = O::add(s_x[i], O::mul(alpha, s_p[i]));
                    const R rn = O::sub(s_r[i], O::mul(alpha, s_y[i]));
                    s_r[i] = rn;
                    const R z = O::mul(s_d[i], rn);
                    t2[0] += rn * rn;
                    t2[1] += rn * z;
                }
                CL_STAMP(3)
                allreduce(t2, std::integral_constant<int, 2>());
                CL_STAMP(4)
                if (prof) a.prof[9] += 1;
                rr = t2[0];
                const bool conv = rr < thr;
                if (conv || it >= a.maxit || !(rr <= O::huge())) {
                    iters = conv ? it - 1 : it;
                    if (!conv) abs_old = t2[1];
                    break;
                }
                // ---- p = D^-1 r + beta p   (:77-80) ----
                const R beta = t2[1] / abs_old;
                abs_old = t2[1];
#pragma unroll 4
                for (int i = t; i < rows; i += kClusterThreads) {
                    const R pn = O::add(O::mul(s_d[i], s_r[i]), O::mul(beta, s_p[i]));
                    s_p[i] = pn;
                    pg[lo + i] = pn;
                }
                CL_STAMP(6)
                cluster.sync();  // p complete
                CL_STAMP(7)
            }
            // ---- solution -> state vector ----
            for (int i = t; i < rows; i += kClusterThreads) xg[lo + i] = s_x[i];
            if (c == 0 && t == 0 && step == a.nsteps - 1) {
                CgState<R> cs;
                cs.abs_new[0] = cs.abs_new[1] = abs_old;
                cs.thr = thr;
                cs.rhs2 = bb;
                cs.rn2 = bb == (R)0 ? (R)0 : rr;
                cs.tol = a.tol;
                cs.iter1 = it;
                cs.done = 1;
                cs.iters = iters;
                cs.maxit = a.maxit;
                a.st[s] = cs;
            }
            __syncthreads();  // the shared vectors are reused by the next solve
        }
        // ---------------- epilogue: wave u += v dt, values <- solution, instability check  (solver.cuh finish_step) ----------------
        if (pair) {  // each cluster looks at the field it solved for (written by the same threads a moment ago); values == v
            for (int i = t; i < rows; i += kClusterThreads) {
                const int r = lo + i;
                const R xi = myc == 0 ? a.u[r] : a.v[r];
                if (myc == 1) a.values[a.node_of[r]] = (float)xi;
                bad = bad || xi > (R)1e4f;
            }
        } else {
            for (int i = t; i < rows; i += kClusterThreads) {
                const int r = lo + i;
                R ui = a.u[r], vi = a.v[r];
                if (eq == 1) {
                    ui = O::add(through_float(ui), O::mul(vi, dt));
                    a.u[r] = ui;
                }
                a.values[a.node_of[r]] = (float)(eq == 3 ? vi : ui);
                bad = bad || ui > (R)1e4f || vi > (R)1e4f;
            }
        }
        if (myc == 0)
            for (int64_t i = (int64_t)c * kClusterThreads + t; i < a.n_dir; i += (int64_t)C * kClusterThreads) a.values[a.dir_nodes[i]] = 0.0f;
        if (__any_sync(0xffffffffu, bad) && (t & 31) == 0) atomicMax(a.unstable, a.stamp0 + step);
        if (pair)
            xsync();         // u and v complete, in both clusters, before the next step's right-hand sides gather them
        else
            cluster.sync();  // u, v complete before the next step's right-hand side gathers them
        CL_STAMP(8)
    }
#undef CL_STAMP
}

}  // namespace feab
