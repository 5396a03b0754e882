// Multi-GPU plumbing of the partitioned solver: scalar all-reduces and halo pushes done by the
// solver's own kernels through peer memory over NVLink / NVSwitch -- no host round trip, no separate collective
// launch.  Every rank (one process per GPU, or one "virtual rank" per handle in the single-process test driver)
// owns one arena allocation that starts with a CommCtrl block; peers map the arena (cudaIpc or same-process
// pointers) and write into it:
//   * all-reduce: the last CTA of a producer kernel sums the kernel's per-CTA partials in a fixed order and
//     stores the rank total + a sequence number into slot[seq & 1][rank] of EVERY rank's CommCtrl; the consumer
//     kernel spins on its own (local) slots and adds the totals in rank order -> every rank obtains bit-identical
//     sums, so all ranks take identical convergence decisions without talking to the host.
//   * halo: a push kernel stores the boundary rows of a vector straight into the ghost tail of each neighbour's
//     vector, then raises that neighbour's flag; the neighbour's SpMV waits for the flag only before its first
//     boundary tile, so the interior rows overlap the transfer.
// Double buffering by sequence parity is sufficient: a rank can publish all-reduce s+2 only after it consumed
// s+1, which needs every peer's s+1, which each peer publishes after consuming s.
// All spins are bounded (kSpinTimeoutNs) and FAIL CLOSED: on timeout the waiting kernel records CommLocal::error, stops
// the solve (every CTA of a persistent kernel takes the same exit at the same grid sync), later kernels of the step
// return at once without touching values / u / v or pushing anything to a peer, and the host raises FEAB_ECOMM.
#pragma once
#include "common.cuh"

namespace feab {

constexpr int kMaxRanks = 8;
constexpr unsigned long long kSpinTimeoutNs = 10ull * 1000 * 1000 * 1000;

// One rank's contribution to one all-reduce, in "LL" form: every double travels as two 8-byte words, each carrying 32
// data bits and the low 32 bits of the sequence number.  An 8-byte store is single-copy atomic, so a word whose flag
// matches is complete by itself -- no fence between data and flag, one NVLink hop per all-reduce.
struct ArSlot {
    unsigned long long w[8];  // [2 k], [2 k + 1]: low / high half of value k, k < 4
};
struct CommCtrl {  // written by peers
    ArSlot slot[2][kMaxRanks];
    unsigned long long halo_flag[kMaxRanks];   // [source rank] number of direction-halo pushes received
    unsigned long long state_flag[kMaxRanks];  // [source rank] number of end-of-step state pushes received
    unsigned long long full_flag[kMaxRanks];   // [source rank] number of full value broadcasts received
    unsigned long long root_flag;              // number of root pushes (rank 0's full surface->values) received
    unsigned long long pad[7];
};
struct CommLocal {  // private to the rank
    unsigned long long ar_seq;
    unsigned long long halo_seq, state_seq, full_seq;
    unsigned long long root_seq;
    unsigned int ticket;
    int error;
    // grid-wide syncs of the persistent kernels (gsync_* below): arrivals at the current sync (atomicInc wraps it back to 0
    // with the last arrival, so it never needs a reset) and the number of barrier-only syncs released so far
    unsigned int gs_arrive;
    int err_kind;                  // which wait gave up first: 1 halo flag, 2 state flag, 3 full-values flag, 4 root flag,
                                   // 5 mailbox word of an all-reduce, 6 grid sync of the own rank
    unsigned long long gs_release;
    unsigned long long err_want;   // the flag value / sequence number that never arrived
    int err_peer;                  // the rank it was expected from
    int err_pad;
    // one GPU (world == 1): the grid syncs count arrivals on a MONOTONE counter every CTA spins on directly (gsync_* below);
    // gs_base = its value at the last completed sync, written by that sync's last arrival, read once at kernel start
    unsigned long long gs_count, gs_base;
};
__device__ __forceinline__ void comm_fail(CommLocal* loc, int kind, unsigned long long want, int peer) {
    if (atomicCAS(&loc->error, 0, 1) == 0) {
        loc->err_kind = kind;
        loc->err_want = want;
        loc->err_peer = peer;
    }
}
constexpr size_t kCtrlBytes = (sizeof(CommCtrl) + 255) / 256 * 256;

// Passed by value to kernels.  world == 1: nothing below is touched.
struct Comm {
    int rank = 0, world = 1;
    CommCtrl* ctrl[kMaxRanks] = {};  // ctrl[rank] is local
    CommLocal* loc = nullptr;
};

// Which neighbours receive which of my rows, and where they land in the neighbour's vectors.
struct HaloPlan {
    int nnb = 0;
    int nb_rank[kMaxRanks] = {};
    int send_off[kMaxRanks + 1] = {};   // ranges in send_rows / send_nodes
    int dst_col[kMaxRanks] = {};        // first ghost column of my block on that neighbour
    const int* send_rows = nullptr;     // my local rows, ascending global unknown id per neighbour
    const int* send_nodes = nullptr;    // their mesh nodes
    // inverse of the send lists, for pushes fused into the kernel that produces the values: for boundary row
    // r >= n_int, push_tab[(r - n_int) * nnb + b] = ghost column on neighbour b that mirrors it, or -1
    const int* push_tab = nullptr;
    int n_int = 0;
    char* base[kMaxRanks] = {};         // neighbour arena base, indexed by RANK
    unsigned long long off_vec[5][kMaxRanks] = {};  // byte offsets of p[0], p[1], u, v, values in that arena, by RANK
};
enum { kVecP0 = 0, kVecP1 = 1, kVecU = 2, kVecV = 3, kVecValues = 4 };

__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// push_tab from the send lists (one thread per send entry; memset to -1 beforehand)
__global__ void build_push_table(HaloPlan hp, int* __restrict__ tab) {
    const int total = hp.send_off[hp.nnb];
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
        int b = 0;
        while (e >= hp.send_off[b + 1]) b++;
        tab[(size_t)(hp.send_rows[e] - hp.n_int) * hp.nnb + b] = hp.dst_col[b] + (e - hp.send_off[b]);
    }
}

// Spin until *flag >= want (one thread).  Returns false on timeout.
__device__ __forceinline__ bool spin_until_ge(const unsigned long long* flag, unsigned long long want, CommLocal* loc, int kind = 0,
                                              int peer = -1) {
    if (ld_acquire_sys(flag) >= want) return true;
    const unsigned long long t0 = globaltimer_ns();
    while (ld_acquire_sys(flag) < want) {
        __nanosleep(64);
        if (globaltimer_ns() - t0 > kSpinTimeoutNs || *(volatile int*)&loc->error) {  // a wait that already failed fails the rest at once
            comm_fail(loc, kind, want, peer);
            return false;
        }
    }
    return true;
}
__device__ __forceinline__ bool comm_failed(const Comm& cm) { return *(volatile int*)&cm.loc->error != 0; }

// LL mailbox: store value k of all-reduce `seq` into a (possibly remote) slot / wait for it in a local slot.
__device__ __forceinline__ void ll_store(ArSlot* sl, int k, double v, unsigned long long seq) {
    const unsigned long long bits = (unsigned long long)__double_as_longlong(v);
    const unsigned long long flag = (seq & 0xffffffffull) << 32;
    st_volatile_u64(&sl->w[2 * k], flag | (bits & 0xffffffffull));
    st_volatile_u64(&sl->w[2 * k + 1], flag | (bits >> 32));
}
// *ok is cleared (and CommLocal::error set) when the word never arrived; the caller must not use the value then.
__device__ __forceinline__ double ll_wait(const ArSlot* sl, int k, unsigned long long seq, CommLocal* loc, bool* ok = nullptr, int peer = -1) {
    const unsigned long long flag = seq & 0xffffffffull;
    unsigned long long lo = ld_volatile_u64(&sl->w[2 * k]), hi = ld_volatile_u64(&sl->w[2 * k + 1]);
    if ((lo >> 32) != flag || (hi >> 32) != flag) {
        const unsigned long long t0 = globaltimer_ns();
        while (true) {
            lo = ld_volatile_u64(&sl->w[2 * k]);
            hi = ld_volatile_u64(&sl->w[2 * k + 1]);
            if ((lo >> 32) == flag && (hi >> 32) == flag) break;
            if (globaltimer_ns() - t0 > kSpinTimeoutNs || *(volatile int*)&loc->error) {
                comm_fail(loc, 5, seq, peer);
                if (ok) *ok = false;
                break;
            }
        }
    }
    return __longlong_as_double((long long)((hi << 32) | (lo & 0xffffffffull)));
}

// Fixed-order sum of per-CTA partials written earlier IN THIS KERNEL by other CTAs (L2 loads, never the
// non-coherent path).
template <typename T, int NW>
__device__ __forceinline__ T sum_partials_cg(const T* part, int count, T* sm) {
    T s = T(0);
    for (int i = threadIdx.x; i < count; i += blockDim.x) s += __ldcg(part + i);
    T r = block_sum_n<T, NW>(s, sm);
    __syncthreads();
    return r;
}

// Producer side.  Call from ALL threads of EVERY CTA after thread 0 stored part[k * stride + blockIdx.x], k < K.
// The last CTA to arrive publishes the rank totals to every rank (its own mailbox included, so the consumer is the same
// code on 1..8 ranks).  raise_halo: the kernel also stored boundary rows into the neighbours' ghost tails (every CTA that
// did ran a system-scope fence before calling this); the last CTA then raises the neighbours' halo flags.
template <typename R, int K, int NW = kBlock / 32>
__device__ __forceinline__ void ar_publish(const Comm& cm, const R* part, int stride, R* sm, const HaloPlan* hp = nullptr, bool raise_halo = false) {
    __shared__ int s_last;
    if (threadIdx.x == 0) {
        __threadfence();
        unsigned t = atomicAdd(&cm.loc->ticket, 1u);
        s_last = (t == gridDim.x - 1) ? 1 : 0;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    R tot[K];
#pragma unroll
    for (int k = 0; k < K; k++) tot[k] = sum_partials_cg<R, NW>(part + (size_t)k * stride, (int)gridDim.x, sm);
    if (threadIdx.x == 0) {
        cm.loc->ticket = 0;
        const unsigned long long s = cm.loc->ar_seq + 1;
        cm.loc->ar_seq = s;
        for (int q = 0; q < cm.world; q++) {
            ArSlot* sl = &cm.ctrl[q]->slot[s & 1][cm.rank];
#pragma unroll
            for (int k = 0; k < K; k++) ll_store(sl, k, (double)tot[k], s);
        }
        if (raise_halo && hp != nullptr) {
            const unsigned long long hs = cm.loc->halo_seq + 1;
            cm.loc->halo_seq = hs;
            for (int b = 0; b < hp->nnb; b++) st_release_sys(&cm.ctrl[hp->nb_rank[b]]->halo_flag[cm.rank], hs);
        }
    }
}

// Boundary row `row` (>= hp.n_int) of arena vector `vec`: mirror `val` on every neighbour that keeps the row as a ghost.
template <typename R>
__device__ __forceinline__ void push_boundary_row(const HaloPlan& hp, int vec, int64_t row, R val) {
    const int* t = hp.push_tab + (size_t)(row - hp.n_int) * hp.nnb;
    for (int b = 0; b < hp.nnb; b++) {
        const int d = t[b];
        if (d >= 0) reinterpret_cast<R*>(hp.base[hp.nb_rank[b]] + hp.off_vec[vec][hp.nb_rank[b]])[d] = val;
    }
}

// Consumer side (the kernel launched after the producer on the same stream).  Every CTA calls it; result in all threads.
template <typename R, int K>
__device__ __forceinline__ void ar_consume(const Comm& cm, R (&out)[K]) {
    __shared__ double s_v[4];
    if (threadIdx.x == 0) {
        const unsigned long long s = *(volatile unsigned long long*)&cm.loc->ar_seq;
        const CommCtrl* me = cm.ctrl[cm.rank];
        R acc[K];
#pragma unroll
        for (int k = 0; k < K; k++) acc[k] = (R)0;
        for (int q = 0; q < cm.world; q++) {
            const ArSlot* sl = &me->slot[s & 1][q];
#pragma unroll
            for (int k = 0; k < K; k++) acc[k] += (R)ll_wait(sl, k, s, cm.loc);
        }
#pragma unroll
        for (int k = 0; k < K; k++) s_v[k] = (double)acc[k];
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < K; k++) out[k] = (R)s_v[k];
    __syncthreads();
}

// One CTA right after a producer: stores the all-reduced totals where single-CTA kernels (cg_begin) expect partials.
template <typename R, int K>
__global__ void ar_collect(Comm cm, R* __restrict__ totals) {
    R t[K];
    ar_consume<R, K>(cm, t);
    if (threadIdx.x == 0)
#pragma unroll
        for (int k = 0; k < K; k++) totals[k] = t[k];
}

// K sums of gridDim.x per-CTA partials written before a grid sync (partials of sum k at part + k * stride), added in
// CTA order: one pass over L2, one block reduction.  Result in every thread of the calling CTA.
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

// ---- grid-wide syncs of the persistent (cooperative) kernels, merged with the all-reduce that rides on them ----------
// gsync_allreduce: every CTA of every rank calls it with its K partial sums.  Thread 0 stores the CTA's partials, fences
// and takes a ticket; the LAST CTA of the rank to arrive adds all partials in CTA order and stores the rank totals, as LL
// words, into EVERY rank's mailbox (its own included: one code path for 1..8 ranks); every CTA then waits -- directly on
// the mailbox words, there is no separate arrival counter to spin on -- for the totals of all ranks and adds them in rank
// order.  Bit-identical results in every CTA of every rank, so all of them take the same decisions.  One wait where the
// first version had three (barrier -> every CTA re-sums all partials -> mailbox exchange).  The sync also orders memory:
// what any CTA of the rank wrote before it is visible to every CTA after it (fence / ticket / LL word / fence).
// Returns false when a peer never answered (CTA-uniform, and the same in every CTA of the rank: all of them sit in the
// same sync when a peer dies).  s_ctl[1] must be zero when the kernel starts.
// One GPU: arrivals are counted on a monotone counter (CommLocal::gs_count) that every CTA spins on directly, so a sync is
// released by the last ARRIVAL itself, not by a store the last CTA makes after it noticed being last -- one L2 round trip
// less on the critical path.  `target` = the counter value that completes this sync (kernel-local running value, started
// from CommLocal::gs_base).  Returns false on timeout.
__device__ __forceinline__ bool gsync_count_arrive_and_wait(const Comm& cm, unsigned long long target) {
    const unsigned long long old = atomicAdd(&cm.loc->gs_count, 1ull);
    if (old + 1 == target) st_volatile_u64(&cm.loc->gs_base, target);  // only the next kernel reads it
    bool ok = true;
    if (old + 1 < target && ld_volatile_u64(&cm.loc->gs_count) < target) {
        const unsigned long long t0 = globaltimer_ns();
        while (ld_volatile_u64(&cm.loc->gs_count) < target) {
            if (globaltimer_ns() - t0 > kSpinTimeoutNs || *(volatile int*)&cm.loc->error) {
                comm_fail(cm.loc, 6, target, cm.rank);
                ok = false;
                break;
            }
        }
    }
    return ok;
}

template <typename R, int K, int NW>
__device__ __forceinline__ bool gsync_allreduce(const Comm& cm, unsigned long long& seq, R* __restrict__ part /* K x gridDim.x */,
                                                R (&v)[K] /* in: this CTA's partials (thread 0); out: the global sums */,
                                                R* red /* (NW + 1) * K */, double* s_pk /* kMaxRanks * 4 */, int* s_ctl /* 2 */,
                                                unsigned long long& gtarget /* world == 1: see gsync_count_arrive_and_wait */) {
    static_assert(K >= 1 && K <= 4, "an ArSlot carries four values");
    const unsigned G = gridDim.x;
    seq += 1;
    if (cm.world == 1) {
        // Every CTA adds all partials itself, in the order the last CTA of the multi-rank path uses (same bits).  `part` must not
        // be the array of the PREVIOUS sync of the kernel: a fast CTA writes its next partial while a slow one still reads.
        gtarget += G;
        __syncthreads();
        if (threadIdx.x == 0) {
#pragma unroll
            for (int k = 0; k < K; k++) part[(size_t)k * G + blockIdx.x] = v[k];
            __threadfence();
            if (!gsync_count_arrive_and_wait(cm, gtarget)) s_ctl[1] = 1;
            __threadfence();
        }
        __syncthreads();
        sum_cta_partials_vec<R, K, NW>(part, (size_t)G, v, red);
        const bool failed1 = s_ctl[1] != 0;
        __syncthreads();
        return !failed1;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int k = 0; k < K; k++) part[(size_t)k * G + blockIdx.x] = v[k];
        __threadfence();
        s_ctl[0] = atomicInc(&cm.loc->gs_arrive, G - 1u) == G - 1u ? 1 : 0;  // wraps to 0 with the last arrival
    }
    __syncthreads();
    if (s_ctl[0]) {  // the last CTA of this rank: every other CTA's partials (and phase stores) are visible after the fence
        __threadfence();
        R tot[K];
        sum_cta_partials_vec<R, K, NW>(part, (size_t)G, tot, red);
        if ((int)threadIdx.x < cm.world) {
            ArSlot* sl = &cm.ctrl[threadIdx.x]->slot[seq & 1][cm.rank];
#pragma unroll
            for (int k = 0; k < K; k++) ll_store(sl, k, (double)tot[k], seq);
        }
        if (threadIdx.x == 0) cm.loc->ar_seq = seq;
    }
    if ((int)threadIdx.x < cm.world) {
        const ArSlot* mine = &cm.ctrl[cm.rank]->slot[seq & 1][threadIdx.x];
        bool ok = true;
#pragma unroll
        for (int k = 0; k < K; k++) s_pk[threadIdx.x * 4 + k] = ll_wait(mine, k, seq, cm.loc, &ok, (int)threadIdx.x);
        if (!ok) s_ctl[1] = 1;
        __threadfence();
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < K; k++) {
        R acc = (R)0;
        for (int w = 0; w < cm.world; w++) acc += (R)s_pk[w * 4 + k];
        v[k] = acc;
    }
    const bool failed = s_ctl[1] != 0;
    __syncthreads();
    return !failed;
}

// Barrier-only sync of the rank's grid (no value travels).  halo_seq != 0: the phase before it pushed boundary rows into
// the neighbours' ghost tails (each pushing CTA ran a system-scope fence before arriving); the last CTA to arrive raises
// the neighbours' halo flags with release stores, one thread per neighbour, AFTER letting its own rank go on.
__device__ __forceinline__ bool gsync_barrier(const Comm& cm, const HaloPlan& hp, unsigned long long& bseq, unsigned long long halo_seq,
                                              int* s_ctl /* 2 */, unsigned long long& gtarget) {
    const unsigned G = gridDim.x;
    bseq += 1;
    if (cm.world == 1) {
        gtarget += G;
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            if (!gsync_count_arrive_and_wait(cm, gtarget)) s_ctl[1] = 1;
            __threadfence();
        }
        __syncthreads();
        const bool failed1 = s_ctl[1] != 0;
        __syncthreads();
        return !failed1;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        s_ctl[0] = atomicInc(&cm.loc->gs_arrive, G - 1u) == G - 1u ? 1 : 0;
    }
    __syncthreads();
    if (s_ctl[0]) {
        if (threadIdx.x == 0) {
            __threadfence();
            st_volatile_u64(&cm.loc->gs_release, bseq);
            if (halo_seq) cm.loc->halo_seq = halo_seq;
        }
        if (halo_seq && (int)threadIdx.x < hp.nnb) st_release_sys(&cm.ctrl[hp.nb_rank[threadIdx.x]]->halo_flag[cm.rank], halo_seq);
    } else if (threadIdx.x == 0) {
        if (ld_volatile_u64(&cm.loc->gs_release) < bseq) {
            const unsigned long long t0 = globaltimer_ns();
            while (ld_volatile_u64(&cm.loc->gs_release) < bseq) {
                if (globaltimer_ns() - t0 > kSpinTimeoutNs || *(volatile int*)&cm.loc->error) {
                    comm_fail(cm.loc, 6, bseq, cm.rank);
                    s_ctl[1] = 1;
                    break;
                }
            }
        }
        __threadfence();
    }
    __syncthreads();
    const bool failed = s_ctl[1] != 0;
    __syncthreads();
    return !failed;
}

// Wait (one thread) until every neighbour's push number `want` has landed.  kind: 0 halo, 1 state, 2 full.
__device__ __forceinline__ bool wait_flags(const Comm& cm, const HaloPlan& hp, int kind, unsigned long long want) {
    const CommCtrl* me = cm.ctrl[cm.rank];
    bool ok = true;
    if (kind == 2) {
        for (int q = 0; q < cm.world; q++)
            if (q != cm.rank) ok = spin_until_ge(&me->full_flag[q], want, cm.loc, 3, q) && ok;
        return ok;
    }
    for (int i = 0; i < hp.nnb; i++) {
        const int q = hp.nb_rank[i];
        ok = spin_until_ge(kind == 0 ? &me->halo_flag[q] : &me->state_flag[q], want, cm.loc, kind == 0 ? 1 : 2, q) && ok;
    }
    return ok;
}
__global__ void wait_kernel(Comm cm, HaloPlan hp, int kind) {
    if (threadIdx.x == 0) {
        const unsigned long long want = kind == 0 ? cm.loc->halo_seq : kind == 1 ? cm.loc->state_seq : cm.loc->full_seq;
        wait_flags(cm, hp, kind, want);
    }
}

// Push the boundary rows of up to two vectors (p of solve 0 / 1) into the neighbours' ghost tails, then raise
// their halo flags.  One CTA: cuts of a surface mesh are O(sqrt(n)) rows.
template <typename R>
__global__ void __launch_bounds__(1024) halo_push(Comm cm, HaloPlan hp, const R* __restrict__ src0, int vec0, const R* __restrict__ src1,
                                                  int vec1) {
    for (int b = 0; b < hp.nnb; b++) {
        const int q = hp.nb_rank[b];
        const int beg = hp.send_off[b], cnt = hp.send_off[b + 1] - beg;
        R* d0 = reinterpret_cast<R*>(hp.base[q] + hp.off_vec[vec0][q]) + hp.dst_col[b];
        R* d1 = src1 ? reinterpret_cast<R*>(hp.base[q] + hp.off_vec[vec1][q]) + hp.dst_col[b] : nullptr;
        for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
            const int row = hp.send_rows[beg + i];
            d0[i] = src0[row];
            if (src1) d1[i] = src1[row];
        }
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned long long s = cm.loc->halo_seq + 1;
        cm.loc->halo_seq = s;
        for (int b = 0; b < hp.nnb; b++) st_release_sys(&cm.ctrl[hp.nb_rank[b]]->halo_flag[cm.rank], s);
    }
}

// End of a time step: boundary rows of u and v (hidden solver state) and their surface values go to the neighbours.
template <typename R>
__global__ void __launch_bounds__(1024) state_push(Comm cm, HaloPlan hp, const R* __restrict__ u, const R* __restrict__ v,
                                                   const float* __restrict__ values) {
    for (int b = 0; b < hp.nnb; b++) {
        const int q = hp.nb_rank[b];
        const int beg = hp.send_off[b], cnt = hp.send_off[b + 1] - beg;
        R* du = reinterpret_cast<R*>(hp.base[q] + hp.off_vec[kVecU][q]) + hp.dst_col[b];
        R* dv = reinterpret_cast<R*>(hp.base[q] + hp.off_vec[kVecV][q]) + hp.dst_col[b];
        float* dval = reinterpret_cast<float*>(hp.base[q] + hp.off_vec[kVecValues][q]);
        for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
            const int row = hp.send_rows[beg + i];
            const int node = hp.send_nodes[beg + i];
            du[i] = u[row];
            dv[i] = v[row];
            dval[node] = values[node];
        }
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned long long s = cm.loc->state_seq + 1;
        cm.loc->state_seq = s;
        for (int b = 0; b < hp.nnb; b++) st_release_sys(&cm.ctrl[hp.nb_rank[b]]->state_flag[cm.rank], s);
    }
}

// surface->values is complete on every rank after advance_time() (the reference's data contract): each rank
// broadcasts the values of the nodes it owns to all peers.  Grid-wide; the last CTA raises the flags.
// root_only (FEAB_SYNC_ROOT): only rank 0's host array is read afterwards, so the owners store into rank 0's mirror only.
__global__ void __launch_bounds__(kBlock) values_broadcast(Comm cm, HaloPlan hp, const int* __restrict__ row_node, int64_t n_own,
                                                           const float* __restrict__ values, int root_only) {
    if (!(root_only && cm.rank == 0)) {
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_own; i += (int64_t)gridDim.x * blockDim.x) {
            const int node = row_node[i];
            const float val = values[node];
            if (root_only) {
                reinterpret_cast<float*>(hp.base[0] + hp.off_vec[kVecValues][0])[node] = val;
            } else {
                for (int q = 0; q < cm.world; q++)
                    if (q != cm.rank) reinterpret_cast<float*>(hp.base[q] + hp.off_vec[kVecValues][q])[node] = val;
            }
        }
    }
    __threadfence_system();
    __shared__ int s_last;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned t = atomicAdd(&cm.loc->ticket, 1u);
        s_last = (t == gridDim.x - 1) ? 1 : 0;
    }
    __syncthreads();
    if (s_last && threadIdx.x == 0) {
        cm.loc->ticket = 0;
        __threadfence_system();
        const unsigned long long s = cm.loc->full_seq + 1;
        cm.loc->full_seq = s;
        for (int q = 0; q < cm.world; q++)
            if (q != cm.rank && (!root_only || q == 0)) st_release_sys(&cm.ctrl[q]->full_flag[cm.rank], s);
    }
}

// FEAB_SYNC_ROOT (feab.h): surface->values lives on the host of rank 0 only (the viewer process).  Rank 0 uploads it into its
// own mirror and raises every peer's root flag (this kernel, stream-ordered after the H2D copy); each peer waits for the flag
// and then PULLS the values of its own columns (owned rows + ghosts) from rank 0's mirror over NVLink, inside its gather
// kernel (solver.cuh gather_cols) -- V floats leave rank 0 in total, not V floats per peer.
__global__ void root_signal(Comm cm) {
    if (threadIdx.x == 0) {
        __threadfence_system();
        const unsigned long long s = cm.loc->root_seq + 1;
        cm.loc->root_seq = s;
        for (int q = 0; q < cm.world; q++)
            if (q != cm.rank) st_release_sys(&cm.ctrl[q]->root_flag, s);
    }
}
__global__ void root_wait(Comm cm) {
    if (threadIdx.x == 0) {
        const unsigned long long s = cm.loc->root_seq + 1;
        cm.loc->root_seq = s;
        spin_until_ge(&cm.ctrl[cm.rank]->root_flag, s, cm.loc, 4, 0);
    }
}

}  // namespace feab
