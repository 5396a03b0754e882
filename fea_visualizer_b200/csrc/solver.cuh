// Per-timestep kernels: operator build, fused RHS + PCG initialisation, the three PCG iteration kernels
// (SpMV fused with p.Ap, vector update fused with |r|^2 and r.z, direction update fused with the
// device-side convergence test) and the step epilogue.  Restates src/FEM/CPUSolver.cpp:64-147 and
// lib/eigen/Eigen/src/IterativeLinearSolvers/ConjugateGradient.h:30-85 for a CSR matrix resident in HBM.
//
// All streaming kernels are HBM-bound: matrix values / column indices are read once with streaming
// (evict-first) loads, vectors with 16-byte loads, the x-gather of the SpMV is served by L1/L2.
// Grid-wide reductions are deterministic: every CTA writes one partial, the consumer kernel re-sums the
// partials in a fixed order (no atomics, no host round trip).
#pragma once
#include "comm.cuh"
#include "common.cuh"

namespace feab {

constexpr int kTileRows = kBlock;       // CSR rows per tile (one row per thread in the reduce phase)
constexpr int kTileCapPerRow = 12;      // shared-memory capacity in entries per row of a tile
constexpr int kTileCap = kTileRows * kTileCapPerRow;

template <typename R>
struct CgState {
    R abs_new[2];   // r.z, double-buffered by iteration parity
    R thr;          // max(tol^2 |b|^2, tiny)   ConjugateGradient.h:51-52
    R rhs2;         // |b|^2
    R rn2;          // |r|^2 of the last test
    R tol;
    int iter1;      // SpMV launches so far in this solve
    int done;
    int iters;      // Eigen's `iters` on return
    int maxit;
};

struct StepScalars {  // per-step coefficients, promoted from the reference's floats on the host
    double dt, c0, c1, c2, c3;
};

template <typename R>
__device__ __forceinline__ R ld_stream(const R* p) {
    return __ldcs(p);
}

// ------------------------------------------------------------------------------------------------
// CPUSolver::get_surface_value_vector (src/FEM/CPUSolver.cpp:13-21)
// ------------------------------------------------------------------------------------------------
// surface->values is float: the vector it feeds passes through float once per step.  Between steps the solver keeps its
// own (unrounded) solution and the RHS kernels apply this rounding when they read it, so gather_values only has to run when
// somebody else wrote values (brush, set_values, host upload); same bits either way.
template <typename R>
__device__ __forceinline__ R through_float(R x) {
    return (R)(float)x;
}

template <typename R>
__global__ void gather_values(const float* __restrict__ values, const int* __restrict__ colmap, int64_t V, R* __restrict__ x, Comm cm,
                              HaloPlan hp) {
    // partitioned: the ghost entries of values come from the neighbours' end-of-step pushes; the caller launches
    // wait_kernel(state) first (one CTA spins, not V / 256 of them)
    if (comm_failed(cm)) return;
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < V) {
        int m = colmap[i];  // idx_map on one GPU; owned + ghost columns on a partition
        if (m >= 0) x[m] = (R)values[i];
    }
}

// ------------------------------------------------------------------------------------------------
// K2: operator values on the shared pattern (src/FEM/CPUSolver.cpp:77, :95, :113, :133, :136).
// The reference re-forms these sparse sums every step; here they are cached until the parameters change.
//   eq 0 heat   A0 = M/dt + kappa K
//   eq 1 wave   A0 = M/dt + ((c c) K) dt
//   eq 2 adv    A0 = (M/dt + c K) - Adv  (transposed == 1: the value Eigen's A^T-walk sees at this CSR slot)
//   eq 3 GS     A0 = M/dt + Du K,  A1 = M/dt + Dv K
// ------------------------------------------------------------------------------------------------
template <typename R>
__global__ void __launch_bounds__(kBlock) build_operator(int eq, int transposed, int64_t nnz, const R* __restrict__ Mv,
                                                          const R* __restrict__ Kv, const R* __restrict__ Av,
                                                          const R* __restrict__ AvT, StepScalars sc, R* __restrict__ Md,
                                                          R* __restrict__ A0, R* __restrict__ A1) {
    using O = Ops<R>;
    const R dt = (R)sc.dt, c0 = (R)sc.c0, c1 = (R)sc.c1;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += (int64_t)gridDim.x * blockDim.x) {
        R md = O::div(Mv[k], dt);
        R kv = Kv[k];
        Md[k] = md;
        if (eq == 0) {
            A0[k] = O::add(md, O::mul(c0, kv));
        } else if (eq == 1) {
            R cc = O::mul(c0, c0);
            A0[k] = O::add(md, O::mul(O::mul(cc, kv), dt));
        } else if (eq == 2) {
            R a = transposed ? AvT[k] : Av[k];
            A0[k] = O::sub(O::add(md, O::mul(c0, kv)), a);
        } else {
            A0[k] = O::add(md, O::mul(c0, kv));
            A1[k] = O::add(md, O::mul(c1, kv));
        }
    }
}

// DiagonalPreconditioner::factorize (BasicPreconditioners.h:63-75)
template <typename R>
__global__ void inv_diag(const R* __restrict__ A, const int* __restrict__ diag, int64_t n, R* __restrict__ dinv) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r < n) {
        int d = diag[r];
        R a = d >= 0 ? A[d] : (R)0;
        dinv[r] = (d >= 0 && a != (R)0) ? Ops<R>::div((R)1, a) : (R)1;
    }
}

// ------------------------------------------------------------------------------------------------
// Tile helpers (CSR-stream): a CTA owns kTileRows consecutive rows; their entries are one contiguous range
// of the value / column arrays, streamed with fully coalesced loads, multiplied with the gathered vector and
// parked in shared memory; then each thread reduces its own row from shared memory in the reference's order.
// ------------------------------------------------------------------------------------------------
struct TileRange {
    int kb, ke;   // this thread's row
    int tb, te;   // whole tile
};

__device__ __forceinline__ TileRange tile_range(const int* __restrict__ rowptr, int64_t n, int tile, int* s_k) {
    TileRange t;
    int64_t r = (int64_t)tile * kTileRows + threadIdx.x;
    t.kb = t.ke = 0;
    if (r < n) {
        t.kb = rowptr[r];
        t.ke = rowptr[r + 1];
        if (threadIdx.x == 0) s_k[0] = t.kb;
        if (r == n - 1 || threadIdx.x == kTileRows - 1) s_k[1] = t.ke;
    }
    __syncthreads();
    t.tb = s_k[0];
    t.te = s_k[1];
    return t;
}

template <typename R>
__device__ __forceinline__ void stage_products(const R* __restrict__ val, const int* __restrict__ col, const R* __restrict__ x,
                                               int tb, int cnt, R* __restrict__ prod) {
#pragma unroll 4
    for (int i = threadIdx.x; i < cnt; i += kBlock) prod[i] = ld_stream(val + tb + i) * x[ld_stream(col + tb + i)];
}

// Eigen's row-major SpMV walk: two alternating accumulators (SparseDenseProduct.h:66-88)
template <typename R>
__device__ __forceinline__ R row_sum_two_acc(const R* __restrict__ prod, int b, int e) {
    R ta = (R)0, tb = (R)0;
    int k = b;
    for (; k + 1 < e; k += 2) {
        ta += prod[k];
        tb += prod[k + 1];
    }
    if (k < e) ta += prod[k];
    return ta + tb;
}
// Eigen's column-major sparse * dense: per row, one accumulator in ascending column order (:96-118)
template <typename R>
__device__ __forceinline__ R row_sum_seq(const R* __restrict__ prod, int b, int e, R acc) {
    for (int k = b; k < e; k++) acc += prod[k];
    return acc;
}

// ------------------------------------------------------------------------------------------------
// K3: y = A p fused with the partial of p.y.  One launch per PCG iteration.
// ------------------------------------------------------------------------------------------------
template <typename R>
__global__ void __launch_bounds__(kBlock) spmv_dot(const int* __restrict__ rowptr, const int* __restrict__ col,
                                                   const R* __restrict__ val, const R* __restrict__ x, R* __restrict__ y,
                                                   int64_t n, int num_tiles, CgState<R>* __restrict__ st,
                                                   R* __restrict__ part_pap) {
    extern __shared__ __align__(16) unsigned char smraw[];
    R* prod = reinterpret_cast<R*>(smraw);
    __shared__ int s_k[2];
    __shared__ R red[kBlock / 32 + 1];
    if (st != nullptr && st->done) return;
    R acc = (R)0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        TileRange t = tile_range(rowptr, n, tile, s_k);
        const int cnt = t.te - t.tb;
        int64_t r = (int64_t)tile * kTileRows + threadIdx.x;
        R sum;
        if (cnt <= kTileCap) {
            stage_products<R>(val, col, x, t.tb, cnt, prod);
            __syncthreads();
            sum = row_sum_two_acc<R>(prod, t.kb - t.tb, t.ke - t.tb);
        } else {  // very long rows in this tile: walk global memory directly
            R ta = (R)0, tb2 = (R)0;
            int k = t.kb;
            for (; k + 1 < t.ke; k += 2) {
                ta += val[k] * x[col[k]];
                tb2 += val[k + 1] * x[col[k + 1]];
            }
            if (k < t.ke) ta += val[k] * x[col[k]];
            sum = ta + tb2;
        }
        if (r < n) {
            y[r] = sum;
            acc += x[r] * sum;
        }
        __syncthreads();  // prod / s_k are reused by the next tile
    }
    acc = block_sum(acc, red);
    if (threadIdx.x == 0) {
        part_pap[blockIdx.x] = acc;
        if (blockIdx.x == 0 && st != nullptr) st->iter1 += 1;  // nobody reads iter1 inside this kernel
    }
}

// ------------------------------------------------------------------------------------------------
// K4: alpha = absNew / p.Ap;  x += alpha p;  r -= alpha Ap;  partials of |r|^2 and r.(dinv r)
// (ConjugateGradient.h:68-76; z itself is never stored).
// ------------------------------------------------------------------------------------------------
template <typename R>
struct Vec;
template <>
struct Vec<double> {
    using T = double2;
    static constexpr int N = 2;
};
template <>
struct Vec<float> {
    using T = float4;
    static constexpr int N = 4;
};
template <typename R>
__device__ __forceinline__ void vload(const R* p, R* out);
template <>
__device__ __forceinline__ void vload<double>(const double* p, double* out) {
    double2 v = *reinterpret_cast<const double2*>(p);
    out[0] = v.x;
    out[1] = v.y;
}
template <>
__device__ __forceinline__ void vload<float>(const float* p, float* out) {
    float4 v = *reinterpret_cast<const float4*>(p);
    out[0] = v.x;
    out[1] = v.y;
    out[2] = v.z;
    out[3] = v.w;
}
template <typename R>
__device__ __forceinline__ void vstore(R* p, const R* in);
template <>
__device__ __forceinline__ void vstore<double>(double* p, const double* in) {
    *reinterpret_cast<double2*>(p) = make_double2(in[0], in[1]);
}
template <>
__device__ __forceinline__ void vstore<float>(float* p, const float* in) {
    *reinterpret_cast<float4*>(p) = make_float4(in[0], in[1], in[2], in[3]);
}

// Streaming ("evict first") variants for data touched once per iteration, so that it does not push re-used vectors out of L2.
template <typename R>
__device__ __forceinline__ void vload_cs(const R* p, R* out);
template <>
__device__ __forceinline__ void vload_cs<double>(const double* p, double* out) {
    double2 v = __ldcs(reinterpret_cast<const double2*>(p));
    out[0] = v.x;
    out[1] = v.y;
}
template <>
__device__ __forceinline__ void vload_cs<float>(const float* p, float* out) {
    float4 v = __ldcs(reinterpret_cast<const float4*>(p));
    out[0] = v.x;
    out[1] = v.y;
    out[2] = v.z;
    out[3] = v.w;
}
template <typename R>
__device__ __forceinline__ void vstore_cs(R* p, const R* in);
template <>
__device__ __forceinline__ void vstore_cs<double>(double* p, const double* in) {
    __stcs(reinterpret_cast<double2*>(p), make_double2(in[0], in[1]));
}
template <>
__device__ __forceinline__ void vstore_cs<float>(float* p, const float* in) {
    __stcs(reinterpret_cast<float4*>(p), make_float4(in[0], in[1], in[2], in[3]));
}

template <typename R>
__global__ void __launch_bounds__(kBlock) cg_update(R* __restrict__ x, R* __restrict__ r, const R* __restrict__ p,
                                                    const R* __restrict__ y, const R* __restrict__ dinv, int64_t n,
                                                    CgState<R>* __restrict__ st, const R* __restrict__ part_pap, int n_part,
                                                    R* __restrict__ part_rr, R* __restrict__ part_rz, Comm cm) {
    using O = Ops<R>;
    __shared__ R red[kBlock / 32 + 1];
    if (st->done) return;
    R pap;
    if (cm.world > 1) {
        R t[1];
        ar_consume<R, 1>(cm, t);
        pap = t[0];
    } else {
        pap = sum_partials<R>(part_pap, n_part, red);
    }
    const int it = st->iter1;
    const R alpha = st->abs_new[(it - 1) & 1] / pap;
    constexpr int N = Vec<R>::N;
    R rr = (R)0, rz = (R)0;
    const int64_t nvec = n / N;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
        R xv[N], rv[N], pv[N], yv[N], dv[N];
        vload<R>(x + i * N, xv);
        vload<R>(r + i * N, rv);
        vload<R>(p + i * N, pv);
        vload<R>(y + i * N, yv);
        vload<R>(dinv + i * N, dv);
#pragma unroll
        for (int q = 0; q < N; q++) {
            xv[q] = O::add(xv[q], O::mul(alpha, pv[q]));
            rv[q] = O::sub(rv[q], O::mul(alpha, yv[q]));
            R z = O::mul(dv[q], rv[q]);
            rr += rv[q] * rv[q];
            rz += rv[q] * z;
        }
        vstore<R>(x + i * N, xv);
        vstore<R>(r + i * N, rv);
    }
    if (blockIdx.x == 0) {  // scalar tail
        for (int64_t i = nvec * N + threadIdx.x; i < n; i += blockDim.x) {
            R xn = O::add(x[i], O::mul(alpha, p[i]));
            R rn = O::sub(r[i], O::mul(alpha, y[i]));
            x[i] = xn;
            r[i] = rn;
            R z = O::mul(dinv[i], rn);
            rr += rn * rn;
            rz += rn * z;
        }
    }
    rr = block_sum(rr, red);
    __syncthreads();
    rz = block_sum(rz, red);
    if (threadIdx.x == 0) {
        part_rr[blockIdx.x] = rr;
        part_rz[blockIdx.x] = rz;
    }
    if (cm.world > 1) {
        __syncthreads();
        ar_publish<R, 2>(cm, part_rr, (int)(part_rz - part_rr), red);
    }
}

// ------------------------------------------------------------------------------------------------
// K5: convergence test on the device, beta = absNew / absOld, p = dinv r + beta p
// (ConjugateGradient.h:72-81).  Thread 0 of CTA 0 publishes the loop state and, when the step runs as a CUDA
// graph, drives the WHILE node's condition.
// ------------------------------------------------------------------------------------------------
template <typename R>
__global__ void __launch_bounds__(kBlock) cg_direction(const R* __restrict__ r, R* __restrict__ p, const R* __restrict__ dinv,
                                                       int64_t n, CgState<R>* __restrict__ st, const R* __restrict__ part_rr,
                                                       const R* __restrict__ part_rz, int n_part,
                                                       cudaGraphConditionalHandle cond, int use_cond, Comm cm) {
    using O = Ops<R>;
    __shared__ R red[kBlock / 32 + 1];
    if (st->done) {
        if (use_cond && blockIdx.x == 0 && threadIdx.x == 0) cudaGraphSetConditional(cond, 0);
        return;
    }
    R rr, rz;
    if (cm.world > 1) {
        R t[2];
        ar_consume<R, 2>(cm, t);
        rr = t[0];
        rz = t[1];
    } else {
        rr = sum_partials<R>(part_rr, n_part, red);
        rz = sum_partials<R>(part_rz, n_part, red);
    }
    const int it = st->iter1;
    const bool conv = rr < st->thr;
    const R abs_old = st->abs_new[(it - 1) & 1];
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        st->rn2 = rr;
        st->abs_new[it & 1] = rz;
        int done = 0;
        if (conv) {
            done = 1;
            st->iters = it - 1;
        } else if (it >= st->maxit || !(rr <= Ops<R>::huge())) {  // also stop on NaN / Inf instead of spinning to 2n
            done = 1;
            st->iters = it;
        }
        if (done) st->done = 1;
        if (use_cond) cudaGraphSetConditional(cond, done ? 0u : 1u);
    }
    if (conv) return;
    const R beta = rz / abs_old;
    constexpr int N = Vec<R>::N;
    const int64_t nvec = n / N;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
        R rv[N], pv[N], dv[N];
        vload<R>(r + i * N, rv);
        vload<R>(p + i * N, pv);
        vload<R>(dinv + i * N, dv);
#pragma unroll
        for (int q = 0; q < N; q++) pv[q] = O::add(O::mul(dv[q], rv[q]), O::mul(beta, pv[q]));
        vstore<R>(p + i * N, pv);
    }
    if (blockIdx.x == 0)
        for (int64_t i = nvec * N + threadIdx.x; i < n; i += blockDim.x) p[i] = O::add(O::mul(dinv[i], r[i]), O::mul(beta, p[i]));
}

// Single CTA after the RHS kernel: |b|^2, threshold, r.z of the initial direction, rhs == 0 early-out
// (ConjugateGradient.h:43-60; the residual equals b because the initial guess is zero,
// IterativeSolverBase.h:357-360).
template <typename R>
__global__ void __launch_bounds__(kBlock) cg_begin(CgState<R>* __restrict__ st, const R* __restrict__ part_bb,
                                                   const R* __restrict__ part_rz, int n_part, R tol, int maxit,
                                                   cudaGraphConditionalHandle cond, int use_cond) {
    __shared__ R red[kBlock / 32 + 1];
    const R bb = sum_partials<R>(part_bb, n_part, red);
    const R rz = sum_partials<R>(part_rz, n_part, red);
    if (threadIdx.x == 0) {
        R thr = tol * tol * bb;
        if (thr < Ops<R>::tiny()) thr = Ops<R>::tiny();
        st->rhs2 = bb;
        st->thr = thr;
        st->rn2 = bb;
        st->tol = tol;
        st->abs_new[0] = rz;
        st->abs_new[1] = (R)0;
        st->iter1 = 0;
        st->iters = 0;
        st->maxit = maxit;
        int done = (bb == (R)0 || bb < thr || maxit <= 0) ? 1 : 0;
        if (bb == (R)0) st->rn2 = (R)0;
        st->done = done;
        if (use_cond) cudaGraphSetConditional(cond, done ? 0u : 1u);
    }
}

// ------------------------------------------------------------------------------------------------
// K6: right-hand side fused with the PCG start (r = b, p = dinv b, partials of |b|^2 and b.p).
//   eq 0 / 2: b = (M/dt) u                                        src/FEM/CPUSolver.cpp:78, :96
//   eq 1    : b = (M/dt) v - ((c c) K) u, the K.u terms accumulate onto the M.v sum    :114
//   eq 3    : b_u = ((M/dt) u - u v v) + f (1 - u);  b_v = ((M/dt) v + u v v) - (f + k) v   :134, :137
// Row sums use one accumulator in ascending column order, like Eigen's column-major sparse * dense.
// ------------------------------------------------------------------------------------------------
template <typename R, int EQ>
__global__ void __launch_bounds__(kBlock) rhs_init(const int* __restrict__ rowptr, const int* __restrict__ col,
                                                   const R* __restrict__ Md, const R* __restrict__ Kv, const R* __restrict__ u,
                                                   const R* __restrict__ v, const R* __restrict__ dinv0,
                                                   const R* __restrict__ dinv1, int64_t n, int num_tiles, StepScalars sc,
                                                   R* __restrict__ r0, R* __restrict__ p0, R* __restrict__ r1, R* __restrict__ p1,
                                                   R* __restrict__ part /* 4 x gridDim.x: bb0, rz0, bb1, rz1 */, Comm cm, HaloPlan hp,
                                                   int publish) {
    using O = Ops<R>;
    extern __shared__ __align__(16) unsigned char smraw[];
    R* prod0 = reinterpret_cast<R*>(smraw);
    R* prod1 = prod0 + kTileCap;
    __shared__ int s_k[2];
    __shared__ R red[kBlock / 32 + 1];
    R bb0 = (R)0, rz0 = (R)0, bb1 = (R)0, rz1 = (R)0;
    const R cc = O::mul((R)sc.c0, (R)sc.c0);
    bool pushed = false;
    if (cm.world > 1) {  // ghost entries of u / v come from the neighbours' end-of-step pushes
        if (comm_failed(cm)) return;
        if (threadIdx.x == 0) wait_flags(cm, hp, 1, *(volatile unsigned long long*)&cm.loc->state_seq);
        __syncthreads();
        if (comm_failed(cm)) return;
    }
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        TileRange t = tile_range(rowptr, n, tile, s_k);
        const int cnt = t.te - t.tb;
        int64_t r = (int64_t)tile * kTileRows + threadIdx.x;
        R s0 = (R)0, s1 = (R)0;
        if (cnt <= kTileCap) {
            // the vector surface->values feeds (u; Gray-Scott: v) is read through float (see through_float)
            if (EQ == 0 || EQ == 2) {
#pragma unroll 4
                for (int i = threadIdx.x; i < cnt; i += kBlock)
                    prod0[i] = ld_stream(Md + t.tb + i) * through_float(u[ld_stream(col + t.tb + i)]);
            } else if (EQ == 1) {
#pragma unroll 4
                for (int i = threadIdx.x; i < cnt; i += kBlock) {
                    int c = ld_stream(col + t.tb + i);
                    prod0[i] = ld_stream(Md + t.tb + i) * v[c];
                    prod1[i] = O::mul(cc, ld_stream(Kv + t.tb + i)) * O::mul((R)-1, through_float(u[c]));
                }
            } else {
#pragma unroll 4
                for (int i = threadIdx.x; i < cnt; i += kBlock) {
                    int c = ld_stream(col + t.tb + i);
                    R m = ld_stream(Md + t.tb + i);
                    prod0[i] = m * u[c];
                    prod1[i] = m * through_float(v[c]);
                }
            }
            __syncthreads();
            s0 = row_sum_seq<R>(prod0, t.kb - t.tb, t.ke - t.tb, (R)0);
            if (EQ == 1) s0 = row_sum_seq<R>(prod1, t.kb - t.tb, t.ke - t.tb, s0);
            if (EQ == 3) s1 = row_sum_seq<R>(prod1, t.kb - t.tb, t.ke - t.tb, (R)0);
        } else {
            for (int k = t.kb; k < t.ke; k++) {
                int c = col[k];
                if (EQ == 0 || EQ == 2) s0 += Md[k] * through_float(u[c]);
                if (EQ == 1) s0 += Md[k] * v[c];
                if (EQ == 3) {
                    s0 += Md[k] * u[c];
                    s1 += Md[k] * through_float(v[c]);
                }
            }
            if (EQ == 1)
                for (int k = t.kb; k < t.ke; k++) s0 += O::mul(cc, Kv[k]) * O::mul((R)-1, through_float(u[col[k]]));
        }
        if (r < n) {
            if (EQ == 3) {
                const R f = (R)sc.c2, fk = O::add((R)sc.c2, (R)sc.c3);
                R ui = u[r], vi = through_float(v[r]);
                R uvv = O::mul(ui, O::mul(vi, vi));
                R bu = O::add(O::sub(s0, uvv), O::mul(f, O::sub((R)1, ui)));
                R bv = O::sub(O::add(s1, uvv), O::mul(fk, vi));
                R pu = O::mul(dinv0[r], bu), pv = O::mul(dinv1[r], bv);
                r0[r] = bu;
                p0[r] = pu;
                r1[r] = bv;
                p1[r] = pv;
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
                R pz = O::mul(dinv0[r], s0);
                r0[r] = s0;
                p0[r] = pz;
                bb0 += s0 * s0;
                rz0 += s0 * pz;
                if (cm.world > 1 && r >= hp.n_int) {
                    push_boundary_row<R>(hp, kVecP0, r, pz);
                    pushed = true;
                }
            }
        }
        __syncthreads();
    }
    bb0 = block_sum(bb0, red);
    __syncthreads();
    rz0 = block_sum(rz0, red);
    __syncthreads();
    if (EQ == 3) {
        bb1 = block_sum(bb1, red);
        __syncthreads();
        rz1 = block_sum(rz1, red);
    }
    if (threadIdx.x == 0) {
        part[blockIdx.x] = bb0;
        part[gridDim.x + blockIdx.x] = rz0;
        if (EQ == 3) {
            part[2 * gridDim.x + blockIdx.x] = bb1;
            part[3 * gridDim.x + blockIdx.x] = rz1;
        }
    }
    if (publish) {  // totals -> every rank's mailbox (persistent solve kernels and partitioned handles start from there)
        if (cm.world > 1 && __syncthreads_or(pushed ? 1 : 0)) __threadfence_system();
        __syncthreads();
        ar_publish<R, (EQ == 3 ? 4 : 2)>(cm, part, (int)gridDim.x, red, &hp, cm.world > 1);
    }
}

// ------------------------------------------------------------------------------------------------
// Step epilogue, one thread per mesh node: (wave) u += v dt (src/FEM/CPUSolver.cpp:118);
// map_vector_to_surface incl. Dirichlet zeros (:28-36); has_numerical_instability (:41-48) as a sticky flag.
// sol_is_v: Gray-Scott writes v back to the surface (:144); the other equations write u.
// ------------------------------------------------------------------------------------------------
// Partitioned handle: the owner of a boundary row also stores the row's u, v and surface value into the neighbours that
// keep it as a ghost (they read them in the next step's RHS), and the last CTA raises the neighbours' state flags.
// unstable: stamped with the step number (no reset launch per step); feab_unstable compares it with the current step.
template <typename R>
__global__ void __launch_bounds__(kBlock) finish_step(float* __restrict__ values, const int* __restrict__ idx_map,
                                                      const int* __restrict__ rowmap, int64_t V, R* __restrict__ u, R* __restrict__ v,
                                                      int wave, R dt, int sol_is_v, int* __restrict__ unstable, int stamp, Comm cm,
                                                      HaloPlan hp) {
    if (comm_failed(cm)) return;  // fail closed: a solve that lost a peer writes nothing back
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool bad = false, pushed = false;
    if (i < V) {
        int m = rowmap[i];  // == idx_map on one GPU; on a partition only the owned nodes are written
        if (m >= 0) {
            R ui = u[m], vi = v[m];
            if (wave) {  // u here is last step's result: it passed through surface->values (float) in between
                ui = Ops<R>::add(through_float(ui), Ops<R>::mul(vi, dt));
                u[m] = ui;
            }
            const float out = (float)(sol_is_v ? vi : ui);
            values[i] = out;
            bad = ui > (R)1e4f || vi > (R)1e4f;
            if (cm.world > 1 && m >= hp.n_int) {
                const int* t = hp.push_tab + (size_t)(m - hp.n_int) * hp.nnb;
                for (int b = 0; b < hp.nnb; b++) {
                    const int d = t[b];
                    if (d >= 0) {
                        const int q = hp.nb_rank[b];
                        reinterpret_cast<R*>(hp.base[q] + hp.off_vec[kVecU][q])[d] = ui;
                        reinterpret_cast<R*>(hp.base[q] + hp.off_vec[kVecV][q])[d] = vi;
                        reinterpret_cast<float*>(hp.base[q] + hp.off_vec[kVecValues][q])[i] = out;
                        pushed = true;
                    }
                }
            }
        } else if (idx_map[i] < 0) {
            values[i] = 0.0f;
        }
    }
    if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicMax(unstable, stamp);
    if (cm.world > 1) {
        __shared__ int s_last;
        if (__syncthreads_or(pushed ? 1 : 0)) __threadfence_system();
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            s_last = atomicAdd(&cm.loc->ticket, 1u) == gridDim.x - 1 ? 1 : 0;
        }
        __syncthreads();
        if (s_last && threadIdx.x == 0) {
            cm.loc->ticket = 0;
            __threadfence_system();
            const unsigned long long s = cm.loc->state_seq + 1;
            cm.loc->state_seq = s;
            for (int b = 0; b < hp.nnb; b++) st_release_sys(&cm.ctrl[hp.nb_rank[b]]->state_flag[cm.rank], s);
        }
    }
}

// The same epilogue on a PARTITION: one pass over this rank's own rows (not over all V mesh nodes -- every rank would read the
// whole rowmap and V / 256 CTAs would queue up at the ticket), bounded grid.  node_of_row[m] = mesh node of local row m.
// Dirichlet nodes (no row anywhere) are zeroed by every rank from the list dir_nodes, like map_vector_to_surface does
// (src/FEM/CPUSolver.cpp:28-36).  Boundary rows also go to the neighbours' ghosts; the last CTA raises the state flags.
template <typename R>
__global__ void __launch_bounds__(kBlock) finish_rows(float* __restrict__ values, const int* __restrict__ node_of_row, int64_t n,
                                                      const int* __restrict__ dir_nodes, int64_t n_dir, R* __restrict__ u, R* __restrict__ v,
                                                      int wave, R dt, int sol_is_v, int* __restrict__ unstable, int stamp, Comm cm,
                                                      HaloPlan hp) {
    if (comm_failed(cm)) return;  // fail closed: a solve that lost a peer writes nothing back
    bool bad = false, pushed = false;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; m < n; m += stride) {
        const int node = node_of_row[m];
        R ui = u[m], vi = v[m];
        if (wave) {
            ui = Ops<R>::add(through_float(ui), Ops<R>::mul(vi, dt));
            u[m] = ui;
        }
        const float out = (float)(sol_is_v ? vi : ui);
        values[node] = out;
        bad = bad || ui > (R)1e4f || vi > (R)1e4f;
        if (m >= hp.n_int) {
            const int* t = hp.push_tab + (size_t)(m - hp.n_int) * hp.nnb;
            for (int b = 0; b < hp.nnb; b++) {
                const int d = t[b];
                if (d >= 0) {
                    const int q = hp.nb_rank[b];
                    reinterpret_cast<R*>(hp.base[q] + hp.off_vec[kVecU][q])[d] = ui;
                    reinterpret_cast<R*>(hp.base[q] + hp.off_vec[kVecV][q])[d] = vi;
                    reinterpret_cast<float*>(hp.base[q] + hp.off_vec[kVecValues][q])[node] = out;
                    pushed = true;
                }
            }
        }
    }
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_dir; i += stride) values[dir_nodes[i]] = 0.0f;
    if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicMax(unstable, stamp);
    __shared__ int s_last;
    if (__syncthreads_or(pushed ? 1 : 0)) __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        s_last = atomicAdd(&cm.loc->ticket, 1u) == gridDim.x - 1 ? 1 : 0;
    }
    __syncthreads();
    if (s_last && threadIdx.x == 0) {
        cm.loc->ticket = 0;
        __threadfence_system();
        const unsigned long long s = cm.loc->state_seq + 1;
        cm.loc->state_seq = s;
        for (int b = 0; b < hp.nnb; b++) st_release_sys(&cm.ctrl[hp.nb_rank[b]]->state_flag[cm.rank], s);
    }
}

// surface->values -> solver vector on a partition: one pass over this rank's columns (own rows + ghosts)
template <typename R>
__global__ void __launch_bounds__(kBlock) gather_cols(const float* __restrict__ values, const int* __restrict__ node_of_col, int64_t n_cols,
                                                      R* __restrict__ x, Comm cm) {
    if (comm_failed(cm)) return;
    for (int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; m < n_cols; m += (int64_t)gridDim.x * blockDim.x) {
        const int node = node_of_col[m];
        if (node >= 0) x[m] = (R)values[node];
    }
}
// FEAB_SYNC_ROOT on a peer: copy the surface values of this rank's columns (own rows + ghosts) from rank 0's mirror, read over
// NVLink, into the local mirror (comm.cuh root_signal / root_wait order it after rank 0's upload)
__global__ void __launch_bounds__(kBlock) pull_cols(const float* root_values, const int* __restrict__ node_of_col, int64_t n_cols,
                                                    float* __restrict__ mirror, Comm cm) {
    if (comm_failed(cm)) return;
    for (int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; m < n_cols; m += (int64_t)gridDim.x * blockDim.x) {
        const int node = node_of_col[m];
        if (node >= 0) mirror[node] = root_values[node];
    }
}

__global__ void brush_nodes(float* __restrict__ values, const long long* __restrict__ nodes, int64_t count, int64_t V, float strength) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count && nodes[i] >= 0 && nodes[i] < V) values[nodes[i]] = strength;
}

}  // namespace feab
