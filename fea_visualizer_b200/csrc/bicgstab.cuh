// Jacobi right-preconditioned BiCGSTAB for the non-symmetric advection-diffusion operator -- the "correct"
// mode next to the reference-parity PCG-on-A^T.  Follows internal::bicgstab
// (lib/eigen/Eigen/src/IterativeLinearSolvers/BiCGSTAB.h:31-119) with the two stated deviations of the
// oracle (oracle/fea_oracle_impl.h): relative stopping test |r| > tol |b|, and a breakdown restarts
// deterministically with r = b - A x, r0 = r (no setRandom).
//
// One loop trip = 8 launches: two restart kernels (no-ops unless a breakdown was flagged), p/y update,
// SpMV + 2 dots, s/z update, SpMV + 2 dots, x/r update + 2 dots, and a single-CTA scalar kernel that
// closes the trip and drives the graph's WHILE condition.  Scalars that a kernel both reads and replaces are
// double-buffered by trip parity; everything else crosses kernels through per-CTA partials.
#pragma once
#include "common.cuh"
#include "solver.cuh"

namespace feab {

constexpr int kBicgKernelsPerIter = 8;

template <typename R>
struct BicgState {
    R rho[2];       // by trip parity
    R alpha, w;
    R r0_norm, r_norm, rhs_norm;
    R tol;
    int iters;      // Eigen's i
    int trips;      // loop trips executed (for launch accounting)
    int restarts;
    int restart_req, restart_now, skip;
    int done;
    int maxit;
};

template <typename R>
struct BicgVectors {
    R *x, *r, *r0, *v, *p, *y, *s, *z, *t, *b;
    const R* dinv;
};
struct BicgGrid {
    int g_tile, g_vec, g_part, num_tiles;
};

// partial layout inside `part` (each g_part long):
//  0: |r|^2   1: r0.r   2: r0.v   3: v.v   4: t.s   5: t.t   6: restart |r|^2
template <typename R>
__host__ __device__ __forceinline__ R* bpart(R* part, int which, int g_part) {
    return part + (size_t)which * g_part;
}

// y = A x (row sums in ascending column order, one accumulator), optionally fused with a.y and y.y partials
template <typename R>
__global__ void __launch_bounds__(kBlock) spmv_plain(const int* __restrict__ rowptr, const int* __restrict__ col,
                                                     const R* __restrict__ val, const R* __restrict__ x, R* __restrict__ y,
                                                     const R* __restrict__ a, int64_t n, int num_tiles,
                                                     const BicgState<R>* __restrict__ st, int only_on_restart,
                                                     R* __restrict__ part_ay, R* __restrict__ part_yy) {
    extern __shared__ __align__(16) unsigned char smraw[];
    R* prod = reinterpret_cast<R*>(smraw);
    __shared__ int s_k[2];
    __shared__ R red[kBlock / 32 + 1];
    if (st->done) return;
    if (only_on_restart ? !st->restart_now : st->skip) return;
    R d0 = (R)0, d1 = (R)0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        TileRange t = tile_range(rowptr, n, tile, s_k);
        const int cnt = t.te - t.tb;
        int64_t r = (int64_t)tile * kTileRows + threadIdx.x;
        R sum = (R)0;
        if (cnt <= kTileCap) {
            stage_products<R>(val, col, x, t.tb, cnt, prod);
            __syncthreads();
            sum = row_sum_seq<R>(prod, t.kb - t.tb, t.ke - t.tb, (R)0);
        } else {
            for (int k = t.kb; k < t.ke; k++) sum += val[k] * x[col[k]];
        }
        if (r < n) {
            y[r] = sum;
            if (a != nullptr) {
                d0 += a[r] * sum;
                d1 += sum * sum;
            }
        }
        __syncthreads();
    }
    if (a != nullptr) {
        d0 = block_sum(d0, red);
        __syncthreads();
        d1 = block_sum(d1, red);
        if (threadIdx.x == 0) {
            part_ay[blockIdx.x] = d0;
            part_yy[blockIdx.x] = d1;
        }
    }
}

// r0 = r (= b from rhs_init), b saved, v = p = 0, partials |b|^2 (also r0.r)
template <typename R>
__global__ void __launch_bounds__(kBlock) bicg_begin(BicgVectors<R> bv, int64_t n, R* __restrict__ part, int g_part) {
    __shared__ R red[kBlock / 32 + 1];
    R bb = (R)0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        R b = bv.r[i];
        bv.r0[i] = b;
        bv.b[i] = b;
        bv.v[i] = (R)0;
        bv.p[i] = (R)0;
        bb += b * b;
    }
    bb = block_sum(bb, red);
    if (threadIdx.x == 0) {
        bpart(part, 0, g_part)[blockIdx.x] = bb;
        bpart(part, 1, g_part)[blockIdx.x] = bb;
    }
}
template <typename R>
__global__ void __launch_bounds__(kBlock) bicg_begin2(BicgState<R>* __restrict__ st, const R* __restrict__ part, int n_part, R tol,
                                                      int maxit, cudaGraphConditionalHandle cond, int use_cond) {
    __shared__ R red[kBlock / 32 + 1];
    const R bb = sum_partials<R>(part, n_part, red);
    if (threadIdx.x == 0) {
        R nb = Ops<R>::sqrt(bb);
        st->rho[0] = st->rho[1] = (R)1;
        st->alpha = (R)0;
        st->w = (R)1;
        st->r0_norm = st->r_norm = st->rhs_norm = nb;
        st->tol = tol;
        st->iters = 0;
        st->trips = 0;
        st->restarts = 0;
        st->restart_req = st->restart_now = st->skip = 0;
        st->maxit = maxit;
        int done = (nb == (R)0 || !(nb > tol * nb) || maxit <= 0) ? 1 : 0;
        st->done = done;
        if (use_cond) cudaGraphSetConditional(cond, done ? 0u : 1u);
    }
}

// restart: r = b - t (t = A x from the restart SpMV), r0 = r, partial |r|^2
template <typename R>
__global__ void __launch_bounds__(kBlock) bicg_restart_vec(BicgVectors<R> bv, int64_t n, const BicgState<R>* __restrict__ st,
                                                           R* __restrict__ part, int g_part) {
    __shared__ R red[kBlock / 32 + 1];
    if (st->done || !st->restart_now) return;
    R rr = (R)0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        R r = Ops<R>::sub(bv.b[i], bv.t[i]);
        bv.r[i] = r;
        bv.r0[i] = r;
        rr += r * r;
    }
    rr = block_sum(rr, red);
    if (threadIdx.x == 0) bpart(part, 6, g_part)[blockIdx.x] = rr;
}

// rho, breakdown test, beta;  p = r + beta (p - w v);  y = dinv p        (BiCGSTAB.h:66-85)
template <typename R>
__global__ void __launch_bounds__(kBlock) bicg_p(BicgVectors<R> bv, int64_t n, BicgState<R>* __restrict__ st, R* __restrict__ part,
                                                 int g_part, int n_part) {
    using O = Ops<R>;
    __shared__ R red[kBlock / 32 + 1];
    if (st->done) return;
    const int par = st->trips & 1;
    R rho, beta, r0n = st->r0_norm, rn = st->r_norm;
    const R w = st->w, alpha = st->alpha;
    bool brk = false;
    if (st->restart_now) {
        const R rr = sum_partials<R>(bpart(part, 6, g_part), n_part, red);
        rho = rr;
        r0n = rn = O::sqrt(rr);
        beta = (R)0;
    } else {
        rho = sum_partials<R>(bpart(part, 1, g_part), n_part, red);
        const R rho_old = st->rho[par ^ 1];
        const R mx = r0n > rn ? r0n : rn, mn = r0n < rn ? r0n : rn;
        brk = fabs((double)rho) / (double)mx < (double)(Ops<R>::eps() * mn);
        beta = (rho / rho_old) * (alpha / w);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        st->rho[par] = rho;
        if (st->restart_now) {
            st->r0_norm = r0n;
            st->r_norm = rn;
        }
        if (brk) {
            st->restart_req = 1;
            st->skip = 1;
        }
    }
    if (brk) return;
    const bool fresh = st->restart_now != 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        R pn = fresh ? bv.r[i] : O::add(bv.r[i], O::mul(beta, O::sub(bv.p[i], O::mul(w, bv.v[i]))));
        bv.p[i] = pn;
        bv.y[i] = O::mul(bv.dinv[i], pn);
    }
}

// theta = r0.v, small-angle breakdown test, alpha = rho / theta;  s = r - alpha v;  z = dinv s   (:87-104)
template <typename R>
__global__ void __launch_bounds__(kBlock) bicg_s(BicgVectors<R> bv, int64_t n, BicgState<R>* __restrict__ st, R* __restrict__ part,
                                                 int g_part, int n_part) {
    using O = Ops<R>;
    __shared__ R red[kBlock / 32 + 1];
    if (st->done || st->skip) return;
    const R theta = sum_partials<R>(bpart(part, 2, g_part), n_part, red);
    const R vv = sum_partials<R>(bpart(part, 3, g_part), n_part, red);
    const R vn = O::sqrt(vv), r0n = st->r0_norm;
    const R mx = r0n > vn ? r0n : vn, mn = r0n < vn ? r0n : vn;
    const bool brk = fabs((double)theta) / (double)mx < (double)(Ops<R>::eps() * mn);
    const R rho = st->rho[st->trips & 1];
    const R alpha = rho / theta;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        if (brk) {
            st->restart_req = 1;
            st->skip = 1;
        } else {
            st->alpha = alpha;
        }
    }
    if (brk) return;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        R sn = O::sub(bv.r[i], O::mul(alpha, bv.v[i]));
        bv.s[i] = sn;
        bv.z[i] = O::mul(bv.dinv[i], sn);
    }
}

// w = t.s / t.t;  x += alpha y + w z;  r = s - w t;  partials |r|^2 and r0.r      (:106-115)
template <typename R>
__global__ void __launch_bounds__(kBlock) bicg_x(BicgVectors<R> bv, int64_t n, BicgState<R>* __restrict__ st, R* __restrict__ part,
                                                 int g_part, int n_part) {
    using O = Ops<R>;
    __shared__ R red[kBlock / 32 + 1];
    if (st->done || st->skip) return;
    const R ts = sum_partials<R>(bpart(part, 2, g_part), n_part, red);  // a.y with a = s, y = t
    const R tt = sum_partials<R>(bpart(part, 3, g_part), n_part, red);
    const R w = tt > (R)0 ? ts / tt : (R)0;
    const R alpha = st->alpha;
    if (blockIdx.x == 0 && threadIdx.x == 0) st->w = w;
    R rr = (R)0, rho = (R)0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        bv.x[i] = O::add(bv.x[i], O::add(O::mul(alpha, bv.y[i]), O::mul(w, bv.z[i])));
        R rn = O::sub(bv.s[i], O::mul(w, bv.t[i]));
        bv.r[i] = rn;
        rr += rn * rn;
        rho += bv.r0[i] * rn;
    }
    rr = block_sum(rr, red);
    __syncthreads();
    rho = block_sum(rho, red);
    if (threadIdx.x == 0) {
        bpart(part, 0, g_part)[blockIdx.x] = rr;
        bpart(part, 1, g_part)[blockIdx.x] = rho;
    }
}

// single CTA: closes the trip (r_norm, ++i, restart bookkeeping, loop condition)
template <typename R>
__global__ void __launch_bounds__(kBlock) bicg_end(BicgState<R>* __restrict__ st, const R* __restrict__ part, int g_part, int n_part,
                                                   cudaGraphConditionalHandle cond, int use_cond) {
    __shared__ R red[kBlock / 32 + 1];
    if (st->done) {
        if (use_cond && threadIdx.x == 0) cudaGraphSetConditional(cond, 0);
        return;
    }
    const R rr = sum_partials<R>(part, n_part, red);
    if (threadIdx.x == 0) {
        st->trips += 1;
        if (st->skip) {  // breakdown: nothing was advanced; the next trip restarts
            st->skip = 0;
            st->restart_now = st->restart_req;
            st->restart_req = 0;
            if (st->restarts++ == 0) st->iters = 0;
            if (st->restarts > 8) st->done = 1;
        } else {
            st->restart_now = 0;
            st->r_norm = Ops<R>::sqrt(rr);
            st->iters += 1;
        }
        // carry rho into the next trip's "old" slot is implicit: rho[par] was written this trip
        if (!(st->r_norm > st->tol * st->rhs_norm) || st->iters >= st->maxit) st->done = 1;
        if (use_cond) cudaGraphSetConditional(cond, st->done ? 0u : 1u);
    }
}

template <typename R>
void bicg_iteration(cudaStream_t s, const int* rowptr, const int* col, const R* A, BicgVectors<R> bv, int64_t n, BicgGrid g,
                    BicgState<R>* st, R* part, size_t smem, cudaGraphConditionalHandle cond, int use_cond) {
    const int np_vec = g.g_vec, np_tile = g.g_tile;
    // restart pair (no-ops unless restart_now)
    spmv_plain<R><<<g.g_tile, kBlock, smem, s>>>(rowptr, col, A, bv.x, bv.t, (const R*)nullptr, n, g.num_tiles, st, 1, (R*)nullptr, (R*)nullptr);
    DBG(s, "bicg_iteration: restart_spmv");
    bicg_restart_vec<R><<<g.g_vec, kBlock, 0, s>>>(bv, n, st, part, g.g_part);
    DBG(s, "bicg_iteration: restart_vec");
    bicg_p<R><<<g.g_vec, kBlock, 0, s>>>(bv, n, st, part, g.g_part, np_vec);
    DBG(s, "bicg_iteration: bicg_p");
    spmv_plain<R><<<g.g_tile, kBlock, smem, s>>>(rowptr, col, A, bv.y, bv.v, bv.r0, n, g.num_tiles, st, 0, bpart(part, 2, g.g_part),
                                                 bpart(part, 3, g.g_part));
    DBG(s, "bicg_iteration: spmv1");
    bicg_s<R><<<g.g_vec, kBlock, 0, s>>>(bv, n, st, part, g.g_part, np_tile);
    DBG(s, "bicg_iteration: bicg_s");
    spmv_plain<R><<<g.g_tile, kBlock, smem, s>>>(rowptr, col, A, bv.z, bv.t, bv.s, n, g.num_tiles, st, 0, bpart(part, 2, g.g_part),
                                                 bpart(part, 3, g.g_part));
    DBG(s, "bicg_iteration: spmv2");
    bicg_x<R><<<g.g_vec, kBlock, 0, s>>>(bv, n, st, part, g.g_part, np_tile);
    DBG(s, "bicg_iteration: bicg_x");
    bicg_end<R><<<1, kBlock, 0, s>>>(st, part, g.g_part, np_vec, cond, use_cond);
    DBG(s, "bicg_iteration: bicg_end");
}

}  // namespace feab
