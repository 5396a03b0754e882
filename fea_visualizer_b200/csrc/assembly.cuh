// Numeric assembly of the P1 mass, stiffness and advection matrices on the device.
// Per-element arithmetic restates src/FEM/FEMContext.cpp:70-202 operation by operation (including how glm
// and Eigen associate the 3-term sums, see oracle/fea_oracle_impl.h for the derivation), in the handle's
// real type, with the float vertex coordinates promoted first.
#pragma once
#include "common.cuh"

namespace feab {

template <typename R>
struct Element {
    R normal[3];
    R area;   // 0.5 * |e1 x e2|                          FEMContext.cpp:77
    R jdet;   // |e1 x e2| via Eigen cross().norm()        :126, :163
    R G[3][3];  // physical gradients                      :93-95
};

template <typename R>
__device__ __forceinline__ R sum3(R a, R b, R c) {  // Eigen fixed-size redux: t0 + (t1 + t2)
    return Ops<R>::add(a, Ops<R>::add(b, c));
}
template <typename R>
__device__ __forceinline__ R edot3(const R* a, const R* b) {  // Eigen Vector3 dot(): fma chain (InnerProduct.h:109-121)
    return Ops<R>::fma(a[2], b[2], Ops<R>::fma(a[1], b[1], Ops<R>::mul(a[0], b[0])));
}

template <typename R>
__device__ __forceinline__ void compute_element(const float* __restrict__ xyz, uint32_t ia, uint32_t ib, uint32_t ic,
                                                Element<R>& e) {
    using O = Ops<R>;
    R e1[3], e2[3], w[3];
#pragma unroll
    for (int d = 0; d < 3; d++) {
        R a = (R)xyz[3 * (int64_t)ia + d], b = (R)xyz[3 * (int64_t)ib + d], c = (R)xyz[3 * (int64_t)ic + d];
        e1[d] = O::sub(b, a);
        e2[d] = O::sub(c, a);
    }
    // glm::cross, glm::dot (left to right), normalize = v * (1 / sqrt(dot)), length = sqrt(dot)
    w[0] = O::sub(O::mul(e1[1], e2[2]), O::mul(e2[1], e1[2]));
    w[1] = O::sub(O::mul(e1[2], e2[0]), O::mul(e2[2], e1[0]));
    w[2] = O::sub(O::mul(e1[0], e2[1]), O::mul(e2[0], e1[1]));
    R ww = O::add(O::add(O::mul(w[0], w[0]), O::mul(w[1], w[1])), O::mul(w[2], w[2]));
    R sq = O::sqrt(ww);
    R inv = O::div((R)1, sq);
#pragma unroll
    for (int d = 0; d < 3; d++) e.normal[d] = O::mul(w[d], inv);
    e.area = O::mul((R)0.5, sq);
    // Eigen cross().norm(): squaredNorm is the unrolled redux x^2 + (y^2 + z^2)
    R cx = O::sub(O::mul(e1[1], e2[2]), O::mul(e1[2], e2[1]));
    R cy = O::sub(O::mul(e1[2], e2[0]), O::mul(e1[0], e2[2]));
    R cz = O::sub(O::mul(e1[0], e2[1]), O::mul(e1[1], e2[0]));
    e.jdet = O::sqrt(sum3(O::mul(cx, cx), O::mul(cy, cy), O::mul(cz, cz)));
    // J = [e1 | e2 | normal]; cofactor inverse (InverseImpl.h:107-142), transposed: At(i,j) = cof(i,j) / det
    R m[3][3];
#pragma unroll
    for (int d = 0; d < 3; d++) {
        m[d][0] = e1[d];
        m[d][1] = e2[d];
        m[d][2] = e.normal[d];
    }
    R cof[3][3];
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) {
            const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
            cof[i][j] = O::sub(O::mul(m[i1][j1], m[i2][j2]), O::mul(m[i1][j2], m[i2][j1]));
        }
    R det = sum3(O::mul(cof[0][0], m[0][0]), O::mul(cof[1][0], m[1][0]), O::mul(cof[2][0], m[2][0]));
    R invdet = O::div((R)1, det);
    R At[3][3];
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) At[i][j] = O::mul(cof[i][j], invdet);
    // G_i = At * refgrad_i with refgrads (-1,-1,0), (1,0,0), (0,1,0); each coefficient t0 + (t1 + t2)
#pragma unroll
    for (int r = 0; r < 3; r++) {
        e.G[0][r] = sum3(O::mul(At[r][0], (R)-1), O::mul(At[r][1], (R)-1), O::mul(At[r][2], (R)0));
        e.G[1][r] = sum3(O::mul(At[r][0], (R)1), O::mul(At[r][1], (R)0), O::mul(At[r][2], (R)0));
        e.G[2][r] = sum3(O::mul(At[r][0], (R)0), O::mul(At[r][1], (R)1), O::mul(At[r][2], (R)0));
    }
}

// t = normalized(v - (v.n)/(|n| |n|) n)   (src/FEM/FEMContext.cpp:179-180; Eigen normalized() leaves a
// vector whose squared norm is not > 0 unchanged)
template <typename R>
__device__ __forceinline__ void tangent_velocity(const Element<R>& e, const R* vel, R* tv) {
    using O = Ops<R>;
    R vn = edot3(vel, e.normal);
    R nn = O::sqrt(sum3(O::mul(e.normal[0], e.normal[0]), O::mul(e.normal[1], e.normal[1]), O::mul(e.normal[2], e.normal[2])));
    R coef = O::div(vn, O::mul(nn, nn));
#pragma unroll
    for (int d = 0; d < 3; d++) tv[d] = O::sub(vel[d], O::mul(coef, e.normal[d]));
    R tz = sum3(O::mul(tv[0], tv[0]), O::mul(tv[1], tv[1]), O::mul(tv[2], tv[2]));
    if (tz > (R)0) {
        R s = O::sqrt(tz);
#pragma unroll
        for (int d = 0; d < 3; d++) tv[d] = O::div(tv[d], s);
    }
}

// K1 (scatter mode): one thread per triangle, local 3x3 matrices, precomputed CSR scatter map, atomic adds.
// Off-diagonal entries receive at most two contributions, so their sums are order-independent; only the
// diagonal sums depend on the atomic arrival order (last-bit effects).  Value arrays must be zeroed first.
template <typename R>
__global__ void __launch_bounds__(kBlock) assemble_scatter(const float* __restrict__ xyz, const uint32_t* __restrict__ tri, int64_t T,
                                                            const uint8_t* __restrict__ on_boundary, const int* __restrict__ smap,
                                                            R vx, R vy, R vz, R* __restrict__ Mv, R* __restrict__ Kv,
                                                            R* __restrict__ Av, R* __restrict__ AvT) {
    using O = Ops<R>;
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= T) return;
    uint32_t nd[3] = {tri[3 * k], tri[3 * k + 1], tri[3 * k + 2]};
    int pos[9];
    bool any = false;
#pragma unroll
    for (int q = 0; q < 9; q++) {
        pos[q] = smap[9 * k + q];
        any |= pos[q] >= 0;
    }
    if (!any) return;
    Element<R> e;
    compute_element<R>(xyz, nd[0], nd[1], nd[2], e);
    R vel[3] = {vx, vy, vz}, tv[3];
    tangent_velocity<R>(e, vel, tv);
    R ms = O::div(e.jdet, (R)24.0);
    R as = O::div(e.jdet, (R)6.0);
    bool interior[3];
#pragma unroll
    for (int i = 0; i < 3; i++) interior[i] = !on_boundary[nd[i]];
    R adv[3];
#pragma unroll
    for (int j = 0; j < 3; j++) adv[j] = O::mul(as, edot3(tv, e.G[j]));
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) {
            int p = pos[3 * i + j];
            if (p < 0) continue;
            atomicAdd(&Kv[p], O::mul(e.area, edot3(e.G[i], e.G[j])));
            atomicAdd(&Mv[p], O::mul(ms, (R)(i == j ? 2 : 1)));
            if (interior[i] && interior[j]) {
                atomicAdd(&Av[p], adv[j]);
                atomicAdd(&AvT[p], adv[i]);  // slot (i, j) of Adv^T holds Adv(j, i) = adv[i]
            }
        }
}

// K1 (ordered mode): one thread per row; walks the row's incident corners in ascending element order and
// accumulates the matching local-matrix row into the CSR row -- exactly the order in which Eigen's
// setFromTriplets/collapseDuplicates sums duplicate triplets, so the result is deterministic and (for the
// same real type) bit-identical to the CPU oracle.  No atomics, no zero-fill needed.
template <typename R>
__global__ void __launch_bounds__(kBlock) assemble_ordered(const float* __restrict__ xyz, const uint32_t* __restrict__ tri, int64_t n,
                                                            const uint8_t* __restrict__ on_boundary, const int* __restrict__ colmap,
                                                            const int* __restrict__ cptr, const int* __restrict__ corner,
                                                            const int* __restrict__ rowptr, const int* __restrict__ col, R vx, R vy,
                                                            R vz, R* __restrict__ Mv, R* __restrict__ Kv, R* __restrict__ Av,
                                                            R* __restrict__ AvT) {
    using O = Ops<R>;
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    const int rb = rowptr[r], re = rowptr[r + 1];
    for (int p = rb; p < re; p++) {
        Mv[p] = (R)0;
        Kv[p] = (R)0;
        Av[p] = (R)0;
        AvT[p] = (R)0;
    }
    R vel[3] = {vx, vy, vz};
    for (int q = cptr[r]; q < cptr[r + 1]; q++) {
        int cr = corner[q];
        int64_t k = cr / 3;
        int i = cr - 3 * (int)k;
        uint32_t nd[3] = {tri[3 * k], tri[3 * k + 1], tri[3 * k + 2]};
        Element<R> e;
        compute_element<R>(xyz, nd[0], nd[1], nd[2], e);
        R tv[3];
        tangent_velocity<R>(e, vel, tv);
        R ms = O::div(e.jdet, (R)24.0);
        R as = O::div(e.jdet, (R)6.0);
        bool inti = !on_boundary[nd[i]];
#pragma unroll
        for (int j = 0; j < 3; j++) {
            int mj = colmap[nd[j]];
            if (mj < 0) continue;
            int p = rb;
            while (p < re && col[p] != mj) p++;  // rows are short (about 7 entries)
            if (p == re) continue;
            // first contribution is stored, later ones added (0 + x is exact, so plain accumulation is identical)
            Kv[p] = O::add(Kv[p], O::mul(e.area, edot3(e.G[i], e.G[j])));
            Mv[p] = O::add(Mv[p], O::mul(ms, (R)(i == j ? 2 : 1)));
            if (inti && !on_boundary[nd[j]]) {
                Av[p] = O::add(Av[p], O::mul(as, edot3(tv, e.G[j])));
                // Adv^T(row, col) = Adv(col, row): the same triangles in the same ascending order contribute
                // Adv_loc(j, i), which depends only on its column index i (FEMContext.cpp:192)
                AvT[p] = O::add(AvT[p], O::mul(as, edot3(tv, e.G[i])));
            }
        }
    }
}

}  // namespace feab
