// Shared device helpers for libfeab (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <cfloat>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

namespace feab {

constexpr int kBlock = 256;  // threads per CTA for every streaming kernel

// FEAB_DEBUG=1: synchronise and report after instrumented launches (debug aid, off by default)
static inline bool dbg_on() {
    static int v = -1;
    if (v < 0) v = getenv("FEAB_DEBUG") ? 1 : 0;
    return v == 1;
}
#define DBG(stream, ...)                                  \
    do {                                                  \
        if (dbg_on()) {                                   \
            cudaError_t e_ = cudaStreamSynchronize(stream); \
            fprintf(stderr, "[feab] " __VA_ARGS__);       \
            fprintf(stderr, " -> %s\n", cudaGetErrorString(e_)); \
            fflush(stderr);                               \
        }                                                 \
    } while (0)

// Correctly-rounded primitives that nvcc never contracts into FMAs, so that the per-element arithmetic
// follows the reference's operation order exactly (the oracle is built with -ffp-contract=off).
template <typename R>
struct Ops;
template <>
struct Ops<float> {
    static __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
    static __device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
    static __device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
    static __device__ __forceinline__ float div(float a, float b) { return __fdiv_rn(a, b); }
    static __device__ __forceinline__ float sqrt(float a) { return __fsqrt_rn(a); }
    static __device__ __forceinline__ float fma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
    static __host__ __device__ constexpr float eps() { return FLT_EPSILON; }
    static __host__ __device__ constexpr float tiny() { return FLT_MIN; }
    static __host__ __device__ constexpr float huge() { return FLT_MAX; }
};
template <>
struct Ops<double> {
    static __device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
    static __device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
    static __device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
    static __device__ __forceinline__ double div(double a, double b) { return __ddiv_rn(a, b); }
    static __device__ __forceinline__ double sqrt(double a) { return __dsqrt_rn(a); }
    static __device__ __forceinline__ double fma(double a, double b, double c) { return __fma_rn(a, b, c); }
    static __host__ __device__ constexpr double eps() { return DBL_EPSILON; }
    static __host__ __device__ constexpr double tiny() { return DBL_MIN; }
    static __host__ __device__ constexpr double huge() { return DBL_MAX; }
};

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}

// Fixed-order block sum (warp shuffles, then one warp over the per-warp results).  Result valid in every
// thread.  `sm` must hold kBlock/32 + 1 elements.  Contains two __syncthreads.
template <typename T>
__device__ __forceinline__ T block_sum(T v, T* sm) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    v = warp_sum(v);
    if (lane == 0) sm[w] = v;
    __syncthreads();
    if (w == 0) {
        T t = lane < (int)(blockDim.x >> 5) ? sm[lane] : T(0);
        t = warp_sum(t);
        if (lane == 0) sm[kBlock / 32] = t;
    }
    __syncthreads();
    T r = sm[kBlock / 32];
    return r;
}

// Same for a CTA of NW warps (the TMA kernels use 512-thread CTAs).  `sm` must hold NW + 1 elements.
template <typename T, int NW>
__device__ __forceinline__ T block_sum_n(T v, T* sm /* NW + 1 */) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    v = warp_sum(v);
    if (lane == 0) sm[w] = v;
    __syncthreads();
    if (w == 0) {
        T t = lane < NW ? sm[lane] : T(0);
        t = warp_sum(t);
        if (lane == 0) sm[NW] = t;
    }
    __syncthreads();
    return sm[NW];
}

// K sums at once, each added in exactly the order block_sum_n uses (so the bits are the same), with two CTA barriers
// instead of 2 K (+ one trailing barrier so that `sm` can be reused at once).  `sm` must hold (NW + 1) * K elements.
template <typename T, int K, int NW>
__device__ __forceinline__ void block_sum_vec(T (&v)[K], T* sm) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < K; k++) {
        v[k] = warp_sum(v[k]);
        if (lane == 0) sm[w * K + k] = v[k];
    }
    __syncthreads();
    if (w == 0) {
#pragma unroll
        for (int k = 0; k < K; k++) {
            T t = lane < NW ? sm[lane * K + k] : T(0);
            t = warp_sum(t);
            if (lane == 0) sm[NW * K + k] = t;
        }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < K; k++) v[k] = sm[NW * K + k];
    __syncthreads();
}

// Deterministic grid-wide sum of `count` per-CTA partials, computed redundantly (and identically) by every
// CTA of the consuming kernel: no atomics, no "last block" fence, no host round trip.
template <typename T>
__device__ __forceinline__ T sum_partials(const T* __restrict__ part, int count, T* sm) {
    T s = T(0);
    for (int i = threadIdx.x; i < count; i += blockDim.x) s += part[i];
    T r = block_sum(s, sm);
    __syncthreads();  // sm is reused by the caller
    return r;
}

}  // namespace feab
