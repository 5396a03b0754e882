// Device kernels of the Surface-side rows: boundary-edge detection, smooth normals, colour map,
// and the threshold-clipping .ply builder.  Everything here is HBM-/atomic-bound integer and fp32 streaming work.
// fp32 arithmetic that must match the reference bit for bit uses __fmul_rn/__fadd_rn/... so nvcc cannot contract it
// (the reference is built for baseline x86-64: no FMA).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace feas {

constexpr int kBlock = 256;

// ------------------------------------------------------------------------------------------------------------------
// exclusive prefix sum, IN = uint8_t or int32_t, 3 passes, 4096 items per CTA; out has n + 1 entries
// ------------------------------------------------------------------------------------------------------------------
constexpr int kScanItems = 4;
constexpr int kScanTile = 1024 * kScanItems;

__device__ __forceinline__ int cta_excl_scan(int v, int* sm /*33*/, int* total) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) sm[w] = inc;
    __syncthreads();
    if (w == 0) {
        int s = sm[lane], si = s;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, si, o);
            if (lane >= o) si += t;
        }
        sm[lane] = si - s;
        if (lane == 31) sm[32] = si;
    }
    __syncthreads();
    int r = inc - v + sm[w];
    *total = sm[32];
    __syncthreads();
    return r;
}

template <typename IN>
__global__ void __launch_bounds__(1024) scan_tile_sums(const IN* __restrict__ in, int64_t n, long long* __restrict__ tsum) {
    __shared__ int sm[33];
    int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
    int s = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; i++)
        if (base + i < n) s += (int)in[base + i];
    int tot;
    cta_excl_scan(s, sm, &tot);
    if (threadIdx.x == 0) tsum[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(1024) scan_tile_offsets(long long* __restrict__ tsum, int ntiles, long long* __restrict__ total) {
    __shared__ int sm[33];
    __shared__ long long carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < ntiles; base += 1024) {
        int i = base + threadIdx.x;
        int v = i < ntiles ? (int)tsum[i] : 0;
        int tot;
        int ex = cta_excl_scan(v, sm, &tot);
        long long c = carry;
        if (i < ntiles) tsum[i] = c + ex;
        __syncthreads();
        if (threadIdx.x == 0) carry = c + tot;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = carry;
}

template <typename IN>
__global__ void __launch_bounds__(1024) scan_tile_apply(const IN* __restrict__ in, int64_t n, const long long* __restrict__ tsum,
                                                        const long long* __restrict__ total, int* __restrict__ out) {
    __shared__ int sm[33];
    int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
    int v[kScanItems];
    int s = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; i++) {
        v[i] = base + i < n ? (int)in[base + i] : 0;
        s += v[i];
    }
    int tot;
    int ex = cta_excl_scan(s, sm, &tot);
    long long off = tsum[blockIdx.x] + ex;
#pragma unroll
    for (int i = 0; i < kScanItems; i++) {
        if (base + i < n) out[base + i] = (int)off;
        off += v[i];
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) out[n] = (int)(*total);
}

// ------------------------------------------------------------------------------------------------------------------
// N2: boundary-edge detection (src/Utils/Surface.cpp:103-122).  The reference counts edges[a][b]++ and edges[b][a]++
// per triangle edge in a map of maps and flags both end points of every pair whose count is exactly 1.  Here: one
// open-addressing hash table keyed by the undirected pair (min, max); a degenerate edge (a == b) counts 2 like the
// reference's double increment of edges[a][a], so it can never look like a boundary edge.
// ------------------------------------------------------------------------------------------------------------------
constexpr unsigned long long kEmptyKey = ~0ull;

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
    x ^= x >> 33;
    x *= 0xff51afd7ed558ccdull;
    x ^= x >> 33;
    x *= 0xc4ceb9fe1a85ec53ull;
    x ^= x >> 33;
    return x;
}

__global__ void edge_count(const uint32_t* __restrict__ tri, int64_t T, unsigned long long* __restrict__ keys, int* __restrict__ cnt,
                           uint64_t mask) {
    int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= 3 * T) return;
    int64_t t = e / 3;
    int j = (int)(e - 3 * t);
    uint32_t a = tri[3 * t + j], b = tri[3 * t + (j + 1) % 3];
    uint32_t lo = a < b ? a : b, hi = a < b ? b : a;
    unsigned long long key = ((unsigned long long)lo << 32) | hi;
    int inc = a == b ? 2 : 1;
    uint64_t h = mix64(key) & mask;
    while (true) {
        unsigned long long cur = keys[h];
        if (cur == kEmptyKey) {
            unsigned long long prev = atomicCAS(&keys[h], kEmptyKey, key);
            cur = prev == kEmptyKey ? key : prev;
        }
        if (cur == key) {
            atomicAdd(&cnt[h], inc);
            return;
        }
        h = (h + 1) & mask;
    }
}

__global__ void edge_mark(const unsigned long long* __restrict__ keys, const int* __restrict__ cnt, uint64_t slots,
                          uint8_t* __restrict__ onb) {
    uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= slots) return;
    if (cnt[s] == 1) {
        unsigned long long k = keys[s];
        onb[(uint32_t)(k >> 32)] = 1;
        onb[(uint32_t)k] = 1;
    }
}

__global__ void count_flags(const uint8_t* __restrict__ f, int64_t n, unsigned long long* __restrict__ out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int v = i < n ? (f[i] != 0) : 0;
    unsigned b = __ballot_sync(0xffffffffu, v);
    if ((threadIdx.x & 31) == 0 && b) atomicAdd(out, (unsigned long long)__popc(b));
}

__global__ void fill_normals_up(float* __restrict__ nrm, int64_t V) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < V) {
        nrm[3 * i] = 0.0f;
        nrm[3 * i + 1] = 1.0f;
        nrm[3 * i + 2] = 0.0f;
    }
}

// ------------------------------------------------------------------------------------------------------------------
// ColorMap::get_color (src/Utils/ColorMap.cpp:12-14): Horner, every product and sum rounded separately
// ------------------------------------------------------------------------------------------------------------------
struct ColorCoeffs {
    float c[21];
};

__device__ __forceinline__ void color_of(const ColorCoeffs& cm, float t, float* rgb) {
#pragma unroll
    for (int k = 0; k < 3; k++) {
        float acc = __fmul_rn(t, cm.c[18 + k]);           // t * c6
#pragma unroll
        for (int d = 5; d >= 1; d--) acc = __fmul_rn(t, __fadd_rn(cm.c[3 * d + k], acc));   // t * (c_d + ...)
        rgb[k] = __fadd_rn(cm.c[k], acc);
    }
}

__global__ void color_map_kernel(ColorCoeffs cm, const float* __restrict__ t, int64_t n, float* __restrict__ rgb) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float c[3];
    color_of(cm, t[i], c);
    rgb[3 * i] = c[0];
    rgb[3 * i + 1] = c[1];
    rgb[3 * i + 2] = c[2];
}

// ------------------------------------------------------------------------------------------------------------------
// N3: smooth_normals.glsl:30-57.  One thread per triangle, native float atomics instead of the shader's CAS loop.
// ------------------------------------------------------------------------------------------------------------------
__global__ void smooth_normals_kernel(const float* __restrict__ xyz, const float* __restrict__ nrm, const float* __restrict__ val,
                                      const uint32_t* __restrict__ tri, int64_t T, float ve, float* __restrict__ out) {
    int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= T) return;
    uint32_t id[3] = {tri[3 * t], tri[3 * t + 1], tri[3 * t + 2]};
    float p[3][3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        float s = fminf(1.0f, __fmul_rn(ve, fmaxf(0.0f, val[id[k]])));
#pragma unroll
        for (int d = 0; d < 3; d++) p[k][d] = __fadd_rn(xyz[3 * (int64_t)id[k] + d], __fmul_rn(s, nrm[3 * (int64_t)id[k] + d]));
    }
    float e1[3], e2[3];
#pragma unroll
    for (int d = 0; d < 3; d++) {
        e1[d] = __fsub_rn(p[1][d], p[0][d]);
        e2[d] = __fsub_rn(p[2][d], p[0][d]);
    }
    float n[3];
    n[0] = __fsub_rn(__fmul_rn(e1[1], e2[2]), __fmul_rn(e2[1], e1[2]));
    n[1] = __fsub_rn(__fmul_rn(e1[2], e2[0]), __fmul_rn(e2[2], e1[0]));
    n[2] = __fsub_rn(__fmul_rn(e1[0], e2[1]), __fmul_rn(e2[0], e1[1]));
#pragma unroll
    for (int d = 0; d < 3; d++)
#pragma unroll
        for (int k = 0; k < 3; k++) atomicAdd(&out[3 * (int64_t)id[k] + d], n[d]);
}

}  // namespace feas
