// N4: brush / picking.  Reference: BVHNode::ray_triangle_intersection (src/Utils/BVH.cpp:182-216) run
// over the leaves of a depth-limited median-split BVH (:8-40, :52-118), and Application::brush (src/Application.cpp:
// 707-727) which writes the brush value into the triangle corner closest to the hit point.
//
// On the device the hierarchy is not needed: one thread per triangle evaluates the reference's own plane / edge test in
// the reference's operation order (glm cross / dot / normalize with every product and sum rounded separately) and the
// nearest hit is an atomicMin on (distance bits << 32 | triangle index).  32 M triangles stream in ~50 us, less than one
// BVH descent on the host.  Differences from the reference, both only on exact ties or non-finite input:
//   * two triangles hit at exactly the same distance (a ray through a shared edge / vertex): the reference keeps the
//     first one in BVH traversal order, this kernel keeps the lowest triangle index;
//   * the reference's slab test can reject a leaf through a NaN when a ray direction component is 0 and the origin lies
//     exactly on that slab plane; no such rejection here.
#pragma once
#include "surface_kernels.cuh"

namespace feas {

struct F3 {
    float x, y, z;
};
__device__ __forceinline__ F3 sub3(F3 a, F3 b) { return F3{__fsub_rn(a.x, b.x), __fsub_rn(a.y, b.y), __fsub_rn(a.z, b.z)}; }
__device__ __forceinline__ F3 add3(F3 a, F3 b) { return F3{__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y), __fadd_rn(a.z, b.z)}; }
__device__ __forceinline__ F3 scale3(F3 a, float s) { return F3{__fmul_rn(a.x, s), __fmul_rn(a.y, s), __fmul_rn(a.z, s)}; }
// glm::dot(vec3): tmp = a * b; tmp.x + tmp.y + tmp.z   (lib/glm/glm/detail/func_geometric.inl)
__device__ __forceinline__ float dot3(F3 a, F3 b) {
    return __fadd_rn(__fadd_rn(__fmul_rn(a.x, b.x), __fmul_rn(a.y, b.y)), __fmul_rn(a.z, b.z));
}
// glm::cross: (x.y*y.z - y.y*x.z, x.z*y.x - y.z*x.x, x.x*y.y - y.x*x.y)
__device__ __forceinline__ F3 cross3(F3 x, F3 y) {
    return F3{__fsub_rn(__fmul_rn(x.y, y.z), __fmul_rn(y.y, x.z)), __fsub_rn(__fmul_rn(x.z, y.x), __fmul_rn(y.z, x.x)),
              __fsub_rn(__fmul_rn(x.x, y.y), __fmul_rn(y.x, x.y))};
}
// glm::normalize: v * inversesqrt(dot(v, v)), inversesqrt(x) = 1 / sqrt(x)
__device__ __forceinline__ F3 normalize3(F3 v) { return scale3(v, __fdiv_rn(1.0f, __fsqrt_rn(dot3(v, v)))); }
__device__ __forceinline__ F3 load3(const float* p, uint32_t i) { return F3{p[3 * (int64_t)i], p[3 * (int64_t)i + 1], p[3 * (int64_t)i + 2]}; }

constexpr unsigned long long kNoHit = ~0ull;

// best: (float bits of dist) << 32 | triangle index; dist > 0 so the bit pattern orders like the value
__global__ void pick_kernel(const float* __restrict__ xyz, const uint32_t* __restrict__ tri, int64_t T, F3 origin, F3 dir,
                            unsigned long long* __restrict__ best) {
    int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long mine = kNoHit;
    if (t < T) {
        const uint32_t id[3] = {tri[3 * t], tri[3 * t + 1], tri[3 * t + 2]};
        const F3 P[3] = {load3(xyz, id[0]), load3(xyz, id[1]), load3(xyz, id[2])};
        const float third = 3.0f;
        F3 s = add3(add3(P[0], P[1]), P[2]);
        const F3 center = F3{__fdiv_rn(s.x, third), __fdiv_rn(s.y, third), __fdiv_rn(s.z, third)};
        const F3 normal = normalize3(cross3(sub3(P[1], P[0]), sub3(P[2], P[0])));
        const float denom = dot3(dir, normal);
        if ((double)fabsf(denom) > 1e-6) {
            const float dist = __fdiv_rn(-dot3(sub3(origin, center), normal), denom);
            const F3 point = add3(origin, scale3(dir, dist));
            int pos = 0, neg = 0;
#pragma unroll
            for (int j = 0; j < 3; j++) {
                const F3 p1 = P[j], p2 = P[(j + 1) % 3];
                const F3 perp = cross3(normal, sub3(p2, p1));
                if (dot3(perp, sub3(point, p1)) < 0.0f) neg++;
                else pos++;
            }
            if ((pos == 3 || neg == 3) && dist > 0.0f) mine = ((unsigned long long)__float_as_uint(dist) << 32) | (unsigned long long)(uint32_t)t;
        }
    }
    // warp minimum first: one atomic per warp that has a hit
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        unsigned long long other = __shfl_down_sync(0xffffffffu, mine, o);
        mine = other < mine ? other : mine;
    }
    if ((threadIdx.x & 31) == 0 && mine != kNoHit) atomicMin(best, mine);
}

// Application::brush (src/Application.cpp:711-724): the corner of the hit triangle closest to the hit point (strictly
// positive distance, first minimum wins) receives `value`.  One thread.  out[0] = node or -1, out[1] = triangle or -1.
__global__ void brush_apply(const float* __restrict__ xyz, const uint32_t* __restrict__ tri, F3 origin, F3 dir,
                            const unsigned long long* __restrict__ best, float value, float* __restrict__ values, int* __restrict__ out,
                            float* __restrict__ out_point) {
    const unsigned long long b = *best;
    out[0] = out[1] = -1;
    if (b == kNoHit) return;
    const uint32_t t = (uint32_t)b;
    const float dist = __uint_as_float((uint32_t)(b >> 32));
    const F3 point = add3(origin, scale3(dir, dist));
    out[1] = (int)t;
    out_point[0] = point.x;
    out_point[1] = point.y;
    out_point[2] = point.z;
    out_point[3] = dist;
    int idx = -1;
    float mind = 0.0f;
    for (int i = 0; i < 3; i++) {
        const F3 d = sub3(load3(xyz, tri[3 * (int64_t)t + i]), point);
        const float dd = __fsqrt_rn(dot3(d, d));      // glm::distance = length(p0 - p1) = sqrt(dot)
        if ((idx == -1 || dd < mind) && dd > 0.0f) {
            mind = dd;
            idx = i;
        }
    }
    if (idx != -1) {
        const uint32_t node = tri[3 * (int64_t)t + idx];
        if (values) values[node] = value;
        out[0] = (int)node;
    }
}

}  // namespace feas
