// N1: Surface::export_to_ply (src/Utils/Surface.cpp:138-285) as a device pipeline.
//
// The reference walks the triangles in order; each one emits up to 6 (position, value) pairs through
// add_clipped_position -- a first-occurrence-wins insert into an unordered_map keyed by the fp32 position -- and up to
// 4 triangles whose corners are map look-ups of those positions.  Output vertex i is the i-th DISTINCT position in
// emission order.  On the device:
//   ply_classify   per triangle: how many candidates / output triangles it emits             (2 bytes out)
//   scan x2        exclusive offsets (emission order == triangle order, then slot order)
//   ply_emit       candidates (position, value) and output triangles as candidate ids
//   ply_insert     hash table keyed by position: every slot ends up holding the SMALLEST candidate id of its key
//   ply_resolve    candidate -> representative; flag[c] = (representative == c)
//   scan           rank of every first occurrence = output vertex index
//   ply_compact    output vertices (+ ColorMap::get_color), ply_remap: output triangles
// Equality is float == on all three components like glm's operator== (so -0 == +0; the first occurrence's bits are the
// ones written).  NaN positions (never equal to themselves in the reference) are not reproduced: they collapse by bit
// pattern here.  They only arise from non-finite input values.
#pragma once
#include "surface_kernels.cuh"

namespace feas {

struct PlyArgs {
    const float* xyz;
    const float* nrm;
    const float* val;
    const uint32_t* tri;
    int64_t T;
    float ve;         // vertex_extrusion as passed (the projected copies use it raw, Surface.cpp:195,232-236,263)
    float ve_pos;     // std::max(0.0f, vertex_extrusion) (extruded_vertex, Surface.cpp:166-168)
    float thr;
    int mesh_type;    // 0 Open, 1 Closed, 2 Mirrored
};

// (values[i] - threshold) >= 1e-9: float difference compared as double (Surface.cpp:151-153)
__device__ __forceinline__ bool above(float v, float thr) { return (double)__fsub_rn(v, thr) >= 1e-9; }

__device__ __forceinline__ void ply_counts(int count, int mesh_type, int& ncand, int& ntri) {
    const bool open = mesh_type == 0;
    switch (count) {
        case 1: ncand = open ? 3 : 4; ntri = open ? 1 : 2; break;
        case 2: ncand = open ? 4 : 6; ntri = open ? 2 : 4; break;
        case 3: ncand = open ? 3 : 6; ntri = open ? 1 : 2; break;
        default: ncand = 0; ntri = 0; break;
    }
}

__global__ void ply_classify(PlyArgs a, uint8_t* __restrict__ ncand, uint8_t* __restrict__ ntri) {
    int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= a.T) return;
    int c = 0;
#pragma unroll
    for (int k = 0; k < 3; k++) c += above(a.val[a.tri[3 * t + k]], a.thr) ? 1 : 0;
    int nc, nt;
    ply_counts(c, a.mesh_type, nc, nt);
    ncand[t] = (uint8_t)nc;
    ntri[t] = (uint8_t)nt;
}

struct V3 {
    float x, y, z;
};
// scalar * vec3 + vec3 in glm: every component product and sum rounded separately
__device__ __forceinline__ V3 axpy(float s, const float* n, const float* v) {
    return V3{__fadd_rn(__fmul_rn(s, n[0]), v[0]), __fadd_rn(__fmul_rn(s, n[1]), v[1]), __fadd_rn(__fmul_rn(s, n[2]), v[2])};
}
// glm::mix(x, y, a) = x * (1 - a) + y * a  (lib/glm/glm/detail/func_common.inl:140)
__device__ __forceinline__ V3 mix3(V3 x, V3 y, float a) {
    float b = __fsub_rn(1.0f, a);
    return V3{__fadd_rn(__fmul_rn(x.x, b), __fmul_rn(y.x, a)), __fadd_rn(__fmul_rn(x.y, b), __fmul_rn(y.y, a)),
              __fadd_rn(__fmul_rn(x.z, b), __fmul_rn(y.z, a))};
}

__global__ void ply_emit(PlyArgs a, const int* __restrict__ coff, const int* __restrict__ toff, float* __restrict__ cpos,
                         float* __restrict__ cval, int* __restrict__ tslot) {
    int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= a.T) return;
    const int c0 = coff[t];
    if (coff[t + 1] == c0) return;
    const int t0 = toff[t];
    uint32_t id[3];
    float v[3], P[3][3], N[3][3];
    bool ab[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        id[k] = a.tri[3 * t + k];
        v[k] = a.val[id[k]];
        ab[k] = above(v[k], a.thr);
#pragma unroll
        for (int d = 0; d < 3; d++) {
            P[k][d] = a.xyz[3 * (int64_t)id[k] + d];
            N[k][d] = a.nrm[3 * (int64_t)id[k] + d];
        }
    }
    const int count = (int)ab[0] + (int)ab[1] + (int)ab[2];
    const bool open = a.mesh_type == 0, mirrored = a.mesh_type == 2;
    V3 ext[3];
#pragma unroll
    for (int k = 0; k < 3; k++) ext[k] = axpy(__fmul_rn(v[k], a.ve_pos), N[k], P[k]);   // extruded_vertex
    // projected copy of vertex k (Surface.cpp:195-196, 232-236, 263-264)
    auto projected = [&](int k) {
        float s = mirrored ? __fsub_rn(a.thr, __fmul_rn(v[k], a.ve)) : __fmul_rn(a.thr, a.ve);
        return axpy(s, N[k], P[k]);
    };
    auto put = [&](int slot, V3 p, float value) {
        cpos[3 * (int64_t)(c0 + slot)] = p.x;
        cpos[3 * (int64_t)(c0 + slot) + 1] = p.y;
        cpos[3 * (int64_t)(c0 + slot) + 2] = p.z;
        cval[c0 + slot] = value;
    };
    auto face = [&](int m, int s0, int s1, int s2) {
        tslot[3 * (int64_t)(t0 + m)] = c0 + s0;
        tslot[3 * (int64_t)(t0 + m) + 1] = c0 + s1;
        tslot[3 * (int64_t)(t0 + m) + 2] = c0 + s2;
    };
    if (count == 1) {
        const int ai = ab[0] ? 0 : ab[1] ? 1 : 2;
        for (int i = 0; i < 3; i++) {
            if (i != ai) {
                float w = __fdiv_rn(__fsub_rn(a.thr, v[i]), __fsub_rn(v[ai], v[i]));
                put(i, mix3(ext[i], ext[ai], w), a.thr);
            } else {
                put(i, ext[i], v[ai]);
            }
        }
        face(0, 0, 1, 2);
        if (!open) {
            put(3, projected(ai), a.thr);
            face(1, ai == 0 ? 3 : 0, ai == 1 ? 3 : 1, ai == 2 ? 3 : 2);
        }
    } else if (count == 2) {
        const int a1 = ab[0] ? 0 : 1;
        const int a2 = (ab[0] && ab[1]) ? 1 : 2;
        const int bi = 3 - (a1 + a2);
        float w1 = __fdiv_rn(__fsub_rn(a.thr, v[bi]), __fsub_rn(v[a1], v[bi]));
        float w2 = __fdiv_rn(__fsub_rn(a.thr, v[bi]), __fsub_rn(v[a2], v[bi]));
        put(0, ext[a1], v[a1]);
        put(1, ext[a2], v[a2]);
        put(2, mix3(ext[bi], ext[a1], w1), a.thr);
        put(3, mix3(ext[bi], ext[a2], w2), a.thr);
        face(0, 2, 0, 3);
        face(1, 3, 0, 1);
        if (!open) {
            put(4, projected(a1), v[a1]);
            put(5, projected(a2), v[a2]);
            face(2, 2, 4, 3);
            face(3, 3, 4, 5);
        }
    } else {  // 3
        for (int i = 0; i < 3; i++) put(i, ext[i], v[i]);
        face(0, 0, 1, 2);
        if (!open) {
            for (int i = 0; i < 3; i++) put(3 + i, projected(i), a.thr);
            face(1, 3, 4, 5);
        }
    }
}

// position key: +0 and -0 compare equal
__device__ __forceinline__ uint32_t canon(float f) { return f == 0.0f ? 0u : __float_as_uint(f); }
__device__ __forceinline__ uint64_t key_hash(uint32_t x, uint32_t y, uint32_t z) {
    return mix64(((uint64_t)x << 32 | y) ^ mix64((uint64_t)z * 0x9e3779b97f4a7c15ull + 0x7f4a7c15ull));
}

__global__ void ply_insert(const float* __restrict__ cpos, int64_t nc, int* __restrict__ rep, uint64_t mask, int* __restrict__ cslot) {
    int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= nc) return;
    const uint32_t kx = canon(cpos[3 * c]), ky = canon(cpos[3 * c + 1]), kz = canon(cpos[3 * c + 2]);
    uint64_t h = key_hash(kx, ky, kz) & mask;
    while (true) {
        int cur = *(volatile int*)&rep[h];
        if (cur < 0) {
            int prev = atomicCAS(&rep[h], -1, (int)c);
            if (prev < 0) break;   // claimed: this slot's key class is ours from now on
            cur = prev;
        }
        // the slot's occupant may be replaced by a smaller id of the SAME key at any time; its key never changes
        if (canon(cpos[3 * (int64_t)cur]) == kx && canon(cpos[3 * (int64_t)cur + 1]) == ky && canon(cpos[3 * (int64_t)cur + 2]) == kz) {
            atomicMin(&rep[h], (int)c);
            break;
        }
        h = (h + 1) & mask;
    }
    cslot[c] = (int)h;
}

// cslot[c]: table slot -> representative (smallest id with the same position); first[c] = 1 for representatives
__global__ void ply_resolve(const int* __restrict__ rep, int64_t nc, int* __restrict__ cslot, uint8_t* __restrict__ first) {
    int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= nc) return;
    int r = rep[cslot[c]];
    cslot[c] = r;
    first[c] = r == (int)c ? 1 : 0;
}

__global__ void ply_compact(const float* __restrict__ cpos, const float* __restrict__ cval, const uint8_t* __restrict__ first,
                            const int* __restrict__ rank, int64_t nc, ColorCoeffs cm, float* __restrict__ opos, float* __restrict__ oval,
                            float* __restrict__ ocol) {
    int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= nc || !first[c]) return;
    const int64_t o = rank[c];
    const float value = cval[c];
    float col[3];
    color_of(cm, value, col);
#pragma unroll
    for (int d = 0; d < 3; d++) {
        opos[3 * o + d] = cpos[3 * c + d];
        ocol[3 * o + d] = col[d];
    }
    oval[o] = value;
}

__global__ void ply_remap(const int* __restrict__ tslot, int64_t n3, const int* __restrict__ crep, const int* __restrict__ rank,
                          uint32_t* __restrict__ otri) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n3) otri[i] = (uint32_t)rank[crep[tslot[i]]];
}

}  // namespace feas
