// Row partition of the unknowns over `world` ranks and the local numbering each rank derives
// from it.  Every rank holds the whole mesh and the same part[] array, so all of this is computed locally and
// identically -- no communication at setup.
//   local rows    = owned unknowns: interior ones first (no neighbour on another rank), boundary ones last,
//                   each group in ascending global order (the SpMV starts on the interior while halos fly);
//   local columns = [rows | padding to a 128-byte line | ghosts grouped by owner rank, ascending global id].
// mask[u] has bit q set iff unknown u shares a triangle with an unknown owned by rank q != owner(u): u is a
// ghost on q, and a "send row" of its owner towards q.  Because both sides enumerate these sets in ascending
// global order, the sender's i-th send row is the receiver's i-th ghost of that owner.
#pragma once
#include <algorithm>
#include <cstring>
#include <numeric>
#include <vector>

#include "comm.cuh"

namespace feab {

__global__ void ghost_mask_build(const uint32_t* __restrict__ tri, int64_t T, const int* __restrict__ idx_map,
                                 const int* __restrict__ part, unsigned* __restrict__ mask) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= T) return;
    int m[3], p[3];
#pragma unroll
    for (int i = 0; i < 3; i++) {
        uint32_t nd = tri[3 * k + i];
        m[i] = idx_map[nd];
        p[i] = part[nd];
    }
#pragma unroll
    for (int i = 0; i < 3; i++) {
        if (m[i] < 0) continue;
        unsigned bits = 0;
#pragma unroll
        for (int j = 0; j < 3; j++)
            if (j != i && m[j] >= 0 && p[j] != p[i]) bits |= 1u << p[j];
        if (bits) atomicOr(&mask[m[i]], bits);
    }
}

// sel 0: owner && mask == 0;  1: owner && mask != 0;  2: owner && bit set
__global__ void flag_unknowns(const int* __restrict__ node_of, const int* __restrict__ part, const unsigned* __restrict__ mask,
                              int64_t nglob, int owner, int sel, int bit, int* __restrict__ flag) {
    int64_t u = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= nglob) return;
    bool f = part[node_of[u]] == owner;
    const unsigned mk = mask[u];
    if (sel == 0) f = f && mk == 0;
    if (sel == 1) f = f && mk != 0;
    if (sel == 2) f = f && ((mk >> bit) & 1u);
    flag[u] = f ? 1 : 0;
}
__global__ void assign_ids(const int* __restrict__ flag, const int* __restrict__ excl, int64_t nglob, int base, int* __restrict__ id) {
    int64_t u = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (u < nglob && flag[u]) id[u] = base + excl[u];
}
__global__ void gather_list(const int* __restrict__ flag, const int* __restrict__ excl, int64_t nglob, const int* __restrict__ id,
                            const int* __restrict__ node_of, int* __restrict__ rows, int* __restrict__ nodes) {
    int64_t u = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (u < nglob && flag[u]) {
        rows[excl[u]] = id[u];
        nodes[excl[u]] = node_of[u];
    }
}
// cnt[o * kMaxRanks + q] = number of unknowns owned by o that are ghosts on q
__global__ void count_ghost_pairs(const int* __restrict__ node_of, const int* __restrict__ part, const unsigned* __restrict__ mask,
                                  int64_t nglob, int* __restrict__ cnt) {
    __shared__ int h[kMaxRanks * kMaxRanks];
    for (int i = threadIdx.x; i < kMaxRanks * kMaxRanks; i += blockDim.x) h[i] = 0;
    __syncthreads();
    for (int64_t u = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; u < nglob; u += (int64_t)gridDim.x * blockDim.x) {
        unsigned mk = mask[u];
        if (!mk) continue;
        const int o = part[node_of[u]];
        while (mk) {
            const int q = __ffs(mk) - 1;
            mk &= mk - 1;
            atomicAdd(&h[o * kMaxRanks + q], 1);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < kMaxRanks * kMaxRanks; i += blockDim.x)
        if (h[i]) atomicAdd(&cnt[i], h[i]);
}
// rowmap / colmap per mesh node and the inverse maps per local column
__global__ void make_local_maps(const int* __restrict__ idx_map, const int* __restrict__ lrow, const int* __restrict__ lcol, int64_t V,
                                int* __restrict__ rowmap, int* __restrict__ colmap, int* __restrict__ node_of_col,
                                int* __restrict__ glob_of_col) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= V) return;
    const int u = idx_map[i];
    int r = -1, c = -1;
    if (u >= 0) {
        r = lrow[u];
        c = lcol[u];
        if (c >= 0) {
            node_of_col[c] = (int)i;
            glob_of_col[c] = u;
        }
    }
    rowmap[i] = r;
    colmap[i] = c;
}

// ------------------------------------------------------------------------------------------------
// Host: recursive coordinate bisection of the vertices into `parts` compact pieces of (nearly) equal size.
// Deterministic (ties broken by vertex index), so every rank computes the same partition.
// ------------------------------------------------------------------------------------------------
inline void rcb_split(const float* xyz, int* idx, int64_t count, int parts, int first_part, int* part) {
    if (parts <= 1 || count == 0) {
        for (int64_t i = 0; i < count; i++) part[idx[i]] = first_part;
        return;
    }
    float lo[3] = {1e30f, 1e30f, 1e30f}, hi[3] = {-1e30f, -1e30f, -1e30f};
    for (int64_t i = 0; i < count; i++)
        for (int d = 0; d < 3; d++) {
            const float c = xyz[3 * (int64_t)idx[i] + d];
            lo[d] = std::min(lo[d], c);
            hi[d] = std::max(hi[d], c);
        }
    int axis = 0;
    for (int d = 1; d < 3; d++)
        if (hi[d] - lo[d] > hi[axis] - lo[axis]) axis = d;
    const int pl = parts / 2;
    const int64_t nl = count * pl / parts;
    std::nth_element(idx, idx + nl, idx + count, [&](int a, int b) {
        const float ca = xyz[3 * (int64_t)a + axis], cb = xyz[3 * (int64_t)b + axis];
        return ca < cb || (ca == cb && a < b);
    });
    rcb_split(xyz, idx, nl, pl, first_part, part);
    rcb_split(xyz, idx + nl, count - nl, parts - pl, first_part + pl, part);
}
inline void rcb_partition(int64_t V, const float* xyz, int parts, int* part) {
    std::vector<int> idx((size_t)V);
    std::iota(idx.begin(), idx.end(), 0);
    rcb_split(xyz, idx.data(), V, parts, 0, part);
}


// ------------------------------------------------------------------------------------------------
// The same bisection ON THE DEVICE (the host version costs 0.7 s at 16 M vertices and every rank would run it): per level
// and per segment, bounding box -> widest axis -> radix select (4 passes over an order-preserving 32-bit image of the
// coordinate) of the nl-th smallest coordinate -> ties at that coordinate go left in ascending vertex order (one exclusive
// scan) -- i.e. exactly the set std::nth_element leaves on the left under the comparator above, so rcb_partition_device ==
// rcb_partition bit for bit (tested), and all ranks obtain the same array without talking to each other.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned rcb_key(float f) {
    f = f + 0.0f;  // -0 -> +0: the host comparator treats them as equal
    const unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float rcb_unkey(unsigned k) {
    const unsigned u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
    float f;
#ifdef __CUDA_ARCH__
    f = __uint_as_float(u);
#else
    memcpy(&f, &u, 4);
#endif
    return f;
}
struct RcbLevel {                  // up to kMaxRanks / 2 segments are split per level
    int nseg;
    int label[kMaxRanks];          // part[] value of the segment's vertices
    int axis[kMaxRanks];
    unsigned prefix[kMaxRanks];    // radix select: digits fixed so far (high bits)
    unsigned key[kMaxRanks];       // the selected coordinate key
    int take[kMaxRanks];           // number of vertices with exactly that key that go left
    int right_label[kMaxRanks];
    int seg_of_label[kMaxRanks];   // label -> segment index or -1
};
__global__ void rcb_bbox(const float* __restrict__ xyz, const int* __restrict__ part, int64_t V, RcbLevel lv, unsigned* __restrict__ box /* nseg x 6 */) {
    __shared__ unsigned s_box[kMaxRanks * 6];
    for (int i = threadIdx.x; i < lv.nseg * 6; i += blockDim.x) s_box[i] = (i % 6) < 3 ? 0xffffffffu : 0u;
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < V; i += (int64_t)gridDim.x * blockDim.x) {
        const int sg = lv.seg_of_label[part[i]];
        if (sg < 0) continue;
#pragma unroll
        for (int d = 0; d < 3; d++) {
            const unsigned k = rcb_key(xyz[3 * i + d]);
            atomicMin(&s_box[sg * 6 + d], k);
            atomicMax(&s_box[sg * 6 + 3 + d], k);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < lv.nseg * 6; i += blockDim.x) {
        if ((i % 6) < 3)
            atomicMin(&box[i], s_box[i]);
        else
            atomicMax(&box[i], s_box[i]);
    }
}
// histogram of the 8-bit digit at `shift` among the vertices whose higher digits equal the segment's prefix
__global__ void rcb_hist(const float* __restrict__ xyz, const int* __restrict__ part, int64_t V, RcbLevel lv, int shift,
                         unsigned* __restrict__ hist /* nseg x 256 */) {
    __shared__ unsigned s_h[kMaxRanks / 2 * 256];
    for (int i = threadIdx.x; i < lv.nseg * 256; i += blockDim.x) s_h[i] = 0;
    __syncthreads();
    const unsigned himask = shift == 24 ? 0u : (0xffffffffu << (shift + 8));
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < V; i += (int64_t)gridDim.x * blockDim.x) {
        const int sg = lv.seg_of_label[part[i]];
        if (sg < 0) continue;
        const unsigned k = rcb_key(xyz[3 * i + lv.axis[sg]]);
        if ((k & himask) != (lv.prefix[sg] & himask)) continue;
        atomicAdd(&s_h[sg * 256 + ((k >> shift) & 255u)], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < lv.nseg * 256; i += blockDim.x)
        if (s_h[i]) atomicAdd(&hist[i], s_h[i]);
}
__global__ void rcb_tie_flags(const float* __restrict__ xyz, const int* __restrict__ part, int64_t V, RcbLevel lv, int sg, int* __restrict__ flag) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= V) return;
    flag[i] = (lv.seg_of_label[part[i]] == sg && rcb_key(xyz[3 * i + lv.axis[sg]]) == lv.key[sg]) ? 1 : 0;
}
// vertices of segment sg: key < key* stay (left), key > key* move to the right label, ties by their rank among the ties
__global__ void rcb_assign(const float* __restrict__ xyz, int* __restrict__ part, int64_t V, RcbLevel lv, int sg, const int* __restrict__ tie_rank) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= V) return;
    if (lv.seg_of_label[part[i]] != sg) return;
    const unsigned k = rcb_key(xyz[3 * i + lv.axis[sg]]);
    const bool right = k > lv.key[sg] || (k == lv.key[sg] && tie_rank[i] >= lv.take[sg]);
    if (right) part[i] = lv.right_label[sg];
}

}  // namespace feab
