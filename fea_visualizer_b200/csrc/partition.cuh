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

}  // namespace feab
