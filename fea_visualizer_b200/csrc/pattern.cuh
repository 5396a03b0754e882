// Symbolic phase on the device: unknown numbering (idx_map), CSR sparsity pattern of the P1 system and the
// per-triangle scatter map.  Replaces FEMContext::update_boundary_conditions (src/FEM/FEMContext.cpp:224-240)
// and the pattern half of Eigen's setFromTriplets (lib/eigen/Eigen/src/SparseCore/SparseMatrix.h:1125-1173):
// row i's column set = {i} U {unknown neighbours of i}, ascending, structural zeros kept.
#pragma once
#include "common.cuh"

namespace feab {

// ------------------------------------------------------------------------------------------------
// Exclusive prefix sum of int32 (three passes, 4096 items per CTA).  out has n + 1 entries; out[n] = total.
// ------------------------------------------------------------------------------------------------
constexpr int kScanItems = 4;                      // per thread
constexpr int kScanTile = 1024 * kScanItems;       // per CTA

__device__ __forceinline__ int block_excl_scan_1024(int v, int* sm, int* total) {
    // exclusive scan of one int per thread over a 1024-thread CTA
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
        int s = sm[lane];
        int si = s;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, si, o);
            if (lane >= o) si += t;
        }
        sm[lane] = si - s;  // exclusive warp offsets
        if (lane == 31) sm[32] = si;
    }
    __syncthreads();
    int r = inc - v + sm[w];
    *total = sm[32];
    __syncthreads();
    return r;
}

__global__ void __launch_bounds__(1024) scan_block_sums(const int* __restrict__ in, int64_t n, long long* __restrict__ bsum) {
    __shared__ int sm[33];
    int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
    int s = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; i++)
        if (base + i < n) s += in[base + i];
    int tot;
    block_excl_scan_1024(s, sm, &tot);
    if (threadIdx.x == 0) bsum[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(1024) scan_block_offsets(long long* __restrict__ bsum, int nblocks, long long* __restrict__ total) {
    // single CTA: in-place exclusive scan of the block sums (serial over chunks of 1024)
    __shared__ int sm[33];
    __shared__ long long carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < nblocks; base += 1024) {
        int i = base + threadIdx.x;
        int v = i < nblocks ? (int)bsum[i] : 0;
        int tot;
        int ex = block_excl_scan_1024(v, sm, &tot);
        long long c = carry;
        if (i < nblocks) bsum[i] = c + ex;
        __syncthreads();
        if (threadIdx.x == 0) carry = c + tot;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = carry;
}

__global__ void __launch_bounds__(1024) scan_apply(const int* __restrict__ in, int64_t n, const long long* __restrict__ bsum,
                                                   const long long* __restrict__ total, int* __restrict__ out) {
    __shared__ int sm[33];
    int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
    int v[kScanItems];
    int s = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; i++) {
        v[i] = base + i < n ? in[base + i] : 0;
        s += v[i];
    }
    int tot;
    int ex = block_excl_scan_1024(s, sm, &tot);
    long long off = bsum[blockIdx.x] + ex;
#pragma unroll
    for (int i = 0; i < kScanItems; i++) {
        if (base + i < n) out[base + i] = (int)off;
        off += v[i];
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) out[n] = (int)(*total);
}

// ------------------------------------------------------------------------------------------------
// idx_map
// ------------------------------------------------------------------------------------------------
__global__ void unknown_flags(const uint8_t* __restrict__ on_boundary, int64_t V, int bc, int* __restrict__ flag) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < V) flag[i] = (bc == 1 /*Neumann*/ || !on_boundary[i]) ? 1 : 0;
}
__global__ void make_idx_map(const int* __restrict__ flag, const int* __restrict__ excl, int64_t V, int* __restrict__ idx_map,
                             int* __restrict__ node_of) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < V) {
        if (flag[i]) {
            idx_map[i] = excl[i];
            node_of[excl[i]] = (int)i;
        } else {
            idx_map[i] = -1;
        }
    }
}

__global__ void validate_indices(const uint32_t* __restrict__ idx, int64_t count, uint32_t limit, int* __restrict__ bad) {
    bool b = false;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) b = b || idx[i] >= limit;
    if (b) *bad = 1;
}

// nodes WITHOUT an unknown (Dirichlet boundary nodes), ascending: node i is the (i - excl[i])-th of them
__global__ void make_dirichlet_list(const int* __restrict__ flag, const int* __restrict__ excl, int64_t V, int* __restrict__ dir_nodes) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < V && !flag[i]) dir_nodes[i - excl[i]] = (int)i;
}

// ------------------------------------------------------------------------------------------------
// Row candidates: every triangle corner whose node owns a row contributes its unknown partner nodes.
// An entry (i, j) of M/K exists iff some triangle holds both nodes and both pass the filter
// (src/FEM/FEMContext.cpp:99, :136): with Dirichlet, boundary nodes have idx_map = -1, with Neumann all pass.
// rowmap[node] = row owned by this handle (or -1), colmap[node] = column id (or -1).  On a single GPU both are
// idx_map; on a partition (partition.cuh) rows are the owned unknowns and columns are owned + ghost unknowns.
// ------------------------------------------------------------------------------------------------
__global__ void count_row_candidates(const uint32_t* __restrict__ tri, int64_t T, const int* __restrict__ rowmap,
                                     const int* __restrict__ colmap, int* __restrict__ cnt) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= T) return;
    int rw[3], m[3];
#pragma unroll
    for (int i = 0; i < 3; i++) {
        uint32_t nd = tri[3 * k + i];
        rw[i] = rowmap[nd];
        m[i] = colmap[nd];
    }
#pragma unroll
    for (int i = 0; i < 3; i++) {
        if (rw[i] < 0) continue;
        int c = 1;  // the diagonal (i, i) is emitted by every triangle that holds an unknown node
        c += (m[(i + 1) % 3] >= 0) + (m[(i + 2) % 3] >= 0);
        atomicAdd(&cnt[rw[i]], c);
    }
}
__global__ void fill_row_candidates(const uint32_t* __restrict__ tri, int64_t T, const int* __restrict__ rowmap,
                                    const int* __restrict__ colmap, const int* __restrict__ cand_ptr, int* __restrict__ fill,
                                    int* __restrict__ cand) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= T) return;
    int rw[3], m[3];
#pragma unroll
    for (int i = 0; i < 3; i++) {
        uint32_t nd = tri[3 * k + i];
        rw[i] = rowmap[nd];
        m[i] = colmap[nd];
    }
#pragma unroll
    for (int i = 0; i < 3; i++) {
        if (rw[i] < 0) continue;
        int c = 1 + (m[(i + 1) % 3] >= 0) + (m[(i + 2) % 3] >= 0);
        int slot = cand_ptr[rw[i]] + atomicAdd(&fill[rw[i]], c);
        cand[slot++] = m[i];
        if (m[(i + 1) % 3] >= 0) cand[slot++] = m[(i + 1) % 3];
        if (m[(i + 2) % 3] >= 0) cand[slot++] = m[(i + 2) % 3];
    }
}
// One thread per row: sort the candidate list in place, drop duplicates, report the row length.
// The sorted unique set does not depend on the (atomic) fill order, so the pattern is deterministic.
__global__ void sort_unique_rows(const int* __restrict__ cand_ptr, int* __restrict__ cand, int64_t n, int* __restrict__ rowlen) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    int b = cand_ptr[r], e = cand_ptr[r + 1];
    for (int i = b + 1; i < e; i++) {
        int v = cand[i];
        int q = i;
        while (q > b && cand[q - 1] > v) {
            cand[q] = cand[q - 1];
            q--;
        }
        cand[q] = v;
    }
    int w = b;
    for (int i = b; i < e; i++) {
        int v = cand[i];
        if (w == b || cand[w - 1] != v) cand[w++] = v;
    }
    rowlen[r] = w - b;
}
__global__ void compact_rows(const int* __restrict__ cand_ptr, const int* __restrict__ cand, const int* __restrict__ rowptr,
                             int64_t n, int* __restrict__ col, int* __restrict__ diag, int* __restrict__ maxlen) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int len = 0;
    if (r < n) {
        int b = cand_ptr[r], o = rowptr[r];
        len = rowptr[r + 1] - o;
        int d = -1;
        for (int i = 0; i < len; i++) {
            int c = cand[b + i];
            col[o + i] = c;
            if (c == (int)r) d = o + i;
        }
        diag[r] = d;
    }
    // FEMContext::compute_max_row_nonzeros (src/FEM/FEMContext.cpp:207-219)
    int m = len;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_down_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m > 0) atomicMax(maxlen, m);
}

__device__ __forceinline__ int find_in_row(const int* __restrict__ col, int b, int e, int target) {
    // binary search in the ascending column list [b, e)
    int lo = b, hi = e - 1;
    while (lo <= hi) {
        int mid = (lo + hi) >> 1;
        int c = col[mid];
        if (c == target) return mid;
        if (c < target)
            lo = mid + 1;
        else
            hi = mid - 1;
    }
    return -1;
}

// The precomputed CSR scatter map: for triangle k and local (i, j) the position of (row_i, col_j) in the
// CSR value arrays, or -1 when the entry is filtered out (or the row belongs to another partition).
__global__ void build_scatter_map(const uint32_t* __restrict__ tri, int64_t T, const int* __restrict__ rowmap,
                                  const int* __restrict__ colmap, const int* __restrict__ rowptr, const int* __restrict__ col,
                                  int* __restrict__ smap) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= T) return;
    int rw[3], m[3];
#pragma unroll
    for (int i = 0; i < 3; i++) {
        uint32_t nd = tri[3 * k + i];
        rw[i] = rowmap[nd];
        m[i] = colmap[nd];
    }
#pragma unroll
    for (int i = 0; i < 3; i++) {
        int b = 0, e = 0;
        if (rw[i] >= 0) {
            b = rowptr[rw[i]];
            e = rowptr[rw[i] + 1];
        }
#pragma unroll
        for (int j = 0; j < 3; j++) {
            int pos = -1;
            if (rw[i] >= 0 && m[j] >= 0) pos = find_in_row(col, b, e, m[j]);
            smap[9 * k + 3 * i + j] = pos;
        }
    }
}

// Whether the advection matrix holds the entry: its filter drops every boundary node regardless of the
// boundary condition (src/FEM/FEMContext.cpp:188).  node_of_col[c] = mesh node of column c (rows are the
// first columns).
__global__ void mark_adv_present(const int* __restrict__ rowptr, const int* __restrict__ col, int64_t n,
                                 const int* __restrict__ node_of_col, const uint8_t* __restrict__ on_boundary,
                                 uint8_t* __restrict__ adv_present) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    bool rb = on_boundary[node_of_col[r]] != 0;
    for (int k = rowptr[r]; k < rowptr[r + 1]; k++) adv_present[k] = (!rb && !on_boundary[node_of_col[col[k]]]) ? 1 : 0;
}

// ------------------------------------------------------------------------------------------------
// Ordered adjacency for the deterministic assembly: per row, the incident triangle corners (3k + i) in
// ascending order -- the order in which Eigen's collapseDuplicates sums the triplets of an entry.
// ------------------------------------------------------------------------------------------------
__global__ void count_row_corners(const uint32_t* __restrict__ tri, int64_t T, const int* __restrict__ rowmap, int* __restrict__ cnt) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= T) return;
#pragma unroll
    for (int i = 0; i < 3; i++) {
        int m = rowmap[tri[3 * k + i]];
        if (m >= 0) atomicAdd(&cnt[m], 1);
    }
}
__global__ void fill_row_corners(const uint32_t* __restrict__ tri, int64_t T, const int* __restrict__ rowmap,
                                 const int* __restrict__ cptr, int* __restrict__ fill, int* __restrict__ corner) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= T) return;
#pragma unroll
    for (int i = 0; i < 3; i++) {
        int m = rowmap[tri[3 * k + i]];
        if (m >= 0) corner[cptr[m] + atomicAdd(&fill[m], 1)] = (int)(3 * k + i);
    }
}
__global__ void sort_row_corners(const int* __restrict__ cptr, int* __restrict__ corner, int64_t n) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    int b = cptr[r], e = cptr[r + 1];
    for (int i = b + 1; i < e; i++) {
        int v = corner[i];
        int q = i;
        while (q > b && corner[q - 1] > v) {
            corner[q] = corner[q - 1];
            q--;
        }
        corner[q] = v;
    }
}

}  // namespace feab
