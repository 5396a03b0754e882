// Packed column indices for the TMA-fed SpMV (K3 / the persistent PCG kernel): 2 bytes per entry instead of 4.
//
// The CSR column array stays what it is (int32, bit-exact against Eigen's innerIndexPtr, returned by feab_get_csr and
// read by every other kernel).  For the 512-row tiles the SpMV streams, a second, derived copy is kept:
//   tile t (rows r0 .. r0+511, entries [tb, te), tba = tb & ~3 like the value stream)
//     codes : one uint16 per entry of [tba, te) (count rounded up to 4, bytes rounded up to 16)
//               code <  kEscBase : column = r0 - kBias + code         (the column lies within +-32 K rows of the tile)
//               code >= kEscBase : column = overflow[code - kEscBase]  (a far column: icosphere face borders, ghosts)
//     overflow : int32 list, <= kMaxEsc entries, bytes rounded up to 16
//   cpoff[t] : offset of the tile's block in 16-byte units (exclusive scan); cpoff[t+1] == cpoff[t] marks a tile that
//              is not packed (more than kMaxEsc far columns, or more entries than the stage holds) -> int32 stream.
// One bulk-TMA copy brings codes + overflow of a tile into the stage's column slot.  Per row of a closed triangle mesh
// this moves 14 B instead of 28 B: 104 -> 90 B/row/SpMV in fp64.
#pragma once
#include "common.cuh"

namespace feab {

constexpr int kPackRows = 512;
constexpr int kEscBase = 0xFF00;
constexpr int kMaxEsc = 0x100;
constexpr int kBias = 0x7F80;

__device__ __forceinline__ int pack_code_bytes(int cnt_al) { return (cnt_al * 2 + 15) & ~15; }

// one CTA per tile; units[t] = size of the tile's block in 16-byte units, 0 = not packed
__global__ void __launch_bounds__(256) colpack_count(const int* __restrict__ rowptr, const int* __restrict__ col, int64_t n, int num_tiles,
                                                     int cap_entries, int* __restrict__ units) {
    const int tile = blockIdx.x;
    if (tile >= num_tiles) return;
    __shared__ int s_esc;
    if (threadIdx.x == 0) s_esc = 0;
    __syncthreads();
    const int64_t r0 = (int64_t)tile * kPackRows;
    const int rows = (int)min((int64_t)kPackRows, n - r0);
    const int tb = rowptr[r0], te = rowptr[r0 + rows];
    const int tba = tb & ~3;
    int mine = 0;
    for (int k = tb + threadIdx.x; k < te; k += blockDim.x) {
        const long long d = (long long)col[k] - r0 + kBias;
        if (d < 0 || d >= kEscBase) mine++;
    }
    if (mine) atomicAdd(&s_esc, mine);
    __syncthreads();
    if (threadIdx.x == 0) {
        const int cnt_al = ((te - tba) + 3) & ~3;
        const int nesc = s_esc;
        const bool ok = nesc <= kMaxEsc && (te - tba) <= cap_entries && te > tb;
        units[tile] = ok ? (pack_code_bytes(cnt_al) + ((nesc * 4 + 15) & ~15)) / 16 : 0;
    }
}

__global__ void __launch_bounds__(256) colpack_fill(const int* __restrict__ rowptr, const int* __restrict__ col, int64_t n, int num_tiles,
                                                    const int* __restrict__ cpoff, unsigned char* __restrict__ packed) {
    const int tile = blockIdx.x;
    if (tile >= num_tiles) return;
    const int u0 = cpoff[tile], u1 = cpoff[tile + 1];
    if (u1 == u0) return;
    __shared__ int s_slot;
    if (threadIdx.x == 0) s_slot = 0;
    __syncthreads();
    const int64_t r0 = (int64_t)tile * kPackRows;
    const int rows = (int)min((int64_t)kPackRows, n - r0);
    const int tb = rowptr[r0], te = rowptr[r0 + rows];
    const int tba = tb & ~3;
    const int cnt_al = ((te - tba) + 3) & ~3;
    unsigned short* codes = reinterpret_cast<unsigned short*>(packed + (size_t)u0 * 16);
    int* ovf = reinterpret_cast<int*>(packed + (size_t)u0 * 16 + pack_code_bytes(cnt_al));
    const int total16 = pack_code_bytes(cnt_al) / 2;
    for (int i = threadIdx.x; i < total16; i += blockDim.x) {
        const int k = tba + i;
        unsigned short code = 0;
        if (k >= tb && k < te) {
            const int c = col[k];
            const long long d = (long long)c - r0 + kBias;
            if (d < 0 || d >= kEscBase) {
                const int slot = atomicAdd(&s_slot, 1);
                ovf[slot] = c;
                code = (unsigned short)(kEscBase + slot);
            } else {
                code = (unsigned short)d;
            }
        }
        codes[i] = code;
    }
    __syncthreads();
    // zero the padding of the overflow list (deterministic bytes)
    const int nesc = s_slot;
    const int ovf_total = ((nesc * 4 + 15) & ~15) / 4;
    for (int i = nesc + threadIdx.x; i < ovf_total; i += blockDim.x) ovf[i] = 0;
}

// column of entry j of a packed tile (consumer side)
__device__ __forceinline__ int unpack_col(const unsigned short* __restrict__ codes, const int* __restrict__ ovf, int j, int colbase) {
    const int code = codes[j];
    return code < kEscBase ? colbase + code : ovf[code - kEscBase];
}

}  // namespace feab
