// libfeab.so: feas_detect_boundary_noalloc (include/feab_surface.h) -- boundary-edge detection for callers that are not
// allowed to allocate GPU memory, and that do not want this library to allocate any on their behalf.
//
// Who that is.  feas_detect_boundary (surface.cu) cudaMallocs a fresh edge hash table on every call and frees it again.
// That is fine for a tool that owns the GPU, and wrong for a host program that does not:
//   * a viewer / render loop: cudaMalloc and cudaFree synchronise the whole device, so a surface reload in the middle
//     of an animation stalls every frame in flight;
//   * a plugin living inside an application that handed ALL device memory to its own pool at start-up (a framework's
//     caching allocator, a fixed per-tenant budget): a raw cudaMalloc from a library either fails or takes memory the
//     pool's accounting never sees;
//   * code that runs where the CUDA runtime's allocator must not be entered at all (inside another library's
//     callback, under a memory-budget audit that forbids allocation after initialisation).
// Such a caller brings its own device buffers for the triangles (input) and the flags (output).  The only other memory
// the algorithm needs -- the open-addressing table that counts how many triangles use each edge -- lives in a
// file-scope __device__ array below: reserved when the module is loaded, never grown, never freed, so the entry point
// calls no cudaMalloc, no cudaFree and no library routine that allocates.
//
// The price, stated once:
//   * every GPU on which this library's module gets loaded gives up kScratchBytes (24 MiB + 256 B) until the process
//     exits, used or not;
//   * the table is fixed: meshes with more than FEAS_NOALLOC_MAX_TRIANGLES triangles are refused (FEAB_EINVAL), use
//     feas_detect_boundary for those;
//   * there is ONE table per device and every call uses it.  It still holds the previous call's edges when the next
//     call starts, so each call wipes the part it is going to use first.  Two calls whose kernels ran at the same time
//     on one GPU would count into each other's slots and both results would be garbage; the entry point therefore
//     holds a process-wide lock and waits for its stream before it returns.  Do not expect concurrency from it.
//
// Set-up.  In host code `g_edge_scratch` is only a handle for the runtime, not an address a kernel could use.  The
// first call on a device (with that device current) asks cudaGetSymbolAddress for the real device address and keeps
// it; later calls reuse it.  That first call is the one that can trigger module loading, so make it early -- in
// particular before the process starts any stream capture.
//
// The reference has no counterpart of this entry point; the algorithm is the one of Surface::init_from_obj
// (src/Utils/Surface.cpp:103-127) exactly as feas_detect_boundary implements it: an edge used by exactly one
// triangle is a boundary edge, both its end points are boundary vertices, a degenerate edge (a == a) counts twice.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <mutex>

#include "../../include/feab_surface.h"

namespace feas_noalloc {

constexpr int kBlock = 256;
constexpr uint64_t kMaxSlots = 1ull << 21;            // table slots; 6 T <= slots keeps the load factor <= 0.5
constexpr uint64_t kMinSlots = 1024;
constexpr size_t kKeysBytes = kMaxSlots * sizeof(unsigned long long);   // 16 MiB
constexpr size_t kCntBytes = kMaxSlots * sizeof(int);                   //  8 MiB
constexpr size_t kTailBytes = 256;                    // [0] number of boundary vertices, [1] "index out of range" flag
constexpr size_t kScratchBytes = kKeysBytes + kCntBytes + kTailBytes;
constexpr int kMaxDevices = 64;
constexpr unsigned long long kEmptyKey = ~0ull;

static_assert(6ull * FEAS_NOALLOC_MAX_TRIANGLES <= kMaxSlots, "header limit does not fit the table");
static_assert(6ull * (FEAS_NOALLOC_MAX_TRIANGLES + 1) > kMaxSlots, "header limit is smaller than the table allows");

// The scratch.  Layout: keys[kMaxSlots] (u64) | cnt[kMaxSlots] (i32) | tail (2 x u64, padded).
static __device__ __align__(256) unsigned char g_edge_scratch[kScratchBytes];

// Device address of g_edge_scratch per CUDA ordinal, filled by the first call on that device.
static void* g_scratch_addr[kMaxDevices] = {};
static std::mutex g_lock;

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
    x ^= x >> 33;
    x *= 0xff51afd7ed558ccdull;
    x ^= x >> 33;
    x *= 0xc4ceb9fe1a85ec53ull;
    x ^= x >> 33;
    return x;
}

// One thread per triangle edge: count the undirected pair (min, max) in the table.  Unlike surface.cu's edge_count this
// variant cannot rely on a host-side range check of the triangle array (the array is the caller's device memory), so
// an index >= V raises tail[1] and the edge is skipped instead of being marked out of bounds later.
__global__ void edge_count_noalloc(const uint32_t* __restrict__ tri, int64_t T, uint32_t V, unsigned long long* __restrict__ keys,
                                   int* __restrict__ cnt, uint64_t mask, unsigned long long* __restrict__ tail) {
    int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= 3 * T) return;
    int64_t t = e / 3;
    int j = (int)(e - 3 * t);
    uint32_t a = tri[3 * t + j], b = tri[3 * t + (j + 1) % 3];
    if (a >= V || b >= V) {
        tail[1] = 1;
        return;
    }
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

__global__ void edge_mark_noalloc(const unsigned long long* __restrict__ keys, const int* __restrict__ cnt, uint64_t slots,
                                  uint8_t* __restrict__ onb) {
    uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= slots) return;
    if (cnt[s] == 1) {
        unsigned long long k = keys[s];
        onb[(uint32_t)(k >> 32)] = 1;
        onb[(uint32_t)k] = 1;
    }
}

__global__ void count_flags_noalloc(const uint8_t* __restrict__ f, int64_t n, unsigned long long* __restrict__ tail) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int v = i < n ? (f[i] != 0) : 0;
    unsigned b = __ballot_sync(0xffffffffu, v);
    if ((threadIdx.x & 31) == 0 && b) atomicAdd(&tail[0], (unsigned long long)__popc(b));
}

static inline unsigned cdiv(int64_t a, int64_t b) { return (unsigned)((a + b - 1) / b); }

struct DeviceGuard {   // puts the caller's current device back, whatever happens
    int prev = -1;
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

static int fail(char* errbuf, int64_t len, int code, const char* what, cudaError_t ce = cudaSuccess) {
    if (errbuf && len > 0) {
        if (ce != cudaSuccess) std::snprintf(errbuf, (size_t)len, "feas_detect_boundary_noalloc: %s: %s", what, cudaGetErrorString(ce));
        else std::snprintf(errbuf, (size_t)len, "feas_detect_boundary_noalloc: %s", what);
    }
    return ce != cudaSuccess ? (int)ce : code;
}

}  // namespace feas_noalloc

extern "C" int feas_detect_boundary_noalloc(int device, void* stream_, int64_t V, int64_t T, const uint32_t* tri_device,
                                            uint8_t* on_boundary_device, int64_t* num_boundary_points, char* errbuf, int64_t errbuf_len) {
    using namespace feas_noalloc;
    if (errbuf && errbuf_len > 0) errbuf[0] = 0;
    if (V <= 0 || V >= (int64_t)1 << 31 || T < 0 || (T > 0 && !tri_device) || !on_boundary_device)
        return fail(errbuf, errbuf_len, FEAB_EINVAL, "bad arguments (need 0 < num_vertices < 2^31, num_triangles >= 0, non-null device pointers)");
    if (T > (int64_t)FEAS_NOALLOC_MAX_TRIANGLES) {
        char msg[200];
        std::snprintf(msg, sizeof msg, "%lld triangles exceed the fixed table (FEAS_NOALLOC_MAX_TRIANGLES = %lld); use feas_detect_boundary",
                      (long long)T, (long long)FEAS_NOALLOC_MAX_TRIANGLES);
        return fail(errbuf, errbuf_len, FEAB_EINVAL, msg);
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(errbuf, errbuf_len, FEAB_ENOGPU, "no CUDA device: libfeab has no CPU fallback");
    if (device < 0 || device >= ndev || device >= kMaxDevices) return fail(errbuf, errbuf_len, FEAB_EINVAL, "bad device ordinal");

    std::lock_guard<std::mutex> lock(g_lock);   // one table per device, shared by every call: never two at once
    DeviceGuard guard;
    cudaError_t ce;
    int prev = -1;
    if ((ce = cudaGetDevice(&prev)) != cudaSuccess) return fail(errbuf, errbuf_len, 0, "cudaGetDevice", ce);
    if (prev != device) {
        if ((ce = cudaSetDevice(device)) != cudaSuccess) return fail(errbuf, errbuf_len, 0, "cudaSetDevice", ce);
        guard.prev = prev;
    }
    // one-time set-up per device: the device address behind the host-side symbol
    if (!g_scratch_addr[device]) {
        void* p = nullptr;
        if ((ce = cudaGetSymbolAddress(&p, g_edge_scratch)) != cudaSuccess) return fail(errbuf, errbuf_len, 0, "cudaGetSymbolAddress", ce);
        if (!p) return fail(errbuf, errbuf_len, FEAB_EINVAL, "cudaGetSymbolAddress returned a null address");
        g_scratch_addr[device] = p;
    }
    unsigned char* base = (unsigned char*)g_scratch_addr[device];
    unsigned long long* keys = (unsigned long long*)base;
    int* cnt = (int*)(base + kKeysBytes);
    unsigned long long* tail = (unsigned long long*)(base + kKeysBytes + kCntBytes);

    uint64_t slots = kMinSlots;
    while (slots < (uint64_t)(6 * T)) slots <<= 1;   // <= kMaxSlots by the check above
    cudaStream_t stream = (cudaStream_t)stream_;

    // the table still holds the previous call's edges: wipe the part this call uses, the tail and the output
    if ((ce = cudaMemsetAsync(keys, 0xff, slots * sizeof(unsigned long long), stream)) != cudaSuccess) return fail(errbuf, errbuf_len, 0, "clear keys", ce);
    if ((ce = cudaMemsetAsync(cnt, 0, slots * sizeof(int), stream)) != cudaSuccess) return fail(errbuf, errbuf_len, 0, "clear counts", ce);
    if ((ce = cudaMemsetAsync(tail, 0, kTailBytes, stream)) != cudaSuccess) return fail(errbuf, errbuf_len, 0, "clear tail", ce);
    if ((ce = cudaMemsetAsync(on_boundary_device, 0, (size_t)V, stream)) != cudaSuccess) return fail(errbuf, errbuf_len, 0, "clear output", ce);
    if (T) {
        edge_count_noalloc<<<cdiv(3 * T, kBlock), kBlock, 0, stream>>>(tri_device, T, (uint32_t)V, keys, cnt, slots - 1, tail);
        edge_mark_noalloc<<<cdiv((int64_t)slots, kBlock), kBlock, 0, stream>>>(keys, cnt, slots, on_boundary_device);
    }
    count_flags_noalloc<<<cdiv(V, kBlock), kBlock, 0, stream>>>(on_boundary_device, V, tail);
    if ((ce = cudaGetLastError()) != cudaSuccess) return fail(errbuf, errbuf_len, 0, "kernel launch", ce);
    unsigned long long h_tail[2] = {0, 0};
    if ((ce = cudaMemcpyAsync(h_tail, tail, sizeof h_tail, cudaMemcpyDeviceToHost, stream)) != cudaSuccess) return fail(errbuf, errbuf_len, 0, "read back count", ce);
    if ((ce = cudaStreamSynchronize(stream)) != cudaSuccess) return fail(errbuf, errbuf_len, 0, "cudaStreamSynchronize", ce);
    if (h_tail[1]) return fail(errbuf, errbuf_len, FEAB_EINVAL, "a triangle index is >= num_vertices (on_boundary_device is incomplete)");
    if (num_boundary_points) *num_boundary_points = (int64_t)h_tail[0];
    return FEAB_OK;
}
