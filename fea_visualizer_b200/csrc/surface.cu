// libfeab.so, Surface side: the device-resident mirror of the reference's `Surface` (include/Utils/Surface.hpp:34-73)
// behind the C ABI of include/feab_surface.h.  Kernels: surface_kernels.cuh, ply_kernels.cuh.  Host text I/O:
// ply_writer.cpp.  No CPU fallback: without a CUDA device feas_create fails.
#include <algorithm>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "../../include/feab_surface.h"
#include "ply_kernels.cuh"
#include "pick_kernels.cuh"
#include "ply_writer.hpp"
#include "surface_kernels.cuh"

namespace feas {

struct Err {
    int code;
    std::string msg;
};
#define SCK(call)                                                                                             \
    do {                                                                                                      \
        cudaError_t e_ = (call);                                                                              \
        if (e_ != cudaSuccess) throw Err{(int)e_, std::string(#call) + ": " + cudaGetErrorString(e_)};        \
    } while (0)
#define SREQ(cond, code, msg) \
    do {                      \
        if (!(cond)) throw Err{(code), (msg)}; \
    } while (0)

template <typename T>
struct Buf {
    T* p = nullptr;
    size_t cap = 0;
    void alloc(size_t n) {
        n = std::max<size_t>(n, 1);
        if (n <= cap) return;
        release();
        cudaError_t e = cudaMalloc(&p, n * sizeof(T));
        if (e != cudaSuccess) {
            p = nullptr;
            throw Err{FEAB_ENOMEM, std::string("cudaMalloc: ") + cudaGetErrorString(e)};
        }
        cap = n;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    ~Buf() { release(); }
};

static inline int cdiv(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

class SurfaceDev {
public:
    explicit SurfaceDev(int device) : device_(device) {
        SCK(cudaSetDevice(device_));
        SCK(cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking));
        SCK(cudaEventCreate(&ev0_));
        SCK(cudaEventCreate(&ev1_));
        d_total_.alloc(2);
        // Viridis: src/Utils/AssetStorage.cpp:141-152 (double literals narrowed to float by glm::vec3)
        const double vir[21] = {0.274344, 0.004462, 0.331359, 0.108915, 1.397291, 1.388110, -0.319631, 0.243490, 0.156419,
                                -4.629188, -5.882803, -19.646115, 6.181719, 14.388598, 57.442181, 4.876952, -13.955112, -66.125783,
                                -5.513165, 4.709245, 26.582180};
        for (int i = 0; i < 21; i++) cm_.c[i] = (float)vir[i];
    }
    ~SurfaceDev() {
        cudaSetDevice(device_);
        cudaStreamSynchronize(stream_);
        cudaEventDestroy(ev0_);
        cudaEventDestroy(ev1_);
        cudaStreamDestroy(stream_);
    }
    void use() { SCK(cudaSetDevice(device_)); }

    void set_mesh(int64_t V, const float* xyz, const float* nrm, int64_t T, const uint32_t* tri) {
        SREQ(V > 0 && xyz && T >= 0 && (T == 0 || tri), FEAB_EINVAL, "set_mesh: bad arguments");
        SREQ(V < (int64_t)1 << 31 && 6 * T < (int64_t)1 << 31, FEAB_EINVAL, "set_mesh: mesh too large for 32-bit indices");
        use();
        V_ = V;
        T_ = T;
        d_xyz_.alloc(3 * V);
        d_nrm_.alloc(3 * V);
        d_val_.alloc(V);
        d_tri_.alloc(3 * std::max<int64_t>(T, 1));
        d_onb_.alloc(V);
        SCK(cudaMemcpyAsync(d_xyz_.p, xyz, 3 * V * sizeof(float), cudaMemcpyHostToDevice, stream_));
        if (nrm) {
            SCK(cudaMemcpyAsync(d_nrm_.p, nrm, 3 * V * sizeof(float), cudaMemcpyHostToDevice, stream_));
        } else {
            fill_normals_up<<<cdiv(V, kBlock), kBlock, 0, stream_>>>(d_nrm_.p, V);
            launches_++;
        }
        if (T) SCK(cudaMemcpyAsync(d_tri_.p, tri, 3 * T * sizeof(uint32_t), cudaMemcpyHostToDevice, stream_));
        SCK(cudaMemsetAsync(d_val_.p, 0, V * sizeof(float), stream_));
        SCK(cudaMemsetAsync(d_onb_.p, 0, V, stream_));
        SCK(cudaStreamSynchronize(stream_));
        // range check on the host copy: the reference would index out of bounds; we refuse
        for (int64_t i = 0; i < 3 * T; i++) SREQ(tri[i] < (uint64_t)V, FEAB_EINVAL, "set_mesh: triangle index out of range");
        nb_ = 0;
        ply_nv_ = ply_nt_ = 0;
    }

    void detect_boundary(uint8_t* out, int64_t* nb, int* closed) {
        need_mesh();
        use();
        uint64_t slots = 1024;
        while (slots < (uint64_t)(6 * T_)) slots <<= 1;   // <= 3T distinct pairs: load factor <= 0.5
        Buf<unsigned long long> keys;
        Buf<int> cnt;
        keys.alloc(slots);
        cnt.alloc(slots);
        SCK(cudaEventRecord(ev0_, stream_));
        SCK(cudaMemsetAsync(keys.p, 0xff, slots * sizeof(unsigned long long), stream_));
        SCK(cudaMemsetAsync(cnt.p, 0, slots * sizeof(int), stream_));
        SCK(cudaMemsetAsync(d_onb_.p, 0, V_, stream_));
        SCK(cudaMemsetAsync(d_total_.p, 0, sizeof(long long), stream_));
        if (T_) edge_count<<<cdiv(3 * T_, kBlock), kBlock, 0, stream_>>>(d_tri_.p, T_, keys.p, cnt.p, slots - 1);
        edge_mark<<<cdiv((int64_t)slots, kBlock), kBlock, 0, stream_>>>(keys.p, cnt.p, slots, d_onb_.p);
        count_flags<<<cdiv(V_, kBlock), kBlock, 0, stream_>>>(d_onb_.p, V_, (unsigned long long*)d_total_.p);
        launches_ += 3;
        SCK(cudaGetLastError());
        SCK(cudaEventRecord(ev1_, stream_));
        long long tot = 0;
        SCK(cudaMemcpyAsync(&tot, d_total_.p, sizeof(long long), cudaMemcpyDeviceToHost, stream_));
        if (out) SCK(cudaMemcpyAsync(out, d_onb_.p, V_, cudaMemcpyDeviceToHost, stream_));
        SCK(cudaStreamSynchronize(stream_));
        stamp();
        nb_ = tot;
        if (nb) *nb = tot;
        if (closed) *closed = tot == 0;
    }

    void load_obj_file(const char* path) {
        ObjMesh m;
        std::string e;
        SREQ(path && load_obj(path, m, e), FEAB_EINVAL, e.empty() ? "load_obj: bad path" : e);
        set_mesh((int64_t)m.xyz.size() / 3, m.xyz.data(), m.nrm.data(), (int64_t)m.tri.size() / 3, m.tri.data());
        detect_boundary(nullptr, nullptr, nullptr);
    }

    void get_mesh(float* xyz, float* nrm, uint32_t* tri, uint8_t* onb) {
        need_mesh();
        use();
        if (xyz) SCK(cudaMemcpyAsync(xyz, d_xyz_.p, 3 * V_ * sizeof(float), cudaMemcpyDeviceToHost, stream_));
        if (nrm) SCK(cudaMemcpyAsync(nrm, d_nrm_.p, 3 * V_ * sizeof(float), cudaMemcpyDeviceToHost, stream_));
        if (tri && T_) SCK(cudaMemcpyAsync(tri, d_tri_.p, 3 * T_ * sizeof(uint32_t), cudaMemcpyDeviceToHost, stream_));
        if (onb) SCK(cudaMemcpyAsync(onb, d_onb_.p, V_, cudaMemcpyDeviceToHost, stream_));
        SCK(cudaStreamSynchronize(stream_));
    }

    void set_values(const float* v, bool device) {
        need_mesh();
        SREQ(v, FEAB_EINVAL, "set_values: null");
        use();
        SCK(cudaMemcpyAsync(d_val_.p, v, V_ * sizeof(float), device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, stream_));
        SCK(cudaStreamSynchronize(stream_));
    }

    void get_values(float* v) {
        need_mesh();
        SREQ(v, FEAB_EINVAL, "get_values: null");
        use();
        SCK(cudaMemcpyAsync(v, d_val_.p, V_ * sizeof(float), cudaMemcpyDeviceToHost, stream_));
        SCK(cudaStreamSynchronize(stream_));
    }

    void set_color_map(const float* c21) {
        SREQ(c21, FEAB_EINVAL, "set_color_map: null");
        std::memcpy(cm_.c, c21, sizeof(cm_.c));
    }

    void color_map(int64_t n, const float* t, float* rgb) {
        SREQ(n >= 0 && (n == 0 || (t && rgb)), FEAB_EINVAL, "color_map: bad arguments");
        if (!n) return;
        use();
        Buf<float> dt, dc;
        dt.alloc(n);
        dc.alloc(3 * n);
        SCK(cudaMemcpyAsync(dt.p, t, n * sizeof(float), cudaMemcpyHostToDevice, stream_));
        color_map_kernel<<<cdiv(n, kBlock), kBlock, 0, stream_>>>(cm_, dt.p, n, dc.p);
        launches_++;
        SCK(cudaGetLastError());
        SCK(cudaMemcpyAsync(rgb, dc.p, 3 * n * sizeof(float), cudaMemcpyDeviceToHost, stream_));
        SCK(cudaStreamSynchronize(stream_));
    }

    void smooth_normals(float ve, float* out) {
        need_mesh();
        use();
        d_snrm_.alloc(3 * V_);
        SCK(cudaEventRecord(ev0_, stream_));
        SCK(cudaMemsetAsync(d_snrm_.p, 0, 3 * V_ * sizeof(float), stream_));   // glClearBufferData, Surface.cpp:301
        if (T_) smooth_normals_kernel<<<cdiv(T_, kBlock), kBlock, 0, stream_>>>(d_xyz_.p, d_nrm_.p, d_val_.p, d_tri_.p, T_, ve, d_snrm_.p);
        launches_++;
        SCK(cudaGetLastError());
        SCK(cudaEventRecord(ev1_, stream_));
        if (out) SCK(cudaMemcpyAsync(out, d_snrm_.p, 3 * V_ * sizeof(float), cudaMemcpyDeviceToHost, stream_));
        SCK(cudaStreamSynchronize(stream_));
        stamp();
    }

    // N4: returns the picked triangle (or -1); with do_brush also writes `value` into values[closest corner]
    void pick(const float* o, const float* d, bool do_brush, float value, int* tri_idx, int* node, float* point3, float* dist) {
        need_mesh();
        SREQ(o && d, FEAB_EINVAL, "pick: null ray");
        use();
        d_pick_.alloc(2);
        d_pick_out_.alloc(8);
        const F3 origin{o[0], o[1], o[2]}, dir{d[0], d[1], d[2]};
        SCK(cudaEventRecord(ev0_, stream_));
        SCK(cudaMemsetAsync(d_pick_.p, 0xff, sizeof(unsigned long long), stream_));
        if (T_) pick_kernel<<<cdiv(T_, kBlock), kBlock, 0, stream_>>>(d_xyz_.p, d_tri_.p, T_, origin, dir, d_pick_.p);
        brush_apply<<<1, 1, 0, stream_>>>(d_xyz_.p, d_tri_.p, origin, dir, d_pick_.p, value, do_brush ? d_val_.p : nullptr,
                                          (int*)d_pick_out_.p, d_pick_out_.p + 2);
        launches_ += 2;
        SCK(cudaGetLastError());
        SCK(cudaEventRecord(ev1_, stream_));
        float h[8];
        SCK(cudaMemcpyAsync(h, d_pick_out_.p, sizeof(h), cudaMemcpyDeviceToHost, stream_));
        SCK(cudaStreamSynchronize(stream_));
        stamp();
        int ih[2];
        std::memcpy(ih, h, sizeof(ih));
        if (node) *node = ih[0];
        if (tri_idx) *tri_idx = ih[1];
        if (ih[1] >= 0) {
            if (point3) std::memcpy(point3, h + 2, 3 * sizeof(float));
            if (dist) *dist = h[5];
        }
    }

    template <typename IN>
    long long scan(const IN* in, int64_t n, int* out) {
        int nt = std::max(1, cdiv(n, kScanTile));
        d_tsum_.alloc(nt);
        scan_tile_sums<IN><<<nt, 1024, 0, stream_>>>(in, n, d_tsum_.p);
        scan_tile_offsets<<<1, 1024, 0, stream_>>>(d_tsum_.p, nt, d_total_.p);
        scan_tile_apply<IN><<<nt, 1024, 0, stream_>>>(in, n, d_tsum_.p, d_total_.p, out);
        launches_ += 3;
        SCK(cudaGetLastError());
        long long tot = 0;
        SCK(cudaMemcpyAsync(&tot, d_total_.p, sizeof(long long), cudaMemcpyDeviceToHost, stream_));
        SCK(cudaStreamSynchronize(stream_));
        return tot;
    }

    void ply_build(float ve, float thr, int mesh_type, int64_t* nv, int64_t* nt) {
        need_mesh();
        SREQ(mesh_type >= 0 && mesh_type <= 2, FEAB_EINVAL, "ply_build: mesh_type must be 0, 1 or 2");
        use();
        const int64_t T = T_;
        PlyArgs a{d_xyz_.p, d_nrm_.p, d_val_.p, d_tri_.p, T, ve, (0.0f < ve) ? ve : 0.0f, thr, mesh_type};
        Buf<uint8_t> ncand, ntri, first;
        Buf<int> coff, toff, tslot, cslot, rep, rank;
        Buf<float> cpos, cval;
        ncand.alloc(T);
        ntri.alloc(T);
        coff.alloc(T + 1);
        toff.alloc(T + 1);
        SCK(cudaEventRecord(ev0_, stream_));
        const int gT = std::max(1, cdiv(T, kBlock));
        long long nc = 0, nto = 0;
        if (T) {
            ply_classify<<<gT, kBlock, 0, stream_>>>(a, ncand.p, ntri.p);
            launches_++;
            nc = scan<uint8_t>(ncand.p, T, coff.p);
            nto = scan<uint8_t>(ntri.p, T, toff.p);
        }
        long long nvo = 0;
        d_opos_.alloc(1);
        if (nc) {
            cpos.alloc(3 * nc);
            cval.alloc(nc);
            cslot.alloc(nc);
            first.alloc(nc);
            rank.alloc(nc + 1);
            tslot.alloc(3 * nto);
            uint64_t slots = 1024;
            while (slots < (uint64_t)(2 * nc)) slots <<= 1;
            rep.alloc(slots);
            SCK(cudaMemsetAsync(rep.p, 0xff, slots * sizeof(int), stream_));
            ply_emit<<<gT, kBlock, 0, stream_>>>(a, coff.p, toff.p, cpos.p, cval.p, tslot.p);
            const int gC = cdiv(nc, kBlock);
            ply_insert<<<gC, kBlock, 0, stream_>>>(cpos.p, nc, rep.p, slots - 1, cslot.p);
            ply_resolve<<<gC, kBlock, 0, stream_>>>(rep.p, nc, cslot.p, first.p);
            launches_ += 3;
            nvo = scan<uint8_t>(first.p, nc, rank.p);
            d_opos_.alloc(3 * nvo);
            d_oval_.alloc(nvo);
            d_ocol_.alloc(3 * nvo);
            d_otri_.alloc(3 * nto);
            ply_compact<<<gC, kBlock, 0, stream_>>>(cpos.p, cval.p, first.p, rank.p, nc, cm_, d_opos_.p, d_oval_.p, d_ocol_.p);
            ply_remap<<<cdiv(3 * nto, kBlock), kBlock, 0, stream_>>>(tslot.p, 3 * nto, cslot.p, rank.p, d_otri_.p);
            launches_ += 2;
            SCK(cudaGetLastError());
        }
        SCK(cudaEventRecord(ev1_, stream_));
        SCK(cudaStreamSynchronize(stream_));
        stamp();
        ply_nv_ = nvo;
        ply_nt_ = nc ? nto : 0;
        ply_candidates_ = nc;
        if (nv) *nv = ply_nv_;
        if (nt) *nt = ply_nt_;
    }

    void ply_fetch(float* pos, float* val, float* col, uint32_t* tri) {
        use();
        if (ply_nv_) {
            if (pos) SCK(cudaMemcpyAsync(pos, d_opos_.p, 3 * ply_nv_ * sizeof(float), cudaMemcpyDeviceToHost, stream_));
            if (val) SCK(cudaMemcpyAsync(val, d_oval_.p, ply_nv_ * sizeof(float), cudaMemcpyDeviceToHost, stream_));
            if (col) SCK(cudaMemcpyAsync(col, d_ocol_.p, 3 * ply_nv_ * sizeof(float), cudaMemcpyDeviceToHost, stream_));
        }
        if (ply_nt_ && tri) SCK(cudaMemcpyAsync(tri, d_otri_.p, 3 * ply_nt_ * sizeof(uint32_t), cudaMemcpyDeviceToHost, stream_));
        SCK(cudaStreamSynchronize(stream_));
    }

    void export_ply(const char* path, float ve, float thr, int mesh_type) {
        SREQ(path, FEAB_EINVAL, "export_ply: null path");
        int64_t nv, nt;
        ply_build(ve, thr, mesh_type, &nv, &nt);
        const double build_ms = last_ms_;
        std::vector<float> pos(3 * std::max<int64_t>(nv, 1)), col(3 * std::max<int64_t>(nv, 1));
        std::vector<uint32_t> tri(3 * std::max<int64_t>(nt, 1));
        ply_fetch(pos.data(), nullptr, col.data(), tri.data());
        std::string e;
        SREQ(write_ply(path, nv, pos.data(), col.data(), nt, tri.data(), e), FEAB_EINVAL, e);
        last_ms_ = build_ms;
    }

    int64_t V() const { return V_; }
    int64_t T() const { return T_; }
    int64_t nb() const { return nb_; }
    double last_ms() const { return last_ms_; }
    int64_t launches() const { return launches_; }

private:
    void need_mesh() const { SREQ(V_ > 0, FEAB_ENOMESH, "no mesh set"); }
    void stamp() {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, ev0_, ev1_) == cudaSuccess) last_ms_ = ms;
    }

    int device_;
    cudaStream_t stream_ = nullptr;
    cudaEvent_t ev0_ = nullptr, ev1_ = nullptr;
    int64_t V_ = 0, T_ = 0, nb_ = 0;
    Buf<float> d_xyz_, d_nrm_, d_val_, d_snrm_;
    Buf<uint32_t> d_tri_;
    Buf<uint8_t> d_onb_;
    Buf<long long> d_tsum_, d_total_;
    Buf<unsigned long long> d_pick_;
    Buf<float> d_pick_out_;
    ColorCoeffs cm_;
    // last ply_build result
    Buf<float> d_opos_, d_oval_, d_ocol_;
    Buf<uint32_t> d_otri_;
    int64_t ply_nv_ = 0, ply_nt_ = 0, ply_candidates_ = 0;
    double last_ms_ = 0.0;
    int64_t launches_ = 0;
};

}  // namespace feas

struct FeasHandle {
    std::string err;
    std::unique_ptr<feas::SurfaceDev> impl;
};

static thread_local std::string g_feas_create_err;

#define FEAS_TRY(h, body)                  \
    if (!(h)) return FEAB_EINVAL;          \
    try {                                  \
        body;                              \
        return FEAB_OK;                    \
    } catch (const feas::Err& e) {         \
        (h)->err = e.msg;                  \
        return e.code;                     \
    } catch (const std::exception& e) {    \
        (h)->err = e.what();               \
        return FEAB_EINVAL;                \
    } catch (...) {                        \
        (h)->err = "unknown error";        \
        return FEAB_EINVAL;                \
    }

extern "C" {

int feas_create(int device, FeasHandle** out) {
    if (!out) return FEAB_EINVAL;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        g_feas_create_err = "no CUDA device: libfeab has no CPU fallback";
        return FEAB_ENOGPU;
    }
    if (device < 0 || device >= ndev) {
        g_feas_create_err = "bad device ordinal";
        return FEAB_EINVAL;
    }
    auto h = std::make_unique<FeasHandle>();
    try {
        h->impl = std::make_unique<feas::SurfaceDev>(device);
    } catch (const feas::Err& e) {
        g_feas_create_err = e.msg;
        return e.code;
    }
    *out = h.release();
    return FEAB_OK;
}
int feas_destroy(FeasHandle* h) {
    delete h;
    return FEAB_OK;
}
const char* feas_last_error(const FeasHandle* h) { return h ? h->err.c_str() : g_feas_create_err.c_str(); }

int feas_set_mesh(FeasHandle* h, int64_t V, const float* xyz, const float* nrm, int64_t T, const uint32_t* tri) {
    FEAS_TRY(h, h->impl->set_mesh(V, xyz, nrm, T, tri));
}
int feas_detect_boundary(FeasHandle* h, uint8_t* out, int64_t* nb, int* closed) { FEAS_TRY(h, h->impl->detect_boundary(out, nb, closed)); }
int feas_load_obj(FeasHandle* h, const char* path) { FEAS_TRY(h, h->impl->load_obj_file(path)); }
int feas_get_counts(const FeasHandle* h, int64_t* V, int64_t* T, int64_t* nb) {
    if (!h) return FEAB_EINVAL;
    if (V) *V = h->impl->V();
    if (T) *T = h->impl->T();
    if (nb) *nb = h->impl->nb();
    return FEAB_OK;
}
int feas_get_mesh(FeasHandle* h, float* xyz, float* nrm, uint32_t* tri, uint8_t* onb) { FEAS_TRY(h, h->impl->get_mesh(xyz, nrm, tri, onb)); }
int feas_set_values(FeasHandle* h, const float* v) { FEAS_TRY(h, h->impl->set_values(v, false)); }
int feas_get_values(FeasHandle* h, float* v) { FEAS_TRY(h, h->impl->get_values(v)); }
int feas_set_values_device(FeasHandle* h, const float* v) { FEAS_TRY(h, h->impl->set_values(v, true)); }
int feas_set_color_map(FeasHandle* h, const float* c21) { FEAS_TRY(h, h->impl->set_color_map(c21)); }
int feas_color_map(FeasHandle* h, int64_t n, const float* t, float* rgb) { FEAS_TRY(h, h->impl->color_map(n, t, rgb)); }
int feas_smooth_normals(FeasHandle* h, float ve, float* out) { FEAS_TRY(h, h->impl->smooth_normals(ve, out)); }
int feas_ply_build(FeasHandle* h, float ve, float thr, int mesh_type, int64_t* nv, int64_t* nt) {
    FEAS_TRY(h, h->impl->ply_build(ve, thr, mesh_type, nv, nt));
}
int feas_ply_fetch(FeasHandle* h, float* pos, float* val, float* col, uint32_t* tri) { FEAS_TRY(h, h->impl->ply_fetch(pos, val, col, tri)); }
int feas_export_ply(FeasHandle* h, const char* path, float ve, float thr, int mesh_type) {
    FEAS_TRY(h, h->impl->export_ply(path, ve, thr, mesh_type));
}
int feas_pick(FeasHandle* h, const float* o, const float* d, int* tri_idx, float* point3, float* dist) {
    FEAS_TRY(h, h->impl->pick(o, d, false, 0.0f, tri_idx, nullptr, point3, dist));
}
int feas_brush(FeasHandle* h, const float* o, const float* d, float value, int* node) {
    FEAS_TRY(h, h->impl->pick(o, d, true, value, nullptr, node, nullptr, nullptr));
}
int feas_write_ply(const char* path, int64_t nv, const float* pos, const float* col, int64_t nt, const uint32_t* tri) {
    if (!path || nv < 0 || nt < 0 || (nv > 0 && (!pos || !col)) || (nt > 0 && !tri)) return FEAB_EINVAL;
    std::string e;
    return feas::write_ply(path, nv, pos, col, nt, tri, e) ? FEAB_OK : FEAB_EINVAL;
}
int feas_parse_obj(const char* path, int64_t* nv, int64_t* nt, float* xyz, float* nrm, uint32_t* tri) {
    if (!path) return FEAB_EINVAL;
    feas::ObjMesh m;
    std::string e;
    if (!feas::load_obj(path, m, e)) {
        g_feas_create_err = e;
        return FEAB_EINVAL;
    }
    if (nv) *nv = (int64_t)m.xyz.size() / 3;
    if (nt) *nt = (int64_t)m.tri.size() / 3;
    if (xyz) std::memcpy(xyz, m.xyz.data(), m.xyz.size() * sizeof(float));
    if (nrm) std::memcpy(nrm, m.nrm.data(), m.nrm.size() * sizeof(float));
    if (tri) std::memcpy(tri, m.tri.data(), m.tri.size() * sizeof(uint32_t));
    return FEAB_OK;
}
int feas_last_ms(const FeasHandle* h, double* ms, int64_t* launches) {
    if (!h) return FEAB_EINVAL;
    if (ms) *ms = h->impl->last_ms();
    if (launches) *launches = h->impl->launches();
    return FEAB_OK;
}

}  // extern "C"
