// Host half of the .ply exporter: the ASCII writer of Surface::export_to_ply (src/Utils/Surface.cpp:263-284) and the
// OBJ reader of Surface::init_from_obj (:58-101).  Plain C++ (no CUDA), part of libfeab.so.
//
// std::format("{}", float) is specified as std::to_chars(first, last, value) -- the shortest decimal string that
// round-trips, fixed or scientific whichever is shorter -- so calling std::to_chars directly reproduces the
// reference's bytes.  Lines are formatted by a pool of threads into per-chunk buffers and written in order.
#include "ply_writer.hpp"

#include <algorithm>
#include <charconv>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <thread>

namespace feas {

namespace {
inline char* put_float(char* p, float v) {
    auto r = std::to_chars(p, p + 32, v);
    return r.ptr;
}
inline char* put_uint(char* p, uint32_t v) {
    auto r = std::to_chars(p, p + 16, v);
    return r.ptr;
}

template <class F>
void format_chunks(FILE* f, int64_t n, size_t max_line, F&& line) {
    if (n <= 0) return;
    const int64_t chunk = 1 << 16;
    const int64_t nchunks = (n + chunk - 1) / chunk;
    unsigned hw = std::thread::hardware_concurrency();
    const int nthreads = (int)std::max<int64_t>(1, std::min<int64_t>({(int64_t)(hw ? hw : 1), nchunks, 64}));
    // process the chunks in waves of nthreads so the memory held is bounded and the file order is kept
    std::vector<std::vector<char>> buf(nthreads);
    std::vector<size_t> len(nthreads);
    for (int64_t base = 0; base < nchunks; base += nthreads) {
        const int wave = (int)std::min<int64_t>(nthreads, nchunks - base);
        auto work = [&](int w) {
            const int64_t lo = (base + w) * chunk, hi = std::min(n, lo + chunk);
            buf[w].resize((size_t)(hi - lo) * max_line);
            char* p = buf[w].data();
            for (int64_t i = lo; i < hi; i++) p = line(p, i);
            len[w] = (size_t)(p - buf[w].data());
        };
        std::vector<std::thread> th;
        for (int w = 1; w < wave; w++) th.emplace_back(work, w);
        work(0);
        for (auto& t : th) t.join();
        for (int w = 0; w < wave; w++) fwrite(buf[w].data(), 1, len[w], f);
    }
}
}  // namespace

bool write_ply(const char* path, int64_t nv, const float* pos, const float* col, int64_t nt, const uint32_t* tri, std::string& err) {
    FILE* f = fopen(path, "wb");
    if (!f) {
        err = std::string("cannot open ") + path;
        return false;
    }
    std::vector<char> big(1 << 22);
    setvbuf(f, big.data(), _IOFBF, big.size());
    // header: src/Utils/Surface.cpp:263-274
    fprintf(f,
            "ply\nformat ascii 1.0\nelement vertex %lld\nproperty float x\nproperty float y\nproperty float z\n"
            "property float red\nproperty float green\nproperty float blue\nelement face %lld\n"
            "property list uchar uint vertex_indices\nend_header\n",
            (long long)nv, (long long)nt);
    // "{} {} {} {} {} {}\n" (:279-282)
    format_chunks(f, nv, 6 * 17 + 2, [&](char* p, int64_t i) {
        for (int k = 0; k < 3; k++) {
            p = put_float(p, pos[3 * i + k]);
            *p++ = ' ';
        }
        for (int k = 0; k < 3; k++) {
            p = put_float(p, col[3 * i + k]);
            *p++ = k == 2 ? '\n' : ' ';
        }
        return p;
    });
    // "3 {} {} {}\n" (:284-285)
    format_chunks(f, nt, 3 * 11 + 4, [&](char* p, int64_t i) {
        *p++ = '3';
        for (int k = 0; k < 3; k++) {
            *p++ = ' ';
            p = put_uint(p, tri[3 * i + k]);
        }
        *p++ = '\n';
        return p;
    });
    bool ok = fflush(f) == 0;
    ok = (fclose(f) == 0) && ok;
    if (!ok) err = std::string("write failed: ") + path;
    return ok;
}

// ------------------------------------------------------------------------------------------------------------------
// OBJ ingest with the semantics of Surface::init_from_obj (src/Utils/Surface.cpp:58-101)
// ------------------------------------------------------------------------------------------------------------------
namespace {
struct Corner {
    int v = -1, n = -1;
};
inline const char* skip_ws(const char* p) {
    while (*p == ' ' || *p == '\t') p++;
    return p;
}
inline const char* parse_int(const char* p, int& out, bool& ok) {
    char* e;
    long v = strtol(p, &e, 10);
    ok = e != p;
    out = (int)v;
    return e;
}
}  // namespace

bool load_obj(const char* path, ObjMesh& m, std::string& err) {
    std::ifstream in(path);
    if (!in) {
        err = "A valid file was not provided.";   // src/Utils/Surface.cpp:59-60
        return false;
    }
    std::vector<float> vn;
    // one index list per shape: the reference pushes ALL file vertices once PER SHAPE (:77-81) and never offsets the
    // face indices, so a multi-shape file yields duplicated, unreferenced vertices.  Reproduced as is.
    std::vector<std::vector<Corner>> shapes;
    std::string line;
    while (std::getline(in, line)) {
        const char* p = skip_ws(line.c_str());
        if (p[0] == 'v' && (p[1] == ' ' || p[1] == '\t')) {
            char* e;
            float x = strtof(p + 1, &e), y = strtof(e, &e), z = strtof(e, &e);
            m.file_xyz.insert(m.file_xyz.end(), {x, y, z});
        } else if (p[0] == 'v' && p[1] == 'n' && (p[2] == ' ' || p[2] == '\t')) {
            char* e;
            float x = strtof(p + 2, &e), y = strtof(e, &e), z = strtof(e, &e);
            vn.insert(vn.end(), {x, y, z});
        } else if ((p[0] == 'o' || p[0] == 'g') && (p[1] == ' ' || p[1] == '\t' || p[1] == 0)) {
            if (!shapes.empty() && !shapes.back().empty()) shapes.emplace_back();
        } else if (p[0] == 'f' && (p[1] == ' ' || p[1] == '\t')) {
            if (shapes.empty()) shapes.emplace_back();
            std::vector<Corner> cs;
            p = skip_ws(p + 1);
            const int nvert = (int)m.file_xyz.size() / 3, nnorm = (int)vn.size() / 3;
            while (*p && *p != '\r' && *p != '\n') {
                Corner c;
                bool ok;
                int v;
                p = parse_int(p, v, ok);
                if (!ok) break;
                c.v = v > 0 ? v - 1 : nvert + v;
                if (*p == '/') {
                    p++;
                    if (*p != '/') {
                        int vt;
                        p = parse_int(p, vt, ok);
                    }
                    if (*p == '/') {
                        p++;
                        int n;
                        p = parse_int(p, n, ok);
                        if (ok) c.n = n > 0 ? n - 1 : nnorm + n;
                    }
                }
                cs.push_back(c);
                p = skip_ws(p);
            }
            for (size_t k = 1; k + 1 < cs.size(); k++) {   // fan triangulation (tinyobj's quad rule is not reproduced)
                shapes.back().push_back(cs[0]);
                shapes.back().push_back(cs[k]);
                shapes.back().push_back(cs[k + 1]);
            }
        }
    }
    if (m.file_xyz.empty()) {
        err = std::string("File ") + path + " does not contain a valid .obj mesh.";   // :71-72
        return false;
    }
    if (vn.empty()) {
        err = std::string("File ") + path + " does not contain normals!";             // :73-74
        return false;
    }
    size_t nshapes = 0;
    for (auto& s : shapes) nshapes += !s.empty();
    const size_t nfile = m.file_xyz.size() / 3;
    const size_t V = nfile * std::max<size_t>(nshapes, 1) * (nshapes ? 1 : 0);
    m.xyz.clear();
    m.tri.clear();
    for (auto& s : shapes) {
        if (s.empty()) continue;
        m.xyz.insert(m.xyz.end(), m.file_xyz.begin(), m.file_xyz.end());
        for (auto& c : s) m.tri.push_back((uint32_t)c.v);
    }
    // normals: sum of the face-corner normals per vertex index, then glm::normalize = v * inversesqrt(dot(v, v))
    // with dot = x*x + y*y + z*z summed left to right and inversesqrt = 1 / sqrt (:92-101)
    m.nrm.assign(3 * V, 0.0f);
    for (auto& s : shapes)
        for (auto& c : s) {
            if (c.v < 0 || (size_t)c.v >= V || c.n < 0 || (size_t)c.n * 3 + 2 >= vn.size()) {
                err = "face index out of range";
                return false;
            }
            for (int d = 0; d < 3; d++) m.nrm[3 * (size_t)c.v + d] += vn[3 * (size_t)c.n + d];
        }
    for (size_t i = 0; i < V; i++) {
        float x = m.nrm[3 * i], y = m.nrm[3 * i + 1], z = m.nrm[3 * i + 2];
        float d = x * x + y * y + z * z;
        float inv = 1.0f / std::sqrt(d);
        m.nrm[3 * i] = x * inv;
        m.nrm[3 * i + 1] = y * inv;
        m.nrm[3 * i + 2] = z * inv;
    }
    return true;
}

}  // namespace feas
