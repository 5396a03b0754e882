// Host helpers of the Surface-side rows (plain C++): .ply ASCII writer and OBJ reader.  See ply_writer.cpp.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace feas {

// Writes the file Surface::export_to_ply writes (src/Utils/Surface.cpp:263-285) from already clipped / de-duplicated /
// coloured arrays.  pos, col: nv x 3; tri: nt x 3.
bool write_ply(const char* path, int64_t nv, const float* pos, const float* col, int64_t nt, const uint32_t* tri, std::string& err);

struct ObjMesh {
    std::vector<float> file_xyz;   // the `v` lines
    std::vector<float> xyz, nrm;   // Surface::vertices / normals
    std::vector<uint32_t> tri;     // Surface::triangles
};
bool load_obj(const char* path, ObjMesh& out, std::string& err);

}  // namespace feas
