// TEST INFRASTRUCTURE ONLY.
//
// Stand-in for the third-party header <tinyobjloader/tiny_obj_loader.h>: the reference vendors tinyobjloader as a git
// submodule (lib/tinyobjloader, .SUBMODULES.json:44-50) and the snapshot under /root/reference has it EMPTY, so
// src/Utils/Surface.cpp cannot be compiled against the real thing.  This file declares only the part of the tinyobj
// API that Surface::init_from_obj touches (src/Utils/Surface.cpp:58-101) and implements it with a small OBJ reader of
// our own, so that the UNMODIFIED Surface.cpp compiles and its own code (vertex/triangle extraction, normal
// accumulation, boundary-edge detection :104-127) runs headless.
//
// What is NOT pinned by this: tinyobjloader's polygon triangulation rule.  Faces with more than 3 corners are split as
// a fan (v0, v(k), v(k+1)); tinyobjloader picks a diagonal by geometry for quads.  Parity fixtures therefore use
// triangle-only OBJ files (assets/fem_meshes/icosphere.obj is one); tube.obj (all quads) is used only for invariants
// that do not depend on the split (vertex count, boundary set).
#pragma once
#include <cstdlib>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

namespace tinyobj {

typedef float real_t;

struct index_t {
    int vertex_index = -1;
    int normal_index = -1;
    int texcoord_index = -1;
};
struct mesh_t {
    std::vector<index_t> indices;
    std::vector<unsigned int> num_face_vertices;  // always 3 (triangulated)
};
struct shape_t {
    std::string name;
    mesh_t mesh;
};
struct attrib_t {
    std::vector<float> vertices;
    std::vector<float> normals;
    std::vector<float> texcoords;
};
struct ObjReaderConfig {
    bool triangulate = true;
};

class ObjReader {
public:
    bool ParseFromFile(const std::string& path, const ObjReaderConfig& cfg = ObjReaderConfig()) {
        (void)cfg;
        std::ifstream in(path);
        if (!in) {
            err_ = "cannot open " + path;
            return false;
        }
        attrib_ = attrib_t();
        shapes_.clear();
        std::string line;
        auto cur = [&]() -> shape_t& {
            if (shapes_.empty()) shapes_.emplace_back();
            return shapes_.back();
        };
        while (std::getline(in, line)) {
            std::istringstream ss(line);
            std::string tag;
            if (!(ss >> tag)) continue;
            if (tag == "v") {
                float x = 0, y = 0, z = 0;
                ss >> x >> y >> z;
                attrib_.vertices.insert(attrib_.vertices.end(), {x, y, z});
            } else if (tag == "vn") {
                float x = 0, y = 0, z = 0;
                ss >> x >> y >> z;
                attrib_.normals.insert(attrib_.normals.end(), {x, y, z});
            } else if (tag == "vt") {
                float u = 0, v = 0;
                ss >> u >> v;
                attrib_.texcoords.insert(attrib_.texcoords.end(), {u, v});
            } else if (tag == "o" || tag == "g") {
                // a new shape starts when the current one already has faces (tinyobj flushes on o/g the same way)
                if (!shapes_.empty() && !shapes_.back().mesh.indices.empty()) shapes_.emplace_back();
                std::string name;
                ss >> name;
                cur().name = name;
            } else if (tag == "f") {
                std::vector<index_t> corners;
                std::string tok;
                while (ss >> tok) corners.push_back(parse_corner(tok));
                for (size_t k = 1; k + 1 < corners.size(); k++) {
                    cur().mesh.indices.push_back(corners[0]);
                    cur().mesh.indices.push_back(corners[k]);
                    cur().mesh.indices.push_back(corners[k + 1]);
                    cur().mesh.num_face_vertices.push_back(3);
                }
            }
        }
        return true;
    }
    const std::string& Error() const { return err_; }
    const attrib_t& GetAttrib() const { return attrib_; }
    const std::vector<shape_t>& GetShapes() const { return shapes_; }

private:
    index_t parse_corner(const std::string& tok) const {
        // v, v/vt, v//vn, v/vt/vn ; 1-based, negative = relative to the end
        index_t r;
        int field = 0;
        size_t start = 0;
        for (size_t i = 0; i <= tok.size(); i++) {
            if (i == tok.size() || tok[i] == '/') {
                if (i > start) {
                    int v = std::atoi(tok.substr(start, i - start).c_str());
                    int count = field == 0 ? (int)attrib_.vertices.size() / 3 : field == 2 ? (int)attrib_.normals.size() / 3 : 0;
                    int idx = v > 0 ? v - 1 : count + v;
                    if (field == 0) r.vertex_index = idx;
                    if (field == 1) r.texcoord_index = idx;
                    if (field == 2) r.normal_index = idx;
                }
                field++;
                start = i + 1;
            }
        }
        return r;
    }
    attrib_t attrib_;
    std::vector<shape_t> shapes_;
    std::string err_;
};

}  // namespace tinyobj
