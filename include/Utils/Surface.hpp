// Standalone mirror of the DATA MEMBERS of the reference's Surface that the time-stepping path reads and writes
// (reference include/Utils/Surface.hpp:11-23 Triangle, :36-52 members).  No OpenGL, no glm, no shaders: this header
// exists so that include/FEM/CUDASolver.hpp can be compiled and tested headless in this repository.  Inside the
// reference tree the reference's own Utils/Surface.hpp is found first and CUDASolver.hpp works with it unchanged --
// it only needs `vertices` (3 packed floats per element, like glm::vec3), `triangles`, `on_boundary`, `values`,
// `num_boundary_points`, `initialized`.
#pragma once
#include <cstddef>
#include <memory>
#include <vector>

namespace feab_mirror {
struct vec3 {
    float x, y, z;
};
}  // namespace feab_mirror

struct Triangle {  // reference include/Utils/Surface.hpp:11-23
    unsigned int idx_a;
    unsigned int idx_b;
    unsigned int idx_c;
    unsigned int& operator[](int idx) { return idx == 0 ? idx_a : idx == 1 ? idx_b : idx_c; }
};

class Surface {
   public:
    std::vector<feab_mirror::vec3> vertices;  // :36
    std::vector<feab_mirror::vec3> normals;   // :37  per-vertex normals (OBJ `vn`, or (0,1,0) for PSLG meshes)
    std::vector<float> values;                // :38  the field shown by the viewer / exported to .ply
    std::vector<Triangle> triangles;          // :39
    std::vector<bool> on_boundary;            // :40
    int num_boundary_points = 0;              // :50
    bool closed = false;
    bool initialized = false;

    // Surface::clear_values (reference src/Utils/Surface.cpp): zero field
    void clear_values() { values.assign(vertices.size(), 0.0f); }

    // Boundary flags exactly as Surface::init_from_obj derives them (reference src/Utils/Surface.cpp:104-122): a vertex
    // is on the boundary iff it touches an edge used by exactly one triangle.  Implemented in Surface_mirror.cpp-free
    // form below (sort-based) so the header stays self-contained.
    void finish_init();
};

#include <algorithm>
#include <cstdint>

inline void Surface::finish_init() {
    std::vector<uint64_t> edges;
    edges.reserve(triangles.size() * 3);
    for (Triangle t : triangles) {
        for (int e = 0; e < 3; e++) {
            uint64_t a = t[e], b = t[(e + 1) % 3];
            if (a > b) std::swap(a, b);
            edges.push_back((a << 32) | b);
            if (a == b) edges.push_back((a << 32) | b);  // the reference increments edges[a][a] twice: never a boundary edge
        }
    }
    std::sort(edges.begin(), edges.end());
    on_boundary.assign(vertices.size(), false);
    for (size_t i = 0; i < edges.size();) {
        size_t j = i;
        while (j < edges.size() && edges[j] == edges[i]) j++;
        if (j - i == 1) {
            on_boundary[(size_t)(edges[i] >> 32)] = true;
            on_boundary[(size_t)(edges[i] & 0xffffffffu)] = true;
        }
        i = j;
    }
    num_boundary_points = (int)std::count(on_boundary.begin(), on_boundary.end(), true);
    closed = num_boundary_points == 0;
    values.assign(vertices.size(), 0.0f);
    initialized = true;
}
