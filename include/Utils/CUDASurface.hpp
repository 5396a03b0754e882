// C++ host side of the Surface-side rows: the methods of the reference's Surface that are
// served by the device (reference include/Utils/Surface.hpp:54-60) -- export_to_ply, the boundary pass of init_from_obj,
// calculate_normals -- as a thin class over the C ABI of include/feab_surface.h.  Works with the reference's own
// Surface (glm::vec3 members) and with the mirror in this repository: both store 3 packed floats per vertex.
//
//   CUDASurface dev(surface);                        // uploads vertices / normals / triangles
//   dev.detect_boundary();                           // fills surface->on_boundary, num_boundary_points, closed
//   dev.export_to_ply("out.ply", extrusion, threshold, mesh_type);   // reads surface->values
//   dev.export_to_ply_from(solver_handle, ...);      // takes the field straight from a CUDASolver, device to device
#pragma once
#include <array>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "Utils/Surface.hpp"
#include "feab_surface.h"

class CUDASurface {
   public:
    std::shared_ptr<Surface> surface;

    explicit CUDASurface(std::shared_ptr<Surface> s, int device = 0) : surface(std::move(s)) {
        if (feas_create(device, &h_) != 0) throw std::runtime_error(std::string("CUDASurface: ") + feas_last_error(nullptr));
        upload();
    }
    ~CUDASurface() { feas_destroy(h_); }
    CUDASurface(const CUDASurface&) = delete;
    CUDASurface& operator=(const CUDASurface&) = delete;

    // after the host Surface changed (new mesh)
    void upload() {
        static_assert(sizeof(surface->vertices[0]) == 12 && sizeof(Triangle) == 12, "packed vec3 / Triangle expected");
        const float* nrm = surface->normals.size() == surface->vertices.size() ? reinterpret_cast<const float*>(surface->normals.data()) : nullptr;
        ck(feas_set_mesh(h_, (int64_t)surface->vertices.size(), reinterpret_cast<const float*>(surface->vertices.data()), nrm,
                         (int64_t)surface->triangles.size(), reinterpret_cast<const uint32_t*>(surface->triangles.data())));
    }
    // the boundary pass of Surface::init_from_obj (reference src/Utils/Surface.cpp:103-129)
    void detect_boundary() {
        std::vector<uint8_t> ob(surface->vertices.size());
        int64_t nb = 0;
        int closed = 0;
        ck(feas_detect_boundary(h_, ob.data(), &nb, &closed));
        surface->on_boundary.assign(ob.begin(), ob.end());
        surface->num_boundary_points = (int)nb;
        surface->closed = closed != 0;
    }
    void set_color_map(const std::array<float, 21>& coeffs) { ck(feas_set_color_map(h_, coeffs.data())); }
    // Surface::export_to_ply (reference src/Utils/Surface.cpp:138-285); mesh_type: 0 Open, 1 Closed, 2 Mirrored
    void export_to_ply(const char* file_path, float vertex_extrusion = 0.25f, float threshold = 0.0f, int mesh_type = FEAS_MESH_OPEN) {
        ck(feas_set_values(h_, surface->values.data()));
        ck(feas_export_ply(h_, file_path, vertex_extrusion, threshold, mesh_type));
    }
    void export_to_ply_from(FeabHandle* solver, const char* file_path, float vertex_extrusion = 0.25f, float threshold = 0.0f,
                            int mesh_type = FEAS_MESH_OPEN) {
        const float* dv = nullptr;
        if (feab_values_device(solver, &dv) != 0) throw std::runtime_error(std::string("CUDASurface: ") + feab_last_error(solver));
        ck(feas_set_values_device(h_, dv));
        ck(feas_export_ply(h_, file_path, vertex_extrusion, threshold, mesh_type));
    }
    // Surface::calculate_normals (reference src/Utils/Surface.cpp:287-306 + shaders/FEM/smooth_normals.glsl): V x 3 floats
    std::vector<float> calculate_normals(float vertex_extrusion) {
        std::vector<float> out(3 * surface->vertices.size());
        ck(feas_set_values(h_, surface->values.data()));
        ck(feas_smooth_normals(h_, vertex_extrusion, out.data()));
        return out;
    }
    FeasHandle* handle() { return h_; }

   private:
    void ck(int rc) {
        if (rc != 0) throw std::runtime_error(std::string("CUDASurface: ") + feas_last_error(h_));
    }
    FeasHandle* h_ = nullptr;
};
