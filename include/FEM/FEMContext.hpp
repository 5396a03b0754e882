// Standalone mirror of the PUBLIC part of the reference's FEMContext (reference include/FEM/FEMContext.hpp:9-42):
// same enums, members and methods.  The three Eigen matrices of the reference (private, :47-49) do not exist on the
// host here -- they are assembled and kept in HBM by libfeab.so (feab_set_mesh / feab_assemble); this class only
// keeps what the rest of the application reads: surface, parameters, idx_map, equation, boundary_condition.
// `revision` counts the events after which the reference re-assembles (init_from_surface, update_boundary_conditions,
// assemble_matrices); CUDASolver compares it with the revision it last uploaded.
#pragma once
#include <memory>
#include <unordered_map>
#include <vector>

#include "FEM/EquationParameters.hpp"
#include "Utils/Surface.hpp"

enum class Equation {  // reference :9-14
    Heat = 0,
    Wave,
    Advection_Diffusion,
    Reaction_Diffusion,
};

enum class BoundaryCondition {  // reference :16-19
    Dirichlet = 0,
    Neumann,
};

class FEMContext {
   public:
    std::shared_ptr<Surface> surface;
    std::unordered_map<Equation, std::shared_ptr<EquationParameters>> parameters;
    std::vector<int> idx_map;

    Equation equation = Equation::Wave;                                   // reference :31
    BoundaryCondition boundary_condition = BoundaryCondition::Dirichlet;  // reference :32

    FEMContext() {  // reference src/FEM/FEMContext.cpp:10-15
        parameters[Equation::Heat] = std::make_shared<HeatParameters>();
        parameters[Equation::Advection_Diffusion] = std::make_shared<AdvectionDiffusionParameters>();
        parameters[Equation::Wave] = std::make_shared<WaveParameters>();
        parameters[Equation::Reaction_Diffusion] = std::make_shared<ReactionDiffusionParameters>();
    }

    // reference src/FEM/FEMContext.cpp:21-27
    void init_from_surface(std::shared_ptr<Surface> s) {
        if (s->initialized) {
            surface = s;
            mesh_revision++;
            update_boundary_conditions();
        }
    }
    // reference src/FEM/FEMContext.cpp:224-240: unknown numbering = node order with Dirichlet boundary nodes removed
    void update_boundary_conditions() {
        idx_map.assign(surface->vertices.size(), -1);
        int next = 0;
        for (size_t i = 0; i < idx_map.size(); i++)
            if (boundary_condition == BoundaryCondition::Neumann || !surface->on_boundary[i]) idx_map[i] = next++;
        assemble_matrices();
    }
    // reference src/FEM/FEMContext.cpp:57-64: here only a request -- the CUDA side assembles on the next solver call
    void assemble_matrices() { revision++; }

    unsigned int num_nodes() { return (unsigned int)surface->vertices.size(); }  // reference :33-35
    unsigned int num_unknowns() {                                                  // reference :46-48
        return (unsigned int)surface->vertices.size() -
               (boundary_condition == BoundaryCondition::Dirichlet ? (unsigned int)surface->num_boundary_points : 0u);
    }
    unsigned int num_max_nonzeros_per_row() { return (unsigned int)max_row_nonzeros; }  // reference :50-52

    // -- not in the reference: bookkeeping for the device-side copy --
    unsigned long long revision = 0;       // bumped whenever the reference would re-assemble
    unsigned long long mesh_revision = 0;  // bumped when the surface itself was replaced
    int max_row_nonzeros = 0;              // filled in by CUDASolver after assembly (feab_get_counts)
};
