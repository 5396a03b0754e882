// Mirror of the reference's abstract solver interface (reference include/FEM/Solver.hpp:8-15): the drop-in boundary
// of the time-stepping path.  CPUSolver and GPUSolver implement it there; CUDASolver (FEM/CUDASolver.hpp) here.
#pragma once
#include <memory>

#include "FEM/FEMContext.hpp"

class Solver {
   public:
    std::shared_ptr<FEMContext> fem_ctx;

    virtual void advance_time() = 0;
    virtual void clear_values() = 0;
    virtual bool has_numerical_instability() = 0;
    virtual ~Solver() = default;
};
