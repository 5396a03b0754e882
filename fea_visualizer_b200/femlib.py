"""Host-side mirror of the reference's solver interface for the hot path, in Python, on top of the C ABI.

Same names, argument meaning and error behaviour as the reference's classes
(/root/reference/include/FEM/Solver.hpp:8-15, include/FEM/FEMContext.hpp:25-59,
include/FEM/EquationParameters.hpp:4-49, include/Utils/Surface.hpp:36-52) so that tests and drivers read like
code written against the reference.  The C++ twin of this file is include/FEM/CUDASolver.hpp.
"""
from __future__ import annotations

import enum
from dataclasses import dataclass, field

import numpy as np

from . import feab as _feab
from . import meshes


class Equation(enum.IntEnum):  # include/FEM/FEMContext.hpp:9-14
    Heat = 0
    Wave = 1
    Advection_Diffusion = 2
    Reaction_Diffusion = 3


class BoundaryCondition(enum.IntEnum):  # include/FEM/FEMContext.hpp:16-19
    Dirichlet = 0
    Neumann = 1


@dataclass
class HeatParameters:  # EquationParameters.hpp:8-15
    time_step: float = 0.01
    conductivity: float = 0.05


@dataclass
class AdvectionDiffusionParameters:  # :17-27
    time_step: float = 0.001
    c: float = 0.25
    velocity: tuple = (1.0, 0.0, 0.0)


@dataclass
class WaveParameters:  # :29-36
    time_step: float = 0.05
    c: float = 0.05


@dataclass
class ReactionDiffusionParameters:  # :38-49
    time_step: float = 0.001
    Du: float = 0.08
    Dv: float = 0.04
    feed_rate: float = 0.035
    kill_rate: float = 0.06


@dataclass
class Surface:
    """The data members of the reference's Surface that the path reads and writes."""
    vertices: np.ndarray = None      # V x 3 float32
    triangles: np.ndarray = None     # T x 3 uint32
    on_boundary: np.ndarray = None   # V uint8
    values: np.ndarray = None        # V float32  (the field shown by the viewer / exported to .ply)
    num_boundary_points: int = 0
    closed: bool = False
    initialized: bool = False

    @staticmethod
    def from_arrays(vertices, triangles, on_boundary=None) -> "Surface":
        s = Surface()
        s.vertices = np.ascontiguousarray(vertices, dtype=np.float32)
        s.triangles = np.ascontiguousarray(triangles, dtype=np.uint32)
        if on_boundary is None:
            on_boundary = meshes.boundary_flags(len(s.vertices), s.triangles)  # Surface.cpp:104-122
        s.on_boundary = np.ascontiguousarray(on_boundary, dtype=np.uint8)
        s.num_boundary_points = int(s.on_boundary.sum())
        s.values = np.zeros(len(s.vertices), dtype=np.float32)
        s.closed = s.num_boundary_points == 0
        s.initialized = True
        return s

    @staticmethod
    def init_from_obj(path: str) -> "Surface":
        v, t = meshes.load_obj(path)
        return Surface.from_arrays(v, t)

    def clear_values(self):
        self.values[:] = 0.0


class FEMContext:
    """FEMContext with the matrices living on the GPU (they are private in the reference too)."""

    def __init__(self, real: int = 64, device: int = 0):
        self.surface: Surface | None = None
        self.parameters = {
            Equation.Heat: HeatParameters(),
            Equation.Advection_Diffusion: AdvectionDiffusionParameters(),
            Equation.Wave: WaveParameters(),
            Equation.Reaction_Diffusion: ReactionDiffusionParameters(),
        }
        self.idx_map = np.zeros(0, dtype=np.int32)
        self.equation = Equation.Wave                      # FEMContext.hpp:31
        self.boundary_condition = BoundaryCondition.Dirichlet  # :32
        self._dev = _feab.Feab(real=real, device=device)
        self._velocity = None

    def init_from_surface(self, surface: Surface):
        if surface.initialized:                            # FEMContext.cpp:22
            self.surface = surface
            self._dev.set_boundary_condition(int(self.boundary_condition))
            vel = self.parameters[Equation.Advection_Diffusion].velocity
            self._dev.set_mesh(surface.vertices, surface.triangles, surface.on_boundary)  # symbolic + assembly
            if tuple(vel) != (1.0, 0.0, 0.0):
                self._dev.assemble(vel)
            self._velocity = tuple(vel)
            self.idx_map = self._dev.idx_map()

    def update_boundary_conditions(self):
        self._dev.set_boundary_condition(int(self.boundary_condition))
        self.idx_map = self._dev.idx_map()

    def assemble_matrices(self):
        vel = tuple(self.parameters[Equation.Advection_Diffusion].velocity)
        self._dev.assemble(vel)
        self._velocity = vel

    def num_nodes(self) -> int:
        return len(self.surface.vertices)

    def num_unknowns(self) -> int:
        return self._dev.counts()[1]

    def num_max_nonzeros_per_row(self) -> int:
        return self._dev.counts()[3]


class Solver:
    """include/FEM/Solver.hpp:8-15"""
    fem_ctx: FEMContext

    def advance_time(self):
        raise NotImplementedError

    def clear_values(self):
        raise NotImplementedError

    def has_numerical_instability(self) -> bool:
        raise NotImplementedError


class CUDASolver(Solver):
    """Drop-in for CPUSolver: same three calls, data through fem_ctx.surface.values."""

    def __init__(self, fem_ctx: FEMContext):
        self.fem_ctx = fem_ctx

    def _push_params(self):
        ctx = self.fem_ctx
        d = ctx._dev
        eq = ctx.equation
        p = ctx.parameters[eq]
        d.set_equation(int(eq))
        if eq == Equation.Heat:
            d.set_params(int(eq), p.time_step, p.conductivity)
        elif eq == Equation.Wave:
            d.set_params(int(eq), p.time_step, p.c)
        elif eq == Equation.Advection_Diffusion:
            d.set_params(int(eq), p.time_step, p.c)
        else:
            d.set_params(int(eq), p.time_step, p.Du, p.Dv, p.feed_rate, p.kill_rate)

    def advance_time(self, nsteps: int = 1):
        ctx = self.fem_ctx
        self._push_params()   # parameters are read fresh at every call, like CPUSolver::advance_time
        ctx._dev.step(nsteps, values=ctx.surface.values, sync_values=True)

    def clear_values(self):
        self.fem_ctx._dev.clear()

    def has_numerical_instability(self) -> bool:
        return self.fem_ctx._dev.unstable()
