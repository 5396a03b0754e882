"""fea_visualizer_b200: B200-native (sm_100a) time-stepping hot path of tvumcc/fea-visualizer.

The product is libfeab.so (CUDA kernels behind the C ABI in include/feab.h); this package holds its sources
(csrc/), the ctypes binding (feab.py), the host-side mirror of the reference's Solver / FEMContext interface
(femlib.py) and the synthetic mesh generators (meshes.py).  There is no CPU fallback.
"""
from . import meshes  # noqa: F401
from .feab import Feab, FeabError, load_library  # noqa: F401
from .femlib import (BoundaryCondition, CUDASolver, Equation, FEMContext, Surface)  # noqa: F401
