"""TEST INFRASTRUCTURE ONLY: ctypes bindings for the CPU oracle (liboracle.so, fea_oracle.c) and for the
compiled, unmodified reference (oracle/_ref/libfea_ref*.so, ref_harness.cpp).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / ``--impl reference`` legs import this
module.  The product path (fea_visualizer_b200 -> libfeab.so) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
EQ_HEAT, EQ_WAVE, EQ_ADVDIFF, EQ_GRAYSCOTT = 0, 1, 2, 3  # include/FEM/FEMContext.hpp:9-14
BC_DIRICHLET, BC_NEUMANN = 0, 1                          # include/FEM/FEMContext.hpp:16-19
MAT_MASS, MAT_STIFFNESS, MAT_ADVECTION = 0, 1, 2


def build(verbose: bool = False) -> None:
    """Compile liboracle.so and, when /root/reference is present, oracle/_ref (building the checker is
    not using it)."""
    out = subprocess.run(["make", "-C", HERE, "all"], capture_output=True, text=True)
    if verbose or out.returncode != 0:
        print(out.stdout, out.stderr)
    if out.returncode != 0:
        raise RuntimeError("oracle build failed")


def _ptr(a, ct):
    return a.ctypes.data_as(C.POINTER(ct))


class _Base:
    prefix = ""
    real = np.float32
    lib = None

    def _f(self, name):
        return getattr(self.lib, self.prefix + name)

    def __init__(self, xyz, tri, on_boundary, bc=BC_DIRICHLET):
        xyz = np.ascontiguousarray(xyz, dtype=np.float32)
        tri = np.ascontiguousarray(tri, dtype=np.uint32)
        ob = np.ascontiguousarray(on_boundary, dtype=np.uint8)
        self.V, self.T = xyz.shape[0], tri.shape[0]
        f = self._f("create")
        f.restype = C.c_void_p
        f.argtypes = [C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int]
        self.h = C.c_void_p(f(self.V, xyz.ctypes.data, self.T, tri.ctypes.data, ob.ctypes.data, bc))
        for name in ("num_unknowns", "num_nodes", "nnz"):
            self._f(name).restype = C.c_int64

    def close(self):
        if self.h:
            self._f("destroy")(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- FEMContext surface --
    def set_boundary_condition(self, bc):
        self._f("set_boundary_condition")(self.h, C.c_int(bc))

    def set_equation(self, eq):
        self._f("set_equation")(self.h, C.c_int(eq))

    def set_params(self, eq, dt, c0=0.0, c1=0.0, c2=0.0, c3=0.0):
        q = np.array([dt, c0, c1, c2, c3], dtype=np.float32)
        self._f("set_params")(self.h, C.c_int(eq), C.c_void_p(q.ctypes.data))

    def set_velocity(self, vel):
        q = np.array(vel, dtype=np.float32)
        self._f("set_velocity")(self.h, C.c_void_p(q.ctypes.data))

    def num_unknowns(self):
        return int(self._f("num_unknowns")(self.h))

    def max_row_nonzeros(self):
        return int(self._f("max_row_nonzeros")(self.h))

    def idx_map(self):
        out = np.empty(self.V, dtype=np.int32)
        self._f("get_idx_map")(self.h, C.c_void_p(out.ctypes.data))
        return out

    def csc(self, which):
        """(outer, inner, val) of Eigen's compressed column-major storage."""
        n = self.num_unknowns()
        nnz = int(self._f("nnz")(self.h, C.c_int(which)))
        outer = np.empty(n + 1, dtype=np.int32)
        inner = np.empty(max(nnz, 1), dtype=np.int32)
        val = np.empty(max(nnz, 1), dtype=self.real)
        self._f("get_csc")(self.h, C.c_int(which), C.c_void_p(outer.ctypes.data), C.c_void_p(inner.ctypes.data),
                           C.c_void_p(val.ctypes.data))
        return outer, inner[:nnz], val[:nnz]

    # -- Solver surface --
    def clear(self):
        self._f("clear")(self.h)

    def set_values(self, v):
        v = np.ascontiguousarray(v, dtype=np.float32)
        assert v.shape[0] == self.V
        self._f("set_values")(self.h, C.c_void_p(v.ctypes.data))

    def get_values(self):
        out = np.empty(self.V, dtype=np.float32)
        self._f("get_values")(self.h, C.c_void_p(out.ctypes.data))
        return out

    def get_state(self, which):
        out = np.empty(self.num_unknowns(), dtype=self.real)
        self._f("get_state")(self.h, C.c_int(which), C.c_void_p(out.ctypes.data))
        return out

    def step(self, nsteps=1):
        self._f("step")(self.h, C.c_int(nsteps))

    def unstable(self):
        return bool(self._f("unstable")(self.h))


_oracle_lib = None


def _load_oracle():
    global _oracle_lib
    if _oracle_lib is None:
        path = os.path.join(HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        _oracle_lib = C.CDLL(path)
    return _oracle_lib


class Oracle(_Base):
    """The CPU restatement; ``real`` = 32 or 64."""

    def __init__(self, xyz, tri, on_boundary, bc=BC_DIRICHLET, real=64):
        self.lib = _load_oracle()
        self.prefix = f"orc{real}_"
        self.real = np.float32 if real == 32 else np.float64
        super().__init__(xyz, tri, on_boundary, bc)

    def set_solver(self, mode=0, tol=-1.0, max_iter=-1):
        """mode 0: parity PCG (multiplies by A^T like Eigen); 1: BiCGSTAB for adv-diff.  tol < 0: eps of real."""
        self._f("set_solver")(self.h, C.c_int(mode), C.c_double(tol), C.c_int64(max_iter))

    def set_state(self, which, x):
        x = np.ascontiguousarray(x, dtype=self.real)
        self._f("set_state")(self.h, C.c_int(which), C.c_void_p(x.ctypes.data))

    def last_stats(self):
        it = np.zeros(2, dtype=np.int64)
        err = np.zeros(2, dtype=np.float64)
        self._f("last_stats")(self.h, C.c_void_p(it.ctypes.data), C.c_void_p(err.ctypes.data))
        return it, err


def reference_available(fast=False) -> bool:
    return os.path.exists(os.path.join(HERE, "_ref", "libfea_ref_fast.so" if fast else "libfea_ref.so"))


_ref_libs = {}


class Reference(_Base):
    """The unmodified reference (FEMContext.cpp + CPUSolver.cpp, Eigen fp32) behind ref_harness.cpp."""

    def __init__(self, xyz, tri, on_boundary, bc=BC_DIRICHLET, fast=False):
        name = "libfea_ref_fast.so" if fast else "libfea_ref.so"
        if name not in _ref_libs:
            _ref_libs[name] = C.CDLL(os.path.join(HERE, "_ref", name))
        self.lib = _ref_libs[name]
        self.prefix = "ref_"
        self.real = np.float32
        super().__init__(xyz, tri, on_boundary, bc)

    def set_threads(self, n):
        f = self.lib.ref_set_threads
        f.restype = C.c_int
        return int(f(C.c_int(n)))


def eigen_bicgstab_available() -> bool:
    return os.path.exists(os.path.join(HERE, "_ref", "libeigen_bicgstab.so"))


def eigen_bicgstab(rowptr, col, val, b, tol_rel=None, max_iter=-1):
    """Eigen's own BiCGSTAB<SparseMatrix<double>, DiagonalPreconditioner<double>> (lib/eigen, BiCGSTAB.h:32-119) on a CSR
    system, x0 = 0, stopping on |r| <= tol_rel |b| (ref_bicgstab_harness.cpp).  Returns (x, iterations, error, info)."""
    name = "libeigen_bicgstab.so"
    if name not in _ref_libs:
        _ref_libs[name] = C.CDLL(os.path.join(HERE, "_ref", name))
    lib = _ref_libs[name]
    rowptr = np.ascontiguousarray(rowptr, dtype=np.int32)
    col = np.ascontiguousarray(col, dtype=np.int32)
    val = np.ascontiguousarray(val, dtype=np.float64)
    b = np.ascontiguousarray(b, dtype=np.float64)
    n = len(b)
    assert len(rowptr) == n + 1 and len(col) == len(val) == rowptr[-1]
    x = np.zeros(n, np.float64)
    it, err = C.c_int64(), C.c_double()
    f = lib.eigen_bicgstab_solve
    f.restype = C.c_int
    info = f(C.c_int64(n), C.c_void_p(rowptr.ctypes.data), C.c_void_p(col.ctypes.data), C.c_void_p(val.ctypes.data),
             C.c_void_p(b.ctypes.data), C.c_double(np.finfo(np.float64).eps if tol_rel is None else tol_rel), C.c_int64(max_iter),
             C.c_void_p(x.ctypes.data), C.byref(it), C.byref(err))
    return x, int(it.value), float(err.value), int(info)


def csc_to_dense(n, outer, inner, val):
    a = np.zeros((n, n), dtype=val.dtype)
    for j in range(n):
        a[inner[outer[j]:outer[j + 1]], j] = val[outer[j]:outer[j + 1]]
    return a
