"""ctypes binding of libfeab.so (include/feab.h).  Thin: every method is one C-ABI call.

The library is the product; if it is missing or cannot create a CUDA context this module raises --
there is no CPU path behind it.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libfeab.so")

EQ_HEAT, EQ_WAVE, EQ_ADVDIFF, EQ_GRAYSCOTT = 0, 1, 2, 3
BC_DIRICHLET, BC_NEUMANN = 0, 1
MAT_MASS, MAT_STIFFNESS, MAT_ADVECTION = 0, 1, 2
SOLVER_PARITY_CG, SOLVER_BICGSTAB = 0, 1
ASSEMBLY_SCATTER, ASSEMBLY_ORDERED = 0, 1
LOOP_GRAPH_WHILE, LOOP_HOST_BATCHED, LOOP_PERSISTENT = 0, 1, 2
SYNC_NONE, SYNC_FULL, SYNC_ROOT = 0, 1, 2

# every symbol include/feab.h declares (checked by tests/test_abi.py without a GPU)
ABI_SYMBOLS = [
    "feab_register_host_values",
    "feab_create", "feab_destroy", "feab_last_error", "feab_set_mesh", "feab_set_mesh_ex", "feab_set_boundary_condition", "feab_assemble",
    "feab_get_counts", "feab_get_idx_map", "feab_get_csr", "feab_set_equation", "feab_set_params", "feab_get_params",
    "feab_set_solver_options", "feab_get_solver_options", "feab_clear", "feab_set_values", "feab_get_values", "feab_brush",
    "feab_step", "feab_unstable", "feab_last_stats", "feab_get_state", "feab_set_state", "feab_time_kernel",
    "feab_device_info", "feab_mg_config", "feab_mg_set_partition", "feab_partition_rcb", "feab_mg_export", "feab_mg_connect",
    "feab_mg_get_layout", "feab_mg_get_maps", "feab_mg_get_partition", "feab_mg_group_step", "feab_pcg_profile",
    "feab_last_solve_kernel",
]


class FeabParams(C.Structure):
    _fields_ = [("time_step", C.c_float), ("c", C.c_float * 4)]


class FeabSolverOptions(C.Structure):
    _fields_ = [("solver", C.c_int), ("tolerance", C.c_double), ("max_iter", C.c_int64), ("loop_mode", C.c_int),
                ("assembly", C.c_int)]


class FeabStats(C.Structure):
    _fields_ = [("iterations", C.c_int64 * 2), ("rel_residual", C.c_double * 2), ("last_call_ms", C.c_double),
                ("last_call_steps", C.c_int64), ("kernel_launches", C.c_int64), ("solve_ms", C.c_double),
                ("shared_iterations", C.c_int64), ("spmv_passes", C.c_int64 * 2)]


class FeabMgExport(C.Structure):
    _fields_ = [("ipc_handle", C.c_ubyte * 64), ("base_ptr", C.c_uint64), ("arena_bytes", C.c_uint64), ("off_vec", C.c_uint64 * 5),
                ("n_own", C.c_int64), ("n_ghost", C.c_int64), ("ghost_base", C.c_int64), ("num_nodes", C.c_int64),
                ("rank", C.c_int32), ("world", C.c_int32), ("device", C.c_int32), ("pid", C.c_int32)]


class FeabError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libfeab error {code}: {msg}")
        self.code = code


_lib = None


def load_library() -> C.CDLL:
    """Loads the in-tree libfeab.so.  Raises if it has not been built (python -m fea_visualizer_b200.build)."""
    global _lib
    if _lib is None:
        # never run a binary older than its sources: rebuild (nvcc, under a file lock so that torchrun ranks do not race)
        # when the recorded source hash differs.  FEAB_NO_REBUILD=1 skips the check.
        if os.environ.get("FEAB_NO_REBUILD") != "1":
            try:
                import fcntl
                from . import build as _build
                with open(os.path.join(HERE, ".build.lock"), "w") as lk:
                    fcntl.flock(lk, fcntl.LOCK_EX)
                    if _build.stale():
                        _build.build()
            except Exception as e:  # no nvcc here: fall through to whatever is built, loudly
                import sys
                sys.stderr.write(f"[feab] libfeab.so could not be checked / rebuilt against its sources: {e}\n")
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(f"{LIB_PATH} not built; run __graft_entry__.build() (nvcc, sm_100a). There is no CPU fallback.")
        lib = C.CDLL(LIB_PATH)
        lib.feab_last_error.restype = C.c_char_p
        lib.feab_last_error.argtypes = [C.c_void_p]
        _lib = lib
    return _lib


class Feab:
    """One mesh on one GPU.  real = 64 (fp64) or 32 (fp32, the reference's arithmetic)."""

    def __init__(self, real: int = 64, device: int = 0):
        self.lib = load_library()
        self.real = np.float64 if real == 64 else np.float32
        self.h = C.c_void_p()
        rc = self.lib.feab_create(C.c_int(device), C.c_int(8 if real == 64 else 4), C.byref(self.h))
        if rc != 0:
            raise FeabError(rc, self.lib.feab_last_error(None).decode())
        self.V = 0

    def _ck(self, rc):
        if rc != 0:
            raise FeabError(rc, self.lib.feab_last_error(self.h).decode())

    def close(self):
        if getattr(self, "h", None):
            self.lib.feab_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- FEMContext --
    def set_mesh(self, xyz, tri, on_boundary):
        xyz = np.ascontiguousarray(xyz, dtype=np.float32)
        tri = np.ascontiguousarray(tri, dtype=np.uint32)
        ob = np.ascontiguousarray(on_boundary, dtype=np.uint8)
        assert xyz.ndim == 2 and xyz.shape[1] == 3 and ob.shape[0] == xyz.shape[0]
        self.V = xyz.shape[0]
        self._ck(self.lib.feab_set_mesh(self.h, C.c_int64(self.V), C.c_void_p(xyz.ctypes.data), C.c_int64(tri.shape[0]),
                                        C.c_void_p(tri.ctypes.data), C.c_void_p(ob.ctypes.data)))

    def set_boundary_condition(self, bc):
        self._ck(self.lib.feab_set_boundary_condition(self.h, C.c_int(bc)))

    def assemble(self, velocity=None):
        if velocity is None:
            self._ck(self.lib.feab_assemble(self.h, None))
        else:
            v = np.array(velocity, dtype=np.float32)
            self._ck(self.lib.feab_assemble(self.h, C.c_void_p(v.ctypes.data)))

    def counts(self):
        V, n, nnz, mr = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int()
        self._ck(self.lib.feab_get_counts(self.h, C.byref(V), C.byref(n), C.byref(nnz), C.byref(mr)))
        return V.value, n.value, nnz.value, mr.value

    def num_unknowns(self):
        return self.counts()[1]

    def idx_map(self):
        out = np.empty(self.V, dtype=np.int32)
        self._ck(self.lib.feab_get_idx_map(self.h, C.c_void_p(out.ctypes.data)))
        return out

    def csr(self, which):
        nnz = C.c_int64()
        self._ck(self.lib.feab_get_csr(self.h, C.c_int(which), C.byref(nnz), None, None, None))
        n = self.num_unknowns()
        rowptr = np.empty(n + 1, dtype=np.int32)
        col = np.empty(max(nnz.value, 1), dtype=np.int32)
        val = np.empty(max(nnz.value, 1), dtype=self.real)
        self._ck(self.lib.feab_get_csr(self.h, C.c_int(which), C.byref(nnz), C.c_void_p(rowptr.ctypes.data),
                                       C.c_void_p(col.ctypes.data), C.c_void_p(val.ctypes.data)))
        return rowptr, col[:nnz.value], val[:nnz.value]

    # -- Solver --
    def set_equation(self, eq):
        self._ck(self.lib.feab_set_equation(self.h, C.c_int(eq)))

    def set_params(self, eq, dt, c0=0.0, c1=0.0, c2=0.0, c3=0.0):
        p = FeabParams(dt, (C.c_float * 4)(c0, c1, c2, c3))
        self._ck(self.lib.feab_set_params(self.h, C.c_int(eq), C.byref(p)))

    def get_params(self, eq):
        p = FeabParams()
        self._ck(self.lib.feab_get_params(self.h, C.c_int(eq), C.byref(p)))
        return p.time_step, list(p.c)

    def set_solver(self, solver=None, tolerance=None, max_iter=None, loop_mode=None, assembly=None):
        o = FeabSolverOptions()
        self._ck(self.lib.feab_get_solver_options(self.h, C.byref(o)))
        if solver is not None:
            o.solver = solver
        if tolerance is not None:
            o.tolerance = tolerance
        if max_iter is not None:
            o.max_iter = max_iter
        if loop_mode is not None:
            o.loop_mode = loop_mode
        if assembly is not None:
            o.assembly = assembly
        self._ck(self.lib.feab_set_solver_options(self.h, C.byref(o)))

    def clear(self):
        self._ck(self.lib.feab_clear(self.h))

    def set_values(self, v):
        v = np.ascontiguousarray(v, dtype=np.float32)
        assert v.shape[0] == self.V
        self._ck(self.lib.feab_set_values(self.h, C.c_void_p(v.ctypes.data)))

    def get_values(self):
        out = np.empty(self.V, dtype=np.float32)
        self._ck(self.lib.feab_get_values(self.h, C.c_void_p(out.ctypes.data)))
        return out

    def brush(self, nodes, strength=1.0):
        nodes = np.ascontiguousarray(nodes, dtype=np.int64)
        self._ck(self.lib.feab_brush(self.h, C.c_void_p(nodes.ctypes.data), C.c_int64(nodes.shape[0]), C.c_float(strength)))

    def step(self, nsteps=1, values=None, sync_values=True):
        """values: optional float32 host array (length V) copied in before and (sync_values) out after."""
        if values is None:
            self._ck(self.lib.feab_step(self.h, None, C.c_int(nsteps), C.c_int(0)))
        else:
            assert values.dtype == np.float32 and values.flags.c_contiguous and values.shape[0] == self.V
            self._ck(self.lib.feab_step(self.h, C.c_void_p(values.ctypes.data), C.c_int(nsteps), C.c_int(int(sync_values))))   # FEAB_SYNC_*

    def register_host_values(self, values):
        """feab_register_host_values: page-locks the (pageable) float32 array the caller will pass to step(values=...) --
        what the C++ binding does with surface->values after a surface load.  None undoes it.  Keep `values` alive."""
        if values is None:
            self._ck(self.lib.feab_register_host_values(self.h, None, C.c_int64(0)))
            return
        assert values.dtype == np.float32 and values.flags.c_contiguous
        self._ck(self.lib.feab_register_host_values(self.h, C.c_void_p(values.ctypes.data), C.c_int64(values.size)))

    def step_ptr(self, host_ptr: int, nsteps: int, sync_values: bool):
        """Same with a raw host pointer (e.g. pinned torch memory)."""
        self._ck(self.lib.feab_step(self.h, C.c_void_p(host_ptr), C.c_int(nsteps), C.c_int(int(sync_values))))   # FEAB_SYNC_*: 0 none, 1 full, 2 root

    def unstable(self):
        f = C.c_int()
        self._ck(self.lib.feab_unstable(self.h, C.byref(f)))
        return bool(f.value)

    def stats(self):
        s = FeabStats()
        self._ck(self.lib.feab_last_stats(self.h, C.byref(s)))
        return {"iterations": list(s.iterations), "rel_residual": list(s.rel_residual), "last_call_ms": s.last_call_ms,
                "last_call_steps": s.last_call_steps, "kernel_launches": s.kernel_launches, "solve_ms": s.solve_ms,
                "shared_iterations": s.shared_iterations, "spmv_passes": list(s.spmv_passes)}

    def get_state(self, which):
        out = np.empty(max(self.num_unknowns(), 1), dtype=self.real)
        self._ck(self.lib.feab_get_state(self.h, C.c_int(which), C.c_void_p(out.ctypes.data)))
        return out[:self.num_unknowns()]

    def set_state(self, which, x):
        x = np.ascontiguousarray(x, dtype=self.real)
        assert x.shape[0] == self.num_unknowns()
        self._ck(self.lib.feab_set_state(self.h, C.c_int(which), C.c_void_p(x.ctypes.data)))

    def time_kernel(self, kernel, reps=20):
        ms, by = C.c_double(), C.c_double()
        self._ck(self.lib.feab_time_kernel(self.h, C.c_int(kernel), C.c_int(reps), C.byref(ms), C.byref(by)))
        return ms.value, by.value

    def pcg_profile(self, reset=True):
        out = (C.c_double * 16)()
        self._ck(self.lib.feab_pcg_profile(self.h, out, C.c_int(1 if reset else 0)))
        names = ["spmv", "sync1", "unused1", "update", "sync2", "unused2", "direction", "sync3", "unused3"]
        steps = ["upload", "gather_rhs", "solves", "epilogue", "broadcast", "download"]
        return {"ns": dict(zip(names, list(out)[:9])), "iterations": int(out[9]), "step_ms": dict(zip(steps, list(out)[10:16]))}

    def last_solve_kernel(self):
        name = C.create_string_buffer(64)
        self._ck(self.lib.feab_last_solve_kernel(self.h, name))
        return name.value.decode()

    def device_info(self):
        sm, mem = C.c_int(), C.c_int64()
        name = C.create_string_buffer(64)
        self._ck(self.lib.feab_device_info(self.h, C.byref(sm), C.byref(mem), name))
        return {"sm_count": sm.value, "total_mem": mem.value, "name": name.value.decode()}

    # -- partitioned (multi-GPU) solve --
    def mg_config(self, rank, world):
        self._ck(self.lib.feab_mg_config(self.h, C.c_int(rank), C.c_int(world)))

    def mg_set_partition(self, part):
        part = np.ascontiguousarray(part, dtype=np.int32)
        self._ck(self.lib.feab_mg_set_partition(self.h, C.c_void_p(part.ctypes.data), C.c_int64(part.shape[0])))

    def mg_export(self) -> bytes:
        e = FeabMgExport()
        self._ck(self.lib.feab_mg_export(self.h, C.byref(e)))
        return bytes(e)

    def mg_connect(self, peer_rank, blob: bytes, same_process=False):
        e = FeabMgExport.from_buffer_copy(blob)
        self._ck(self.lib.feab_mg_connect(self.h, C.c_int(peer_rank), C.byref(e), C.c_int(1 if same_process else 0)))

    def mg_layout(self):
        a, b, c, d = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64()
        self._ck(self.lib.feab_mg_get_layout(self.h, C.byref(a), C.byref(b), C.byref(c), C.byref(d)))
        return {"n_own": a.value, "n_interior": b.value, "n_ghost": c.value, "n_cols": d.value}

    def mg_partition(self):
        out = np.empty(self.V, dtype=np.int32)
        self._ck(self.lib.feab_mg_get_partition(self.h, C.c_void_p(out.ctypes.data)))
        return out

    def mg_maps(self):
        lay = self.mg_layout()
        row_node = np.empty(max(lay["n_own"], 1), dtype=np.int32)
        col_unknown = np.empty(max(lay["n_cols"], 1), dtype=np.int32)
        self._ck(self.lib.feab_mg_get_maps(self.h, C.c_void_p(row_node.ctypes.data), C.c_void_p(col_unknown.ctypes.data)))
        return row_node[:lay["n_own"]], col_unknown[:lay["n_cols"]]


def partition_rcb(xyz, parts):
    """Recursive coordinate bisection (host C++, csrc/partition.cuh); needs no GPU."""
    lib = load_library()
    xyz = np.ascontiguousarray(xyz, dtype=np.float32)
    out = np.empty(xyz.shape[0], dtype=np.int32)
    rc = lib.feab_partition_rcb(C.c_int64(xyz.shape[0]), C.c_void_p(xyz.ctypes.data), C.c_int(parts), C.c_void_p(out.ctypes.data))
    if rc != 0:
        raise FeabError(rc, "feab_partition_rcb: bad arguments")
    return out


def group_step(handles, nsteps=1):
    """Lock-step advance of several partitioned handles living on one device in this process (virtual ranks)."""
    lib = load_library()
    arr = (C.c_void_p * len(handles))(*[h.h for h in handles])
    rc = lib.feab_mg_group_step(arr, C.c_int(len(handles)), C.c_int(nsteps))
    if rc != 0:
        raise FeabError(rc, lib.feab_last_error(handles[0].h).decode())
