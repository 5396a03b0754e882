"""ctypes binding of the Surface-side C ABI (include/feab_surface.h): the device mirror of the reference's `Surface`
(include/Utils/Surface.hpp:34-73) -- OBJ ingest + boundary detection, smooth normals, colour map, .ply exporter.
Method names follow the reference's (init_from_obj, export_to_ply, calculate_normals).  No CPU path behind it.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .feab import FeabError, load_library

MESH_OPEN, MESH_CLOSED, MESH_MIRRORED = 0, 1, 2     # include/Utils/Surface.hpp:25-29

# every symbol include/feab_surface.h declares (checked by tests/test_abi.py without a GPU)
ABI_SYMBOLS = [
    "feas_create", "feas_destroy", "feas_last_error", "feas_set_mesh", "feas_detect_boundary", "feas_load_obj", "feas_get_counts",
    "feas_get_mesh", "feas_set_values", "feas_get_values", "feas_set_values_device", "feab_values_device", "feas_set_color_map", "feas_color_map",
    "feas_smooth_normals", "feas_pick", "feas_brush", "feas_ply_build", "feas_ply_fetch", "feas_export_ply", "feas_write_ply", "feas_parse_obj", "feas_last_ms",
    "feas_detect_boundary_noalloc",
]


NOALLOC_MAX_TRIANGLES = 349525     # FEAS_NOALLOC_MAX_TRIANGLES, include/feab_surface.h


def _p(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


def detect_boundary_noalloc(device, stream, num_vertices, num_triangles, tri_device_ptr, on_boundary_device_ptr):
    """feas_detect_boundary_noalloc: boundary flags for a caller that owns its device buffers and must not allocate.
    `tri_device_ptr` / `on_boundary_device_ptr` are raw device addresses (ints) of memory the CALLER allocated, `stream`
    a raw cudaStream_t or 0.  Nothing is allocated on the way, here or in the library: the edge table is a fixed array
    inside libfeab (at most NOALLOC_MAX_TRIANGLES triangles).  Blocks until the flags are written; returns the number of
    boundary vertices.  Raises FeabError with the library's message otherwise."""
    lib = load_library()
    fn = lib.feas_detect_boundary_noalloc
    fn.restype = C.c_int
    fn.argtypes = [C.c_int, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.POINTER(C.c_int64), C.c_char_p, C.c_int64]
    nb = C.c_int64(-1)
    err = C.create_string_buffer(512)
    rc = fn(device, stream or None, num_vertices, num_triangles, tri_device_ptr or None, on_boundary_device_ptr or None, C.byref(nb), err, 512)
    if rc != 0:
        raise FeabError(rc, err.value.decode())
    return nb.value


def write_ply(path, positions, colors, tri):
    """Host-only: the ASCII writer of the exporter on ready-made arrays (no GPU, no handle)."""
    lib = load_library()
    pos = np.ascontiguousarray(positions, dtype=np.float32)
    col = np.ascontiguousarray(colors, dtype=np.float32)
    t = np.ascontiguousarray(tri, dtype=np.uint32)
    rc = lib.feas_write_ply(os.fsencode(path), C.c_int64(len(pos)), _p(pos), _p(col), C.c_int64(len(t)), _p(t))
    if rc != 0:
        raise FeabError(rc, "feas_write_ply failed")


def parse_obj(path):
    """Host-only OBJ reader with the semantics of Surface::init_from_obj: (xyz, normals, tri)."""
    lib = load_library()
    lib.feas_last_error.restype = C.c_char_p
    lib.feas_last_error.argtypes = [C.c_void_p]
    nv, nt = C.c_int64(), C.c_int64()
    rc = lib.feas_parse_obj(os.fsencode(path), C.byref(nv), C.byref(nt), None, None, None)
    if rc != 0:
        raise FeabError(rc, lib.feas_last_error(None).decode())
    xyz, nrm = np.empty((nv.value, 3), np.float32), np.empty((nv.value, 3), np.float32)
    tri = np.empty((nt.value, 3), np.uint32)
    rc = lib.feas_parse_obj(os.fsencode(path), None, None, _p(xyz), _p(nrm), _p(tri))
    if rc != 0:
        raise FeabError(rc, lib.feas_last_error(None).decode())
    return xyz, nrm, tri


class DeviceSurface:
    def __init__(self, device: int = 0):
        self.lib = load_library()
        self.lib.feas_last_error.restype = C.c_char_p
        self.lib.feas_last_error.argtypes = [C.c_void_p]
        self.h = C.c_void_p()
        rc = self.lib.feas_create(C.c_int(device), C.byref(self.h))
        if rc != 0:
            raise FeabError(rc, self.lib.feas_last_error(None).decode())

    def _ck(self, rc):
        if rc != 0:
            raise FeabError(rc, self.lib.feas_last_error(self.h).decode())

    def close(self):
        if getattr(self, "h", None):
            self.lib.feas_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- mesh --
    def set_mesh(self, xyz, normals, tri):
        xyz = np.ascontiguousarray(xyz, dtype=np.float32)
        tri = np.ascontiguousarray(tri, dtype=np.uint32)
        nrm = None if normals is None else np.ascontiguousarray(normals, dtype=np.float32)
        self._ck(self.lib.feas_set_mesh(self.h, C.c_int64(len(xyz)), _p(xyz), _p(nrm), C.c_int64(len(tri)), _p(tri)))

    def init_from_obj(self, path):
        self._ck(self.lib.feas_load_obj(self.h, os.fsencode(path)))

    def detect_boundary(self):
        V, _, _ = self.counts()
        out = np.empty(V, np.uint8)
        nb, closed = C.c_int64(), C.c_int()
        self._ck(self.lib.feas_detect_boundary(self.h, _p(out), C.byref(nb), C.byref(closed)))
        return out, nb.value, bool(closed.value)

    def counts(self):
        V, T, nb = C.c_int64(), C.c_int64(), C.c_int64()
        self._ck(self.lib.feas_get_counts(self.h, C.byref(V), C.byref(T), C.byref(nb)))
        return V.value, T.value, nb.value

    def get_mesh(self):
        V, T, _ = self.counts()
        xyz, nrm = np.empty((V, 3), np.float32), np.empty((V, 3), np.float32)
        tri, ob = np.empty((T, 3), np.uint32), np.empty(V, np.uint8)
        self._ck(self.lib.feas_get_mesh(self.h, _p(xyz), _p(nrm), _p(tri), _p(ob)))
        return xyz, nrm, tri, ob

    # -- values / colour --
    def set_values(self, v):
        v = np.ascontiguousarray(v, dtype=np.float32)
        assert len(v) == self.counts()[0]
        self._ck(self.lib.feas_set_values(self.h, _p(v)))

    def set_values_from_solver(self, feab):
        """device-to-device hand-over of the solver's surface->values mirror"""
        ptr = C.c_void_p()
        rc = self.lib.feab_values_device(feab.h, C.byref(ptr))
        if rc != 0:
            raise FeabError(rc, self.lib.feab_last_error(feab.h).decode())
        self._ck(self.lib.feas_set_values_device(self.h, ptr))

    def set_color_map(self, coeffs):
        c = np.ascontiguousarray(coeffs, dtype=np.float32).reshape(21)
        self._ck(self.lib.feas_set_color_map(self.h, _p(c)))

    def get_color(self, t):
        t = np.ascontiguousarray(t, dtype=np.float32)
        out = np.empty((len(t), 3), np.float32)
        self._ck(self.lib.feas_color_map(self.h, C.c_int64(len(t)), _p(t), _p(out)))
        return out

    # -- normals --
    def calculate_normals(self, vertex_extrusion):
        out = np.empty((self.counts()[0], 3), np.float32)
        self._ck(self.lib.feas_smooth_normals(self.h, C.c_float(vertex_extrusion), _p(out)))
        return out

    # -- picking / brush (BVH::ray_triangle_intersection, Application::brush) --
    def pick(self, origin, direction):
        o = np.ascontiguousarray(origin, dtype=np.float32)
        d = np.ascontiguousarray(direction, dtype=np.float32)
        tri, dist = C.c_int(-1), C.c_float(0.0)
        pt = np.zeros(3, np.float32)
        self._ck(self.lib.feas_pick(self.h, _p(o), _p(d), C.byref(tri), _p(pt), C.byref(dist)))
        return tri.value, pt, dist.value

    def brush(self, origin, direction, value):
        o = np.ascontiguousarray(origin, dtype=np.float32)
        d = np.ascontiguousarray(direction, dtype=np.float32)
        node = C.c_int(-1)
        self._ck(self.lib.feas_brush(self.h, _p(o), _p(d), C.c_float(value), C.byref(node)))
        return node.value

    def get_values(self):
        out = np.empty(self.counts()[0], np.float32)
        self._ck(self.lib.feas_get_values(self.h, _p(out)))
        return out

    # -- .ply --
    def ply_build(self, vertex_extrusion=0.25, threshold=0.0, mesh_type=MESH_OPEN, fetch=True):
        nv, nt = C.c_int64(), C.c_int64()
        self._ck(self.lib.feas_ply_build(self.h, C.c_float(vertex_extrusion), C.c_float(threshold), C.c_int(mesh_type), C.byref(nv), C.byref(nt)))
        if not fetch:
            return nv.value, nt.value
        pos, col = np.empty((nv.value, 3), np.float32), np.empty((nv.value, 3), np.float32)
        val, tri = np.empty(nv.value, np.float32), np.empty((nt.value, 3), np.uint32)
        self._ck(self.lib.feas_ply_fetch(self.h, _p(pos), _p(val), _p(col), _p(tri)))
        return pos, val, col, tri

    def export_to_ply(self, path, vertex_extrusion=0.25, threshold=0.0, mesh_type=MESH_OPEN):
        self._ck(self.lib.feas_export_ply(self.h, os.fsencode(path), C.c_float(vertex_extrusion), C.c_float(threshold), C.c_int(mesh_type)))

    def last_ms(self):
        ms, n = C.c_double(), C.c_int64()
        self._ck(self.lib.feas_last_ms(self.h, C.byref(ms), C.byref(n)))
        return ms.value, n.value
