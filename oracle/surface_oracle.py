"""TEST INFRASTRUCTURE ONLY: ctypes binding of oracle/_ref/libfea_surface_ref.so -- the reference's UNMODIFIED
src/Utils/Surface.cpp + ColorMap.cpp behind ref_surface_harness.cpp -- plus a numpy restatement of the smooth-normals
compute shader (shaders/FEM/smooth_normals.glsl:30-57; GLSL cannot run here: "unavailable on this box", parity of that
one kernel is therefore UNPINNED and checked against this restatement only).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg import this module.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
MESH_OPEN, MESH_CLOSED, MESH_MIRRORED = 0, 1, 2     # include/Utils/Surface.hpp:25-29

# src/Utils/AssetStorage.cpp:141-200 (the double literals are narrowed to float by glm::vec3)
COLOR_MAPS = {
    "Viridis": [(0.274344, 0.004462, 0.331359), (0.108915, 1.397291, 1.388110), (-0.319631, 0.243490, 0.156419),
                (-4.629188, -5.882803, -19.646115), (6.181719, 14.388598, 57.442181), (4.876952, -13.955112, -66.125783),
                (-5.513165, 4.709245, 26.582180)],
    "Inferno": [(0.000129, 0.001094, -0.041044), (0.083266, 0.574933, 4.155398), (11.783686, -4.013093, -16.439814),
                (-42.246539, 17.689298, 45.210269), (78.087062, -33.838649, -83.264061), (-72.108852, 32.950143, 74.479447),
                (25.378501, -12.368929, -23.407604)],
    "Coolwarm": [(0.227376, 0.286898, 0.752999), (1.204846, 2.314886, 1.563499), (0.102341, -7.369214, -1.860252),
                 (2.218624, 32.578457, -1.643751), (-5.076863, -75.374676, -3.704589), (1.336276, 73.453060, 9.595678),
                 (0.694723, -25.863102, -4.558659)],
}


def lib_path():
    return os.path.join(HERE, "_ref", "libfea_surface_ref.so")


def available() -> bool:
    return os.path.exists(lib_path())


class RefSurface:
    """The reference's Surface object, headless."""

    def __init__(self):
        L = C.CDLL(lib_path())
        L.srf_create.restype = C.c_void_p
        L.srf_last_error.restype = C.c_char_p
        L.srf_last_error.argtypes = [C.c_void_p]
        L.srf_destroy.argtypes = [C.c_void_p]
        L.srf_set_colormap.argtypes = [C.c_void_p, C.c_void_p]
        L.srf_get_color.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
        L.srf_init_from_obj.argtypes = [C.c_void_p, C.c_char_p]
        L.srf_init_from_pslg.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_float]
        L.srf_set_mesh.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
        L.srf_counts.argtypes = [C.c_void_p] + [C.c_void_p] * 4
        L.srf_get_mesh.argtypes = [C.c_void_p] + [C.c_void_p] * 4
        L.srf_set_values.argtypes = [C.c_void_p, C.c_void_p]
        L.srf_export_to_ply.argtypes = [C.c_void_p, C.c_char_p, C.c_float, C.c_float, C.c_int]
        L.srf_bvh_build.argtypes = [C.c_void_p, C.c_int]
        L.srf_pick.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.srf_brush.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float]
        L.srf_get_values.argtypes = [C.c_void_p, C.c_void_p]
        self.L = L
        self.h = C.c_void_p(L.srf_create())

    def close(self):
        if self.h:
            self.L.srf_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_colormap(self, coeffs):
        c = np.ascontiguousarray(coeffs, dtype=np.float32).reshape(21)
        self.L.srf_set_colormap(self.h, c.ctypes.data)

    def get_color(self, t):
        t = np.ascontiguousarray(t, dtype=np.float32)
        out = np.empty((len(t), 3), np.float32)
        self.L.srf_get_color(self.h, len(t), t.ctypes.data, out.ctypes.data)
        return out

    def init_from_obj(self, path):
        if self.L.srf_init_from_obj(self.h, os.fsencode(path)) != 0:
            raise RuntimeError(self.L.srf_last_error(self.h).decode())

    def init_from_pslg(self, xz, triangle_area):
        xz = np.ascontiguousarray(xz, dtype=np.float32)
        if self.L.srf_init_from_pslg(self.h, len(xz), xz.ctypes.data, triangle_area) != 0:
            raise RuntimeError(self.L.srf_last_error(self.h).decode())

    def set_mesh(self, xyz, normals, tri, on_boundary=None):
        xyz = np.ascontiguousarray(xyz, dtype=np.float32)
        tri = np.ascontiguousarray(tri, dtype=np.uint32)
        nrm = None if normals is None else np.ascontiguousarray(normals, dtype=np.float32)
        ob = None if on_boundary is None else np.ascontiguousarray(on_boundary, dtype=np.uint8)
        self.L.srf_set_mesh(self.h, len(xyz), xyz.ctypes.data, None if nrm is None else nrm.ctypes.data, len(tri), tri.ctypes.data,
                            None if ob is None else ob.ctypes.data)

    def counts(self):
        V, T, nb = C.c_int64(), C.c_int64(), C.c_int64()
        closed = C.c_int()
        self.L.srf_counts(self.h, C.byref(V), C.byref(T), C.byref(nb), C.byref(closed))
        return V.value, T.value, nb.value, bool(closed.value)

    def get_mesh(self):
        V, T, _, _ = self.counts()
        xyz = np.empty((V, 3), np.float32)
        nrm = np.zeros((V, 3), np.float32)
        tri = np.empty((T, 3), np.uint32)
        ob = np.empty(V, np.uint8)
        self.L.srf_get_mesh(self.h, xyz.ctypes.data, nrm.ctypes.data, tri.ctypes.data, ob.ctypes.data)
        return xyz, nrm, tri, ob

    def set_values(self, v):
        v = np.ascontiguousarray(v, dtype=np.float32)
        assert len(v) == self.counts()[0]
        self.L.srf_set_values(self.h, v.ctypes.data)

    def bvh_build(self, max_depth=10):
        self.L.srf_bvh_build(self.h, max_depth)

    def pick(self, origin, direction):
        o = np.ascontiguousarray(origin, dtype=np.float32)
        d = np.ascontiguousarray(direction, dtype=np.float32)
        pt = np.zeros(3, np.float32)
        dist = C.c_float(0.0)
        t = self.L.srf_pick(self.h, o.ctypes.data, d.ctypes.data, pt.ctypes.data, C.addressof(dist))
        return t, pt, dist.value

    def brush(self, origin, direction, value):
        o = np.ascontiguousarray(origin, dtype=np.float32)
        d = np.ascontiguousarray(direction, dtype=np.float32)
        return self.L.srf_brush(self.h, o.ctypes.data, d.ctypes.data, value)

    def get_values(self):
        out = np.empty(self.counts()[0], np.float32)
        self.L.srf_get_values(self.h, out.ctypes.data)
        return out

    def export_to_ply(self, path, vertex_extrusion=0.25, threshold=0.0, mesh_type=MESH_OPEN):
        if self.L.srf_export_to_ply(self.h, os.fsencode(path), vertex_extrusion, threshold, mesh_type) != 0:
            raise RuntimeError(self.L.srf_last_error(self.h).decode())


def smooth_normals(xyz, normals, values, tri, vertex_extrusion):
    """shaders/FEM/smooth_normals.glsl:30-57 restated in fp32 numpy: every triangle adds the (un-normalised) cross
    product of its extruded edges to its three vertices.  The shader's float atomics make the summation order
    arbitrary; here it is triangle order (np.add.at is unbuffered and sequential)."""
    f = np.float32
    xyz = np.asarray(xyz, f)
    normals = np.asarray(normals, f)
    values = np.asarray(values, f)
    tri = np.asarray(tri, np.int64)
    s = np.minimum(f(1.0), f(vertex_extrusion) * np.maximum(f(0.0), values)).astype(f)
    pos = (xyz + (s[:, None] * normals).astype(f)).astype(f)
    a, b, c = pos[tri[:, 0]], pos[tri[:, 1]], pos[tri[:, 2]]
    e1, e2 = (b - a).astype(f), (c - a).astype(f)
    n = np.empty_like(e1)
    n[:, 0] = (e1[:, 1] * e2[:, 2]).astype(f) - (e2[:, 1] * e1[:, 2]).astype(f)
    n[:, 1] = (e1[:, 2] * e2[:, 0]).astype(f) - (e2[:, 2] * e1[:, 0]).astype(f)
    n[:, 2] = (e1[:, 0] * e2[:, 1]).astype(f) - (e2[:, 0] * e1[:, 1]).astype(f)
    out = np.zeros_like(xyz)
    for k in range(3):
        np.add.at(out, tri[:, k], n)
    return out


def parse_ply(path):
    """Reads back an ASCII .ply written by Surface::export_to_ply: (V x 6 float32 [x y z r g b], F x 3 uint32)."""
    with open(path, "rb") as f:
        data = f.read()
    head, _, body = data.partition(b"end_header\n")
    nv = nf = 0
    for line in head.split(b"\n"):
        if line.startswith(b"element vertex"):
            nv = int(line.split()[-1])
        if line.startswith(b"element face"):
            nf = int(line.split()[-1])
    lines = body.split(b"\n")
    verts = np.array([[np.float32(x) for x in ln.split()] for ln in lines[:nv]], dtype=np.float32).reshape(nv, 6)
    faces = np.array([[int(x) for x in ln.split()[1:]] for ln in lines[nv:nv + nf]], dtype=np.uint32).reshape(nf, 3)
    return verts, faces
