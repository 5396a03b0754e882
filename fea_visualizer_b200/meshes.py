"""Synthetic surface meshes and mesh ingest helpers (host side, numpy).

These produce exactly the plain data members of the reference's ``Surface`` that the hot path reads
(/root/reference/include/Utils/Surface.hpp:36-52): ``vertices`` (V x 3 float32), ``triangles``
(T x 3 uint32) and ``on_boundary`` (V uint8).  SURVEY.md section 8(d) specifies the generators.
"""
from __future__ import annotations

import numpy as np

# --- base icosahedron (unit sphere, outward winding) -------------------------------------------
_PHI = (1.0 + 5.0 ** 0.5) / 2.0
_ICO_V = np.array(
    [[-1, _PHI, 0], [1, _PHI, 0], [-1, -_PHI, 0], [1, -_PHI, 0],
     [0, -1, _PHI], [0, 1, _PHI], [0, -1, -_PHI], [0, 1, -_PHI],
     [_PHI, 0, -1], [_PHI, 0, 1], [-_PHI, 0, -1], [-_PHI, 0, 1]], dtype=np.float64)
_ICO_V /= np.linalg.norm(_ICO_V, axis=1, keepdims=True)
_ICO_F = np.array(
    [[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11],
     [1, 5, 9], [5, 11, 4], [11, 10, 2], [10, 7, 6], [7, 1, 8],
     [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9],
     [4, 9, 5], [2, 4, 11], [6, 2, 10], [8, 6, 7], [9, 8, 1]], dtype=np.int64)


def icosphere(nu: int, displace: bool = False, radius: float = 1.0):
    """Class-I geodesic subdivision of the icosahedron: every face split into nu^2 triangles,
    vertices normalised to radius 1 and rounded to float32 (the reference's storage type).

    V = 10 nu^2 + 2, T = 20 nu^2, E = 30 nu^2.  Deterministic vertex order: the 12 base vertices,
    then the (nu-1) interior points of each of the 30 edges by edge id (edges numbered in order of
    first appearance over faces, stored low->high vertex), then the face-interior points of each of
    the 20 faces by face id, row-major in the face's barycentric grid.  ``displace`` applies the
    "bunny surrogate" radial displacement r = 1 + 0.15 sin(3x) sin(4y) sin(5z).
    ``radius`` scales the sphere (in float64, before the float32 rounding): radius = nu / 32 keeps the element
    size of the reference's bundled icosphere.obj (nu = 32, radius 1) at any vertex count, which is what keeps
    the reference's non-mass-weighted Gray-Scott reaction terms stable on large meshes.
    """
    nu = int(nu)
    assert nu >= 1
    # edge table
    edge_id = {}
    edges = []
    for f in _ICO_F:
        for a, b in ((f[0], f[1]), (f[1], f[2]), (f[2], f[0])):
            key = (min(a, b), max(a, b))
            if key not in edge_id:
                edge_id[key] = len(edges)
                edges.append(key)
    n_e_int = nu - 1
    n_f_int = (nu - 1) * (nu - 2) // 2
    V = 12 + 30 * n_e_int + 20 * n_f_int
    pos = np.empty((V, 3), dtype=np.float64)
    pos[:12] = _ICO_V
    t = np.arange(1, nu, dtype=np.float64)[:, None]
    for e, (a, b) in enumerate(edges):
        base = 12 + e * n_e_int
        pos[base:base + n_e_int] = ((nu - t) * _ICO_V[a] + t * _ICO_V[b]) / nu
    face_base0 = 12 + 30 * n_e_int

    def edge_vertex(a, b, k):
        """global index of the k-th point (k = 0..nu, vectorised) walking from base vertex a to b"""
        k = np.asarray(k)
        lo, hi = (a, b) if a < b else (b, a)
        e = edge_id[(lo, hi)]
        kk = k if a < b else nu - k
        out = 12 + e * n_e_int + (kk - 1)
        out = np.where(kk == 0, lo, out)
        out = np.where(kk == nu, hi, out)
        return out

    tris = []
    # barycentric grid of a face (A, B, C): point (i, j) = ((nu-i-j) A + i B + j C) / nu, i + j <= nu
    ii, jj = np.meshgrid(np.arange(nu + 1), np.arange(nu + 1), indexing="ij")
    valid = (ii + jj) <= nu
    for f, (A, B, C) in enumerate(_ICO_F):
        idx = np.full((nu + 1, nu + 1), -1, dtype=np.int64)
        # interior: i >= 1, j >= 1, i + j <= nu - 1, row-major over (j, i)
        if n_f_int:
            jn, in_ = np.meshgrid(np.arange(1, nu), np.arange(1, nu), indexing="ij")
            m = (in_ + jn) <= nu - 1
            iv, jv = in_[m], jn[m]
            ids = face_base0 + f * n_f_int + np.arange(n_f_int)
            idx[iv, jv] = ids
            p = ((nu - iv - jv)[:, None] * _ICO_V[A] + iv[:, None] * _ICO_V[B] + jv[:, None] * _ICO_V[C]) / nu
            pos[ids] = p
        k = np.arange(nu + 1)
        idx[k, 0] = edge_vertex(A, B, k)          # j = 0 : A -> B
        idx[0, k] = edge_vertex(A, C, k)          # i = 0 : A -> C
        idx[nu - k, k] = edge_vertex(B, C, k)     # i + j = nu : B -> C
        # upward triangles (i,j),(i+1,j),(i,j+1) for i + j <= nu - 1
        i0, j0 = ii[:-1, :-1], jj[:-1, :-1]
        mu = (i0 + j0) <= nu - 1
        up = np.stack([idx[i0[mu], j0[mu]], idx[i0[mu] + 1, j0[mu]], idx[i0[mu], j0[mu] + 1]], axis=1)
        md = (i0 + j0) <= nu - 2
        dn = np.stack([idx[i0[md] + 1, j0[md]], idx[i0[md] + 1, j0[md] + 1], idx[i0[md], j0[md] + 1]], axis=1)
        tris.append(up)
        tris.append(dn)
    del valid
    tri = np.concatenate(tris, axis=0)
    pos /= np.linalg.norm(pos, axis=1, keepdims=True)
    if displace:
        r = 1.0 + 0.15 * np.sin(3 * pos[:, 0]) * np.sin(4 * pos[:, 1]) * np.sin(5 * pos[:, 2])
        pos *= r[:, None]
    if radius != 1.0:
        pos *= float(radius)
    assert tri.min() >= 0 and tri.shape[0] == 20 * nu * nu
    return np.ascontiguousarray(pos, dtype=np.float32), np.ascontiguousarray(tri, dtype=np.uint32)


def boundary_flags(num_vertices: int, tri: np.ndarray) -> np.ndarray:
    """A vertex is on the boundary iff it touches an edge used by exactly one triangle
    (/root/reference/src/Utils/Surface.cpp:104-122)."""
    t = tri.astype(np.int64)
    e = np.concatenate([t[:, [0, 1]], t[:, [1, 2]], t[:, [2, 0]]], axis=0)
    e.sort(axis=1)
    key = e[:, 0] * np.int64(num_vertices) + e[:, 1]
    uniq, inv = np.unique(key, return_inverse=True)
    # a degenerate edge (a, a) is incremented twice by the reference (edges[a][a]++ twice), so it never counts as 1
    counts = np.bincount(inv.reshape(-1), weights=np.where(e[:, 0] == e[:, 1], 2, 1), minlength=len(uniq))
    b = uniq[counts == 1]
    flags = np.zeros(num_vertices, dtype=np.uint8)
    flags[b // num_vertices] = 1
    flags[b % num_vertices] = 1
    return flags


def grid(nx: int, nz: int, lx: float = 1.0, lz: float = 1.0):
    """Open planar rectangle in the XZ plane (y = 0, like the reference's PSLG surfaces,
    Surface.cpp:455-461), (nx+1) x (nz+1) vertices, right-angled triangles.  The right angles give
    exactly-zero stiffness off-diagonals, which exercises 'explicit zeros are kept'."""
    xs = np.linspace(0.0, lx, nx + 1)
    zs = np.linspace(0.0, lz, nz + 1)
    X, Z = np.meshgrid(xs, zs, indexing="ij")
    pos = np.stack([X.ravel(), np.zeros(X.size), Z.ravel()], axis=1)
    vid = np.arange((nx + 1) * (nz + 1)).reshape(nx + 1, nz + 1)
    a, b, c, d = vid[:-1, :-1].ravel(), vid[1:, :-1].ravel(), vid[1:, 1:].ravel(), vid[:-1, 1:].ravel()
    tri = np.concatenate([np.stack([a, c, b], 1), np.stack([a, d, c], 1)], axis=0)
    return pos.astype(np.float32), tri.astype(np.uint32)


def disk(rings: int, radius: float = 1.0):
    """Open planar disk in the XZ plane: a centre vertex plus `rings` concentric rings with 6 r
    vertices on ring r (V = 1 + 3 rings (rings+1), T = 6 rings^2).  A structured stand-in for the
    reference's PSLG -> Triangle disks."""
    pts = [np.zeros((1, 2))]
    start = [0, 1]
    for r in range(1, rings + 1):
        th = 2 * np.pi * np.arange(6 * r) / (6 * r)
        pts.append(radius * r / rings * np.stack([np.cos(th), np.sin(th)], 1))
        start.append(start[-1] + 6 * r)
    p2 = np.concatenate(pts, 0)
    tris = []
    for r in range(1, rings + 1):
        o, i0 = start[r], start[r - 1]
        no, ni = 6 * r, max(6 * (r - 1), 1)
        for s in range(6):
            for k in range(r):
                a = o + (s * r + k) % no
                b = o + (s * r + k + 1) % no
                c = i0 + ((s * (r - 1) + k) % ni if r > 1 else 0)
                tris.append((a, b, c))
                if k < r - 1:
                    d = i0 + (s * (r - 1) + k + 1) % ni
                    tris.append((c, b, d))
    pos = np.stack([p2[:, 0], np.zeros(len(p2)), p2[:, 1]], 1)
    return pos.astype(np.float32), np.asarray(tris, dtype=np.uint32)


def fan_refined(xyz: np.ndarray, tri: np.ndarray):
    """Every triangle (a, b, c) of a base mesh split at its centroid m into (a, b, m), (b, c, m), (c, a, m); the centroids
    are appended AFTER the base vertices.  A base vertex of valence d gets valence 2 d, so on a regular base mesh the
    first rows of the system have 13 entries each while the centroid rows have 4: consecutive 256 / 512-row tiles of the
    base block hold more than 8 entries per row, which is what pushes the tile kernels (spmv_dot_tma, rhs_init_tma,
    pcg_persistent, pcg_dual) onto their un-staged path (tile entries > the staging capacity)."""
    xyz = np.asarray(xyz, dtype=np.float32)
    t = np.asarray(tri, dtype=np.int64)
    V = len(xyz)
    cen = (xyz[t[:, 0]].astype(np.float64) + xyz[t[:, 1]] + xyz[t[:, 2]]) / 3.0
    m = V + np.arange(len(t), dtype=np.int64)
    out = np.concatenate([np.stack([t[:, 0], t[:, 1], m], 1), np.stack([t[:, 1], t[:, 2], m], 1), np.stack([t[:, 2], t[:, 0], m], 1)], axis=0)
    return (np.ascontiguousarray(np.concatenate([xyz, cen.astype(np.float32)], 0), dtype=np.float32),
            np.ascontiguousarray(out, dtype=np.uint32))


def load_obj(path: str):
    """Minimal OBJ ingest following /root/reference/src/Utils/Surface.cpp:57-102: vertices in file
    order, faces as vertex-index triples, polygons fan-triangulated, a file without 'vn' lines is
    rejected like the reference does (:73-74)."""
    verts, faces, has_vn = [], [], False
    with open(path) as fh:
        for line in fh:
            if line.startswith("v "):
                verts.append([float(x) for x in line.split()[1:4]])
            elif line.startswith("vn "):
                has_vn = True
            elif line.startswith("f "):
                ids = [int(tok.split("/")[0]) for tok in line.split()[1:]]
                ids = [i - 1 if i > 0 else len(verts) + i for i in ids]
                for k in range(1, len(ids) - 1):
                    faces.append((ids[0], ids[k], ids[k + 1]))
    if not verts:
        raise RuntimeError(f"File {path} does not contain a valid .obj mesh.")
    if not has_vn:
        raise RuntimeError(f"File {path} does not contain normals!")
    return np.asarray(verts, dtype=np.float32), np.asarray(faces, dtype=np.uint32)


def seed_nodes(num_vertices: int, count: int = 50) -> np.ndarray:
    """Brush seed nodes i = (k * 2654435761) mod V, k = 0..count-1."""
    k = np.arange(count, dtype=np.uint64)
    return ((k * np.uint64(2654435761)) % np.uint64(num_vertices)).astype(np.int64)
