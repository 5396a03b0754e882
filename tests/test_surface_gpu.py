"""GPU parity tests of the Surface-side rows, through the C ABI of include/feab_surface.h.

Bars: .ply files BYTE-IDENTICAL to the ones the unmodified reference writes (golden fixtures + the compiled reference
on larger meshes); boundary flags, colour map, OBJ ingest bit-exact; smooth normals within 1e-5 relative of the
sequential restatement (the shader's float atomics leave the summation order open -- parity of that kernel is unpinned).
"""
import os

import numpy as np
import pytest

from fea_visualizer_b200 import feab, meshes
from fea_visualizer_b200.surface import MESH_CLOSED, MESH_MIRRORED, MESH_OPEN, DeviceSurface
from oracle import surface_oracle as S
from tests.golden.make_golden_surface import write_obj
from tests.util import load_golden

pytestmark = pytest.mark.gpu
CASES = [(0.25, 0.0), (0.25, 0.1), (0.0, 0.3), (-0.5, 0.05), (1.5, -0.2)]
needs_ref = pytest.mark.skipif(not S.available(), reason="oracle/_ref/libfea_surface_ref.so not built")


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


# ------------------------------------------------------------------------------------------------- N1: .ply exporter
@pytest.mark.parametrize("name", ["ico6d", "grid", "disk"])
@pytest.mark.parametrize("mt", [MESH_OPEN, MESH_CLOSED, MESH_MIRRORED])
def test_ply_bytes_equal_golden(name, mt, tmp_path):
    g = load_golden("ref_surface.npz")
    D = DeviceSurface()
    D.set_mesh(g[f"{name}_xyz"], g[f"{name}_nrm"], g[f"{name}_tri"])
    D.set_values(g[f"{name}_val"])
    for k, (ve, thr) in enumerate(CASES):
        p = str(tmp_path / "o.ply")
        D.export_to_ply(p, ve, thr, mt)
        assert open(p, "rb").read() == g[f"ply_{name}_{mt}_{k}"].tobytes(), (name, mt, k)


@needs_ref
@pytest.mark.parametrize("mt", [MESH_OPEN, MESH_CLOSED, MESH_MIRRORED])
def test_ply_bytes_equal_compiled_reference_100k(mt, tmp_path):
    xyz, tri = meshes.icosphere(100, displace=True)          # 100 002 vertices, 200 000 triangles
    nrm = (xyz / np.linalg.norm(xyz, axis=1, keepdims=True)).astype(np.float32)
    val = (np.sin(7 * xyz[:, 0]) * np.cos(5 * xyz[:, 1]) + 0.3 * np.sin(9 * xyz[:, 2])).astype(np.float32)
    R = S.RefSurface()
    R.set_mesh(xyz, nrm, tri)
    R.set_values(val)
    D = DeviceSurface()
    D.set_mesh(xyz, nrm, tri)
    D.set_values(val)
    for ve, thr in [(0.25, 0.2), (0.1, -5.0)]:       # a clipped field, and everything above the threshold
        pr, pd = str(tmp_path / "r.ply"), str(tmp_path / "d.ply")
        R.export_to_ply(pr, ve, thr, mt)
        D.export_to_ply(pd, ve, thr, mt)
        assert open(pr, "rb").read() == open(pd, "rb").read()


def test_ply_everything_below_threshold_is_empty(tmp_path):
    xyz, tri = meshes.icosphere(4)
    D = DeviceSurface()
    D.set_mesh(xyz, xyz, tri)
    nv, nt = D.ply_build(0.25, 0.5, MESH_OPEN, fetch=False)       # values are all 0 < 0.5
    assert (nv, nt) == (0, 0)
    p = str(tmp_path / "e.ply")
    D.export_to_ply(p, 0.25, 0.5, MESH_CLOSED)
    assert b"element vertex 0\n" in open(p, "rb").read() and b"element face 0\n" in open(p, "rb").read()


def test_ply_open_above_threshold_keeps_the_mesh():
    # size-independent property: with every value above the threshold an Open export is the extruded mesh itself --
    # one output vertex per referenced input vertex in first-use order, one output triangle per input triangle
    xyz, tri = meshes.icosphere(300)                               # 900 002 vertices
    nrm = xyz.copy()
    val = (0.5 + 0.25 * np.sin(3 * xyz[:, 0])).astype(np.float32)
    D = DeviceSurface()
    D.set_mesh(xyz, nrm, tri)
    D.set_values(val)
    pos, v, col, otri = D.ply_build(0.25, 0.0, MESH_OPEN)
    assert len(pos) == len(xyz) and len(otri) == len(tri)
    flat = tri.reshape(-1).astype(np.int64)
    _, first = np.unique(flat, return_index=True)
    order = flat[np.sort(first)]                                    # input vertices in order of first use
    assert (bits(v) == bits(val[order])).all()
    s = (val * np.float32(0.25)).astype(np.float32)
    ext = ((s[:, None] * nrm).astype(np.float32) + xyz).astype(np.float32)
    assert (bits(pos) == bits(ext[order])).all()
    inv = np.empty(len(xyz), np.int64)
    inv[order] = np.arange(len(xyz))
    assert (otri.astype(np.int64) == inv[tri.astype(np.int64)]).all()
    assert (bits(col) == bits(D.get_color(v))).all()


def test_ply_closed_doubles_the_surface():
    xyz, tri = meshes.icosphere(20)
    val = np.full(len(xyz), 0.5, np.float32)
    D = DeviceSurface()
    D.set_mesh(xyz, xyz, tri)
    D.set_values(val)
    pos, v, col, otri = D.ply_build(0.25, 0.1, MESH_CLOSED)
    assert len(otri) == 2 * len(tri) and len(pos) == 2 * len(xyz)
    # the projected copy carries the threshold as its value (Surface.cpp:264)
    assert np.isclose(np.sort(v)[: len(xyz)], 0.1).all() and np.isclose(np.sort(v)[len(xyz):], 0.5).all()


def test_values_hand_over_from_the_solver_is_zero_copy_equal():
    xyz, tri = meshes.icosphere(16)
    ob = meshes.boundary_flags(len(xyz), tri)
    F = feab.Feab(real=64)
    F.set_mesh(xyz, tri, ob)
    F.set_equation(0)
    v0 = np.zeros(len(xyz), np.float32)
    v0[meshes.seed_nodes(len(xyz))] = 1.0
    F.set_values(v0)
    F.step(3)
    A, B = DeviceSurface(), DeviceSurface()
    for X in (A, B):
        X.set_mesh(xyz, xyz, tri)
    A.set_values(F.get_values())
    B.set_values_from_solver(F)
    ra, rb = A.ply_build(0.25, 0.01, MESH_MIRRORED), B.ply_build(0.25, 0.01, MESH_MIRRORED)
    assert len(ra[0]) > 0
    for x, y in zip(ra, rb):
        assert (x.view(np.uint32) == y.view(np.uint32)).all()


# ------------------------------------------------------------------------------------------------- colour map
def test_color_map_bit_exact():
    g = load_golden("ref_surface.npz")
    D = DeviceSurface()
    for cm, coeffs in S.COLOR_MAPS.items():
        D.set_color_map(coeffs)
        assert (bits(D.get_color(g[f"cm_{cm}_t"])) == bits(g[f"cm_{cm}_rgb"])).all(), cm


# ------------------------------------------------------------------------------------------------- N2: ingest
@pytest.mark.parametrize("name", ["ico6d", "grid", "disk"])
def test_boundary_detection_equals_reference(name):
    g = load_golden("ref_surface.npz")
    D = DeviceSurface()
    D.set_mesh(g[f"{name}_xyz"], None, g[f"{name}_tri"])
    ob, nb, closed = D.detect_boundary()
    assert (ob == g[f"{name}_onb"]).all() and nb == int(g[f"{name}_onb"].sum()) and closed == (nb == 0)


def test_boundary_detection_odd_cases():
    # edge shared by 3 triangles (count 3: not a boundary edge), a degenerate triangle (edge a-a counts 2), an isolated vertex
    xyz = np.zeros((7, 3), np.float32)
    tri = np.array([[0, 1, 2], [1, 0, 3], [0, 1, 4], [5, 5, 6]], np.uint32)
    D = DeviceSurface()
    D.set_mesh(xyz, None, tri)
    ob, nb, closed = D.detect_boundary()
    assert (ob == meshes.boundary_flags(7, tri)).all()
    # edge 0-1 is used three times; every other edge of the first three triangles once -> 0..4 flagged via those;
    # triangle (5,5,6): edge 5-5 counts twice, edges 5-6 and 6-5 are the same pair used twice -> nothing flagged
    assert ob.tolist() == [1, 1, 1, 1, 1, 0, 0] and nb == 5 and not closed


def test_boundary_detection_large_mesh_properties():
    xyz, tri = meshes.icosphere(500)                                  # closed: no boundary
    D = DeviceSurface()
    D.set_mesh(xyz, None, tri)
    ob, nb, closed = D.detect_boundary()
    assert nb == 0 and closed and not ob.any()
    xyz, tri = meshes.grid(1500, 1000)                                # open: the rim
    D.set_mesh(xyz, None, tri)
    ob, nb, closed = D.detect_boundary()
    assert nb == 2 * (1500 + 1000) and not closed
    assert (ob == meshes.boundary_flags(len(xyz), tri)).all()


def test_obj_ingest_equals_reference(tmp_path):
    g = load_golden("ref_surface.npz")
    for name in ("ico6d", "disk"):
        p = str(tmp_path / (name + ".obj"))
        write_obj(p, g[f"{name}_xyz"], g[f"{name}_nrm"], g[f"{name}_tri"])
        D = DeviceSurface()
        D.init_from_obj(p)
        xyz, nrm, tri, ob = D.get_mesh()
        assert (bits(xyz) == bits(g[f"{name}_xyz"])).all() and (tri == g[f"{name}_tri"]).all()
        assert (bits(nrm) == bits(g[f"{name}_objnrm"])).all()
        assert (ob == g[f"{name}_onb"]).all() and D.counts()[2] == int(ob.sum())


def test_obj_without_normals_is_rejected(tmp_path):
    p = str(tmp_path / "n.obj")
    open(p, "w").write("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n")
    D = DeviceSurface()
    with pytest.raises(feab.FeabError, match="does not contain normals"):
        D.init_from_obj(p)
    with pytest.raises(feab.FeabError, match="valid file"):
        D.init_from_obj(str(tmp_path / "missing.obj"))


# ------------------------------------------------------------------------------------------------- N3: normals
@pytest.mark.parametrize("ve", [0.0, 0.25, 3.0])
def test_smooth_normals_vs_restatement(ve):
    xyz, tri = meshes.icosphere(24, displace=True)
    nrm = (xyz / np.linalg.norm(xyz, axis=1, keepdims=True)).astype(np.float32)
    val = (np.sin(4 * xyz[:, 0]) + 0.5 * xyz[:, 1]).astype(np.float32)
    D = DeviceSurface()
    D.set_mesh(xyz, nrm, tri)
    D.set_values(val)
    got = D.calculate_normals(ve)
    want = S.smooth_normals(xyz, nrm, val, tri, ve)
    scale = np.abs(want).max()
    assert np.abs(got - want).max() <= 1e-5 * scale     # summation order only


# ------------------------------------------------------------------------------------------------- N4: picking / brush
@pytest.mark.parametrize("name", ["ico6d", "disk"])
def test_pick_and_brush_equal_reference_bvh(name):
    g = load_golden("ref_surface.npz")
    D = DeviceSurface()
    D.set_mesh(g[f"{name}_xyz"], g[f"{name}_nrm"], g[f"{name}_tri"])
    o, d = g[f"pick_{name}_o"], g[f"pick_{name}_d"]
    hits = 0
    for i in range(len(o)):
        t, pt, dist = D.pick(o[i], d[i])
        assert t == int(g[f"pick_{name}_tri"][i]), i
        if t >= 0:
            hits += 1
            assert (bits(pt) == bits(g[f"pick_{name}_point"][i])).all() and bits([dist])[0] == bits([g[f"pick_{name}_dist"][i]])[0], i
        assert D.brush(o[i], d[i], 1.0 + i) == int(g[f"pick_{name}_node"][i]), i
    assert hits > 100
    assert (bits(D.get_values()) == bits(g[f"pick_{name}_values"])).all()      # the painted field, brush after brush


@needs_ref
def test_pick_equals_compiled_reference_200k_triangles():
    from tests.golden.make_golden_surface import pick_rays
    xyz, tri = meshes.icosphere(100, displace=True)
    R = S.RefSurface()
    R.set_mesh(xyz, xyz, tri)
    R.bvh_build(10)
    D = DeviceSurface()
    D.set_mesh(xyz, xyz, tri)
    o, d = pick_rays(xyz, n=300, seed=11)
    for i in range(len(o)):
        rt, rp, rd = R.pick(o[i], d[i])
        t, pt, dist = D.pick(o[i], d[i])
        assert t == rt, i
        if t >= 0:
            assert (bits(pt) == bits(rp)).all() and bits([dist])[0] == bits([rd])[0]
        assert D.brush(o[i], d[i], 0.5) == R.brush(o[i], d[i], 0.5)


def test_pick_ignores_hits_behind_the_origin_and_degenerate_triangles():
    xyz = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 2], [1, 0, 2], [0, 1, 2], [5, 5, 5]], np.float32)
    tri = np.array([[0, 1, 2], [3, 4, 5], [6, 6, 6]], np.uint32)                   # z = 0, z = 2, and a degenerate one
    D = DeviceSurface()
    D.set_mesh(xyz, None, tri)
    t, pt, dist = D.pick([0.2, 0.2, 1.0], [0, 0, 1])                               # from between the two: only z = 2 is ahead
    assert t == 1 and dist == 1.0 and pt.tolist() == [np.float32(0.2), np.float32(0.2), 2.0]
    assert D.pick([0.2, 0.2, 3.0], [0, 0, 1])[0] == -1                             # both behind
    assert D.pick([0.2, 0.2, -1.0], [0, 0, 1])[0] == 0                             # nearest of two
    assert D.pick([0.2, 0.2, -1.0], [1, 0, 0])[0] == -1                            # parallel: |denom| <= 1e-6
    assert D.brush([0.9, 0.05, -1.0], [0, 0, 1], 7.0) == 1                         # closest corner of triangle 0
    assert D.get_values()[1] == 7.0


# ------------------------------------------------------------------------------------------------- PSLG hand-off
@pytest.mark.parametrize("eq", [0, 2])
def test_reference_pslg_mesh_through_the_solver(eq):
    """A mesh produced by the reference's own PSLG hand-off (Surface::init_from_PSLG -> Triangle, winding reversed,
    point-marker boundary flags; fixture from the compiled reference) solved on the device: fp32 build vs the unmodified
    fp32 CPUSolver's field after 20 steps, fp64 build vs the fp64 oracle; the device boundary pass agrees with Triangle's
    point markers on this simply connected domain."""
    from oracle import oracle as O
    from tests.util import rel_l2
    g = load_golden("ref_surface.npz")
    xyz, tri, ob = g["pslg_xyz"], g["pslg_tri"], g["pslg_onb"]
    V = len(xyz)
    D = DeviceSurface()
    D.set_mesh(xyz, None, tri)
    assert (D.detect_boundary()[0] == ob).all()
    v0 = np.zeros(V, np.float32)
    v0[meshes.seed_nodes(V)] = 1.0
    out = {}
    for real in (32, 64):
        F = feab.Feab(real=real)
        F.set_mesh(xyz, tri, ob)
        F.clear()
        F.set_equation(eq)
        F.set_values(v0)
        F.step(20)
        out[real] = F.get_values()
    assert rel_l2(out[32], g[f"pslg_values_eq{eq}"]) <= 2e-5          # fp32 vs fp32 Eigen, 20 steps (drift bar of DESIGN.md 5)
    X = O.Oracle(xyz, tri, ob, 0, real=64)
    X.clear()
    X.set_equation(eq)
    X.set_values(v0)
    X.step(20)
    assert rel_l2(out[64], X.get_values()) <= 1e-7                     # fp64 vs fp64 oracle through float32 surface->values
