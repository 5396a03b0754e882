"""CPU tests (no GPU) for the Surface-side rows: the compiled, unmodified reference
(oracle/_ref/libfea_surface_ref.so) reproduces the committed golden fixtures; the C ABI of include/feab_surface.h is
exported and matches the binding; the numpy restatement of the smooth-normals shader behaves."""
import os
import re

import numpy as np
import pytest

from fea_visualizer_b200 import build as fbuild
from fea_visualizer_b200 import feab, meshes, surface
from oracle import surface_oracle as S
from tests.util import load_golden

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASES = [(0.25, 0.0), (0.25, 0.1), (0.0, 0.3), (-0.5, 0.05), (1.5, -0.2)]   # tests/golden/make_golden_surface.py

needs_ref = pytest.mark.skipif(not S.available(), reason="oracle/_ref/libfea_surface_ref.so not built")


def header_symbols():
    text = open(os.path.join(ROOT, "include", "feab_surface.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(fea[bs]_[a-z_0-9]+)\s*\(", text)))


def test_surface_header_and_binding_agree():
    assert header_symbols() == sorted(surface.ABI_SYMBOLS)


def test_library_exports_every_surface_symbol():
    fbuild.build()
    lib = feab.load_library()
    for name in header_symbols():
        assert hasattr(lib, name), name


def test_golden_fixture_is_complete():
    g = load_golden("ref_surface.npz")
    for name in ("ico6d", "grid", "disk"):
        for mt in (0, 1, 2):
            for k in range(len(CASES)):
                b = g[f"ply_{name}_{mt}_{k}"].tobytes()
                assert b.startswith(b"ply\nformat ascii 1.0\nelement vertex ")
    assert tuple(g["obj_icosphere_counts"]) == (10242, 20480, 0, 1)      # SURVEY.md 8(c) invariants
    assert g["grid_onb"].sum() == 2 * (9 + 7) and g["ico6d_onb"].sum() == 0


@needs_ref
@pytest.mark.parametrize("name", ["ico6d", "grid", "disk"])
def test_compiled_reference_reproduces_golden_ply(name, tmp_path):
    g = load_golden("ref_surface.npz")
    R = S.RefSurface()
    R.set_mesh(g[f"{name}_xyz"], g[f"{name}_nrm"], g[f"{name}_tri"])
    R.set_values(g[f"{name}_val"])
    for mt in (0, 1, 2):
        for k, (ve, thr) in enumerate(CASES):
            p = str(tmp_path / "o.ply")
            R.export_to_ply(p, ve, thr, mt)
            assert open(p, "rb").read() == g[f"ply_{name}_{mt}_{k}"].tobytes(), (name, mt, k)


@needs_ref
def test_compiled_reference_color_maps_match_golden():
    g = load_golden("ref_surface.npz")
    for cm, coeffs in S.COLOR_MAPS.items():
        R = S.RefSurface()
        R.set_colormap(coeffs)
        assert (R.get_color(g[f"cm_{cm}_t"]).view(np.uint32) == g[f"cm_{cm}_rgb"].view(np.uint32)).all()


@needs_ref
def test_reference_boundary_detection_matches_edge_count_definition():
    # the definition SURVEY.md 8(c) states: a vertex is on the boundary iff it touches an edge used by exactly one triangle
    g = load_golden("ref_surface.npz")
    for name in ("ico6d", "grid", "disk"):
        assert (g[f"{name}_onb"] == meshes.boundary_flags(len(g[f"{name}_xyz"]), g[f"{name}_tri"])).all()


def test_ply_golden_counts_are_consistent():
    # every face index is < vertex count; Open exports never have more faces than 2 per input triangle
    g = load_golden("ref_surface.npz")
    for name in ("ico6d", "grid", "disk"):
        T = len(g[f"{name}_tri"])
        for mt in (0, 1, 2):
            for k in range(len(CASES)):
                path = os.path.join("/tmp", f"_g_{os.getpid()}.ply")
                open(path, "wb").write(g[f"ply_{name}_{mt}_{k}"].tobytes())
                v, f = S.parse_ply(path)
                os.remove(path)
                assert len(f) <= (2 if mt == 0 else 4) * T
                assert len(f) == 0 or f.max() < len(v)


def test_smooth_normals_restatement_on_a_sphere():
    xyz, tri = meshes.icosphere(6)
    nrm = xyz.copy()
    val = np.zeros(len(xyz), np.float32)
    out = S.smooth_normals(xyz, nrm, val, tri, 0.25)
    unit = out / np.linalg.norm(out, axis=1, keepdims=True)
    assert np.abs((unit * xyz).sum(1) - 1.0).max() < 2e-2     # area-weighted normals of a sphere point outward
    # extrusion is clamped to [0, 1] and ignores negative values (smooth_normals.glsl:48-50)
    out_neg = S.smooth_normals(xyz, nrm, val - 5.0, tri, 0.25)
    assert (out_neg == out).all()
    big = S.smooth_normals(xyz, nrm, val + 100.0, tri, 0.25)
    two = S.smooth_normals((2 * xyz).astype(np.float32), nrm, val, tri, 0.25)
    assert np.allclose(big, two, rtol=1e-5, atol=1e-7)


# ---------------------------------------------------------------------------------------- host halves of libfeab (no GPU)
@pytest.mark.parametrize("name", ["ico6d", "grid"])
def test_ascii_writer_reproduces_the_reference_bytes(name, tmp_path):
    """feas_write_ply (std::to_chars shortest form) on the arrays parsed back from the reference's own files gives the
    same bytes: the formatting half of the exporter is std::format("{}")-identical."""
    fbuild.build()
    g = load_golden("ref_surface.npz")
    for mt in (0, 2):
        for k in range(len(CASES)):
            ref = g[f"ply_{name}_{mt}_{k}"].tobytes()
            src = str(tmp_path / "ref.ply")
            open(src, "wb").write(ref)
            v, f = S.parse_ply(src)
            out = str(tmp_path / "ours.ply")
            surface.write_ply(out, v[:, :3], v[:, 3:], f)
            assert open(out, "rb").read() == ref, (name, mt, k)


def test_ascii_writer_float_forms(tmp_path):
    # shortest round-trip digits; fixed vs scientific as std::to_chars chooses; negative zero; uint32 range
    pos = np.array([[0.1, -0.0, 1e-5], [123456792.0, 1e10, 3.4028235e38], [1.0, 0.5, 1.1754944e-38]], np.float32)
    col = np.array([[1 / 3, 2 / 3, 1.0], [0.0, 100.0, 1e-3], [16777216.0, 0.30000001192092896, 7.0]], np.float32)
    tri = np.array([[0, 1, 2], [4294967295, 0, 7]], np.uint32)
    out = str(tmp_path / "f.ply")
    surface.write_ply(out, pos, col, tri)
    body = open(out, "rb").read().split(b"end_header\n")[1].decode().split("\n")
    assert body[0] == "0.1 -0 1e-05 0.33333334 0.6666667 1"
    assert body[1] == "123456792 1e+10 3.4028235e+38 0 100 0.001"      # fixed form prints the exact integer
    assert body[2] == "1 0.5 1.1754944e-38 16777216 0.3 7"
    assert body[3] == "3 0 1 2" and body[4] == "3 4294967295 0 7"
    v, _ = S.parse_ply(out)
    assert (v.view(np.uint32) == np.hstack([pos, col]).view(np.uint32)).all()      # every float round-trips


def test_obj_reader_matches_reference_ingest(tmp_path):
    from tests.golden.make_golden_surface import write_obj
    g = load_golden("ref_surface.npz")
    for name in ("ico6d", "disk"):
        p = str(tmp_path / (name + ".obj"))
        write_obj(p, g[f"{name}_xyz"], g[f"{name}_nrm"], g[f"{name}_tri"])
        xyz, nrm, tri = surface.parse_obj(p)
        assert (xyz.view(np.uint32) == g[f"{name}_xyz"].view(np.uint32)).all() and (tri == g[f"{name}_tri"]).all()
        assert (nrm.view(np.uint32) == g[f"{name}_objnrm"].view(np.uint32)).all()   # normalised sums, bit-exact vs the reference
    # quads are fan-triangulated, negative indices are relative, files without normals are rejected like the reference does
    q = tmp_path / "q.obj"
    q.write_text("v 0 0 0\nv 1 0 0\nv 1 0 1\nv 0 0 1\nvn 0 1 0\nf -4//1 -3//1 -2//1 -1//1\n")
    xyz, nrm, tri = surface.parse_obj(str(q))
    assert tri.tolist() == [[0, 1, 2], [0, 2, 3]] and (nrm == [0, 1, 0]).all()
    n = tmp_path / "n.obj"
    n.write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n")
    with pytest.raises(feab.FeabError, match="does not contain normals"):
        surface.parse_obj(str(n))


@needs_ref
def test_pslg_disk_generator_of_config_c4():
    """bench_configs.pslg_disk: the reference's Triangle hand-off ("Qeqpza0.004") on a regular polygon.  Triangle count follows
    SURVEY.md 8(d)'s T ~ pi R^2 / (0.6 area); the boundary flags Triangle's point markers give are the edge-count flags."""
    import bench_configs as B
    xyz, tri, ob = B.pslg_disk(radius=10.0, sides=512)
    T = len(tri)
    assert abs(T / (np.pi * 100.0 / (0.6 * 0.004)) - 1.0) < 0.10
    assert (xyz[:, 1] == 0).all() and tri.max() < len(xyz)
    assert ((ob != 0) == (meshes.boundary_flags(len(xyz), tri) != 0)).all() and int((ob != 0).sum()) >= 512     # Triangle may split long boundary edges
    # winding: Surface::perform_triangulation reverses Triangle's counter-clockwise order (src/Utils/Surface.cpp:467-469),
    # which makes the normals of the XZ-plane disk point along +y
    a, b, c = xyz[tri[:, 0]], xyz[tri[:, 1]], xyz[tri[:, 2]]
    ny = np.cross(b - a, c - a)[:, 1]
    assert (ny > 0).all() or (ny < 0).all()
