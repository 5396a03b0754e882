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
