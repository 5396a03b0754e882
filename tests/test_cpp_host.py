"""The C++ host side of the boundary: include/FEM/CUDASolver.hpp (same interface as the reference's Solver,
include/FEM/Solver.hpp:8-15) driven by a headless C++ program the way Application drives CPUSolver, calling CUDA only
through the C ABI.  CPU part: it compiles against the mirror headers AND against the reference's own headers.
GPU part: its surface->values after every step equal the fp64 oracle's within the per-step bar (1e-8 relative L2)."""
import os
import shutil
import subprocess

import numpy as np
import pytest

from fea_visualizer_b200 import build as fbuild
from fea_visualizer_b200 import meshes
from tests.util import rel_l2

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"


def test_driver_builds_against_mirror_headers():
    exe = fbuild.build_cpp_host(force=True)
    assert os.path.exists(exe)
    out = subprocess.run(["ldd", exe], capture_output=True, text=True).stdout
    assert "libfeab.so" in out and "not found" not in out.split("libfeab.so")[1].split("\n")[0]


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "include", "FEM")), reason="reference tree not mounted")
def test_cuda_solver_header_compiles_against_reference_headers(tmp_path):
    """Inside the reference tree the reference's own FEM/Solver.hpp, FEMContext.hpp and Utils/Surface.hpp are found
    first; CUDASolver.hpp must compile against them unchanged."""
    src = tmp_path / "t.cpp"
    src.write_text('#include "FEM/CUDASolver.hpp"\n#include "Utils/CUDASurface.hpp"\n'
                   "void g(std::shared_ptr<Surface> s) { CUDASurface d(s); d.detect_boundary(); d.export_to_ply(\"x.ply\"); }\n"
                   "int main() { std::shared_ptr<FEMContext> c; std::shared_ptr<Solver> s = std::make_shared<CUDASolver>(c); return s ? 0 : 1; }\n")
    inc = [f"-I{REF}/include", f"-I{REF}/lib", f"-I{REF}/lib/glad/include", f"-I{REF}/lib/glm", f"-I{REF}/lib/eigen", f"-I{ROOT}/include"]
    out = subprocess.run(["g++", "-std=c++20", "-fpermissive", "-w", "-fsyntax-only", *inc, "-H", str(src)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr[-2000:]
    assert f"{REF}/include/FEM/FEMContext.hpp" in out.stderr      # really the reference's headers, not the mirror


def _write_mesh(path, xyz, tri, seeds):
    with open(path, "wb") as f:
        np.array([len(xyz), len(tri), len(seeds)], np.int64).tofile(f)
        np.ascontiguousarray(xyz, np.float32).tofile(f)
        np.ascontiguousarray(tri, np.uint32).tofile(f)
        np.ascontiguousarray(seeds, np.int64).tofile(f)


@pytest.mark.gpu
@pytest.mark.parametrize("name,bc", [("ico6", 0), ("disk8", 0), ("disk8", 1)])
@pytest.mark.parametrize("eq", [0, 1, 2, 3])
def test_cpp_cuda_solver_matches_oracle(tmp_path, name, bc, eq):
    from oracle import oracle as O
    exe = fbuild.CPP_DRIVER
    assert os.path.exists(exe), "tests/cpp/cuda_solver_driver not built (run __graft_entry__.build())"
    xyz, tri = meshes.icosphere(6) if name == "ico6" else meshes.disk(8)
    ob = meshes.boundary_flags(len(xyz), tri)
    V = len(xyz)
    seeds = meshes.seed_nodes(V)
    nsteps = 6
    _write_mesh(tmp_path / "m.bin", xyz, tri, seeds)
    out = subprocess.run([exe, str(tmp_path / "m.bin"), str(eq), str(bc), str(nsteps), "8", str(tmp_path / "o.bin")], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    raw = np.fromfile(tmp_path / "o.bin", dtype=np.float32)
    vals = raw[:nsteps * V].reshape(nsteps, V)
    unstable = raw[nsteps * V:].view(np.int32)
    D = O.Oracle(xyz, tri, ob, bc, real=64)
    D.clear()
    D.set_equation(eq)
    n = V - (int(ob.sum()) if bc == 0 else 0)
    assert f"unknowns={n} " in out.stdout
    for s in range(nsteps):
        if s == 0 or (eq == 3 and s < 5):
            v = D.get_values()
            v[seeds] = 1.0
            D.set_values(v)
        D.step(1)
        assert rel_l2(vals[s], D.get_values()) <= 1e-8, (s, rel_l2(vals[s], D.get_values()))
        assert bool(unstable[s]) == bool(D.unstable())


@pytest.mark.gpu
@pytest.mark.parametrize("mesh_type", [0, 2])
def test_cpp_export_to_ply_equals_reference(tmp_path, mesh_type):
    """The C++ host path end to end: CUDASolver steps, CUDASurface exports the field device-to-device; the file equals
    the one the unmodified reference writes for the same float values."""
    from oracle import surface_oracle as S
    if not S.available():
        pytest.skip("oracle/_ref/libfea_surface_ref.so not built")
    exe = fbuild.CPP_DRIVER
    xyz, tri = meshes.disk(8)
    V = len(xyz)
    nsteps = 4
    _write_mesh(tmp_path / "m.bin", xyz, tri, meshes.seed_nodes(V))
    ply = str(tmp_path / "o.ply")
    out = subprocess.run([exe, str(tmp_path / "m.bin"), "0", "1", str(nsteps), "8", str(tmp_path / "o.bin"), ply, "0.002", str(mesh_type)],
                         capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    vals = np.fromfile(tmp_path / "o.bin", dtype=np.float32)[:nsteps * V].reshape(nsteps, V)[-1]
    R = S.RefSurface()
    R.set_mesh(xyz, None, tri)
    R.set_values(vals)
    ref = str(tmp_path / "r.ply")
    R.export_to_ply(ref, 0.25, 0.002, mesh_type)
    got = open(ply, "rb").read()
    assert got == open(ref, "rb").read()
    assert int(got.split(b"element vertex ")[1].split(b"\n")[0]) > 0
