"""BASELINE.json configs C2, C4, C5 at (or near) their full sizes, through the C ABI.  Small-mesh parity against the
oracle is in test_parity_gpu.py; here the oracle is used where it still finishes in seconds (C2: every step; C4: two
steps) and size-independent properties carry the rest (mass conservation, true residuals, row sums, partition
invariants).  Tolerances are written next to each assertion."""
import json
import os

import numpy as np
import pytest

from fea_visualizer_b200 import feab, meshes
from oracle import oracle as O
from tests.util import rel_l2

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _seeded(V):
    v = np.zeros(V, np.float32)
    v[meshes.seed_nodes(V)] = 1.0
    return v


def _csr_matvec(rowptr, col, val, x):
    return np.add.reduceat(val * x[col], rowptr[:-1])


def test_c2_gray_scott_1000_steps_drift_on_bunny_surrogate():
    """configs[1]: Gray-Scott, 1000 steps, bunny-sized mesh.  The Stanford bunny is not in the reference repo (SURVEY.md
    8d), so this is the documented surrogate: icosphere nu=60 (36 002 V / 72 000 T) with radial displacement, scaled to
    the element size of the bundled icosphere.obj (radius nu/32) -- at unit radius the reference's own Gray-Scott step
    reaches NaN within 10 steps and Eigen's CG then spins through 2n iterations per solve (measured with the oracle).
    With the SURVEY's seed protocol the v field decays (u -> 1); parity is relative, so that does not matter.  Per-step
    bar 1e-8 relative L2 (fp64), drift after 10 / 100 / 1000 steps reported and bounded."""
    xyz, tri = meshes.icosphere(60, displace=True, radius=60 / 32.0)
    V = len(xyz)
    ob = meshes.boundary_flags(V, tri)
    seeds = meshes.seed_nodes(V)
    F = feab.Feab(real=64)
    F.set_mesh(xyz, tri, ob)
    D = O.Oracle(xyz, tri, ob, 0, real=64)
    for X in (F, D):
        X.clear()
        X.set_equation(3)
        X.set_values(_seeded(V))
    drift = {}
    done = 0
    for target in (1, 2, 3, 4, 5, 10, 100, 1000):
        for X in (F, D):
            if 0 < done < 5:                      # the brush is held for the first 5 steps
                v = X.get_values()
                v[seeds] = 1.0
                X.set_values(v)
            X.step(target - done)
        done = target
        drift[target] = max(rel_l2(F.get_values(), D.get_values()), rel_l2(F.get_state(0), D.get_state(0)),
                            rel_l2(F.get_state(1), D.get_state(1)))
        # never walk into a long run with a blown-up field: the oracle (like Eigen) iterates 2n times on NaN
        assert np.isfinite(D.get_state(0)).all() and not D.unstable() and not F.unstable(), target
        assert max(F.stats()["iterations"]) < 200, (target, F.stats())
    assert drift[1] <= 1e-8                                     # the north star's per-step bar
    assert drift[1000] <= 1e-6                                  # drift over 1000 steps (values are float32 in between)
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, "c2_drift.json"), "w") as f:
            json.dump({"mesh": "icosphere nu=60 displaced (bunny surrogate)", "V": V, "rel_l2_vs_fp64_oracle": drift}, f)


def test_c2_gray_scott_1000_steps_drift_of_a_live_pattern():
    """The drift of a field that has NOT decayed.  On the surrogate above the reference's reaction terms (not mass-weighted,
    src/FEM/CPUSolver.cpp:134,137) outweigh M/dt and v dies within a few dozen steps, so its 1000-step drift mostly measures
    agreement on a constant.  The same mesh at radius 30 (element size ~0.55: M_ii/dt >> f + k) keeps the seeded pattern
    alive -- after 1000 steps max v = 0.36, min u = 0.058 in the fp64 oracle -- and the drift bar is the same."""
    xyz, tri = meshes.icosphere(60, displace=True, radius=30.0)
    V = len(xyz)
    ob = meshes.boundary_flags(V, tri)
    seeds = meshes.seed_nodes(V)
    F = feab.Feab(real=64)
    F.set_mesh(xyz, tri, ob)
    D = O.Oracle(xyz, tri, ob, 0, real=64)
    for X in (F, D):
        X.clear()
        X.set_equation(3)
        X.set_values(_seeded(V))
    drift = {}
    done = 0
    for target in (1, 2, 3, 4, 5, 10, 100, 1000):
        for X in (F, D):
            if 0 < done < 5:
                v = X.get_values()
                v[seeds] = 1.0
                X.set_values(v)
            X.step(target - done)
        done = target
        drift[target] = max(rel_l2(F.get_values(), D.get_values()), rel_l2(F.get_state(0), D.get_state(0)),
                            rel_l2(F.get_state(1), D.get_state(1)))
        assert np.isfinite(D.get_state(0)).all() and not D.unstable() and not F.unstable(), target
    v_end, u_end = D.get_state(1), D.get_state(0)
    assert v_end.max() > 0.2 and u_end.min() < 0.2 and u_end.max() > 0.02      # a pattern, not a constant
    assert drift[1] <= 1e-8
    assert drift[1000] <= 1e-6
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, "c2_drift_live_pattern.json"), "w") as f:
            json.dump({"mesh": "icosphere nu=60 displaced, radius 30", "V": V, "v_max_after_1000": float(v_end.max()),
                       "u_min_after_1000": float(u_end.min()), "rel_l2_vs_fp64_oracle": drift}, f)


def test_c4_advection_diffusion_on_a_4m_triangle_disk():
    """configs[3]: advection-diffusion on a triangulated disk with ~4 M triangles, Dirichlet, non-symmetric operator.
    Reference-parity mode (PCG on A^T, ~950 iterations per step at this size) against the oracle at 1 M triangles
    (the CPU oracle needs ~40 s per step on the full mesh); at full size: convergence to Eigen's threshold in parity mode
    and the true residual of the BiCGSTAB mode."""
    xs, ts = meshes.disk(408)
    obs = meshes.boundary_flags(len(xs), ts)
    S = feab.Feab(real=64)
    S.set_mesh(xs, ts, obs)
    Ds = O.Oracle(xs, ts, obs, 0, real=64)
    for X in (S, Ds):
        X.clear()
        X.set_equation(2)
        X.set_values(_seeded(len(xs)))
        X.step(2)
    assert 0.9e6 < len(ts) < 1.1e6
    assert rel_l2(S.get_state(0), Ds.get_state(0)) <= 1e-8     # per-step bar, reference-parity mode
    assert abs(S.stats()["iterations"][0] - int(Ds.last_stats()[0][0])) <= 3
    S.close()
    xyz, tri = meshes.disk(816)
    V, T = len(xyz), len(tri)
    assert 3.9e6 < T < 4.1e6
    ob = meshes.boundary_flags(V, tri)
    v0 = _seeded(V)
    F = feab.Feab(real=64)
    F.set_solver(assembly=feab.ASSEMBLY_ORDERED)
    F.set_mesh(xyz, tri, ob)
    n = F.num_unknowns()
    assert n == V - int(ob.sum())
    # pattern invariants
    rowptr, col, M = F.csr(0)
    assert (np.diff(rowptr) >= 3).all() and np.all(col[rowptr[:-1]] <= np.arange(n))
    rp_a, col_a, A = F.csr(2)
    assert len(A) <= len(M)                                     # the advection filter drops boundary-adjacent entries
    # parity mode at full size: converges to Eigen's threshold (tol = eps), stays finite
    F.clear()
    F.set_equation(2)
    F.set_values(v0)
    F.step(2)
    st = F.stats()
    assert st["rel_residual"][0] <= 2.3e-16 and 0 < st["iterations"][0] < 5000
    assert np.isfinite(F.get_values()).all() and not F.unstable()
    # BiCGSTAB: solves the true system; check ||b - A x|| / ||b|| with A, b rebuilt on the host from the exported CSR
    F.set_solver(solver=feab.SOLVER_BICGSTAB)
    F.clear()
    F.set_equation(2)
    F.set_values(v0)
    idx = F.idx_map()
    u0 = np.zeros(n)
    u0[idx[idx >= 0]] = v0[idx >= 0].astype(np.float64)
    F.step(1)
    x = F.get_state(0)
    dt, c = np.float64(np.float32(0.001)), np.float64(np.float32(0.25))
    K = F.csr(1)[2]
    Ax = _csr_matvec(rowptr, col, M / dt + c * K, x) - _csr_matvec(rp_a, col_a, A, x)
    b = _csr_matvec(rowptr, col, M / dt, u0)
    true_res = float(np.linalg.norm(b - Ax) / np.linalg.norm(b))
    assert true_res <= 1e-10, true_res                        # recursive residual is at eps; the true one drifts with the iteration count
    st = F.stats()
    assert 0 < st["iterations"][0] < 5000 and st["rel_residual"][0] <= 2.3e-16


def test_c5_gray_scott_16m_vertices_properties():
    """configs[4], the benchmark mesh (icosphere nu=1265, 16 002 252 vertices): size-independent properties."""
    nu = 1265
    xyz, tri = meshes.icosphere(nu, radius=nu / 32.0)
    V = len(xyz)
    F = feab.Feab(real=64)
    F.set_mesh(xyz, tri, np.zeros(V, np.uint8))
    _, n, nnz, max_row = F.counts()
    assert (n, nnz, max_row) == (16002252, 7 * V - 12, 7)
    rowptr, col, M = F.csr(0)
    assert (np.diff(rowptr) >= 6).all()
    area = 4 * np.pi * (nu / 32.0) ** 2
    assert abs(M.sum() - area) / area < 1e-6                    # total mass = area of the sphere
    lumped = np.add.reduceat(M, rowptr[:-1])
    del M, col
    # heat step conserves 1^T M u on a closed surface
    v0 = _seeded(V)
    F.set_equation(0)
    F.set_values(v0)
    m0 = float(lumped @ v0.astype(np.float64))
    F.step(1)
    assert abs(float(lumped @ F.get_state(0)) - m0) <= 1e-9 * abs(m0)
    assert F.stats()["rel_residual"][0] < 2.3e-16
    # Gray-Scott: seeded protocol, both solves converge to Eigen's threshold, field stays finite and bounded
    F.clear()
    F.set_equation(3)
    F.set_values(v0)
    for s in range(6):
        if 0 < s < 5:
            F.brush(meshes.seed_nodes(V), 1.0)
        F.step(1)
        st = F.stats()
        assert max(st["rel_residual"]) < 2.3e-16 and 0 < min(st["iterations"]) and max(st["iterations"]) < 200
    vals = F.get_values()
    assert np.isfinite(vals).all() and not F.unstable() and 0 < float(vals.max()) <= 1.0
    # K steps in one device-resident call land on the same bits as K single steps (the bench's timed region)
    u_a = F.get_state(0).copy()
    F.step(3)
    G = feab.Feab(real=64)
    G.set_mesh(xyz, tri, np.zeros(V, np.uint8))
    G.set_equation(3)
    G.set_values(vals)
    G.set_state(0, u_a)
    for _ in range(3):
        G.step(1)
    assert rel_l2(G.get_values(), F.get_values()) <= 1e-6      # scatter assembly: atomic order differs between handles


# ------------------------------------------------------------------------------------------------- full size vs the oracle
def test_c3_wave_1m_vertices_two_steps_against_the_oracle():
    """configs[2] at its full size (icosphere nu = 316, 998 562 rows), wave equation, reference defaults: the field and the
    hidden velocity after each of two steps against the fp64 oracle.  Bar: 1e-8 relative L2 per step (the north star's)."""
    nu = 316
    xyz, tri = meshes.icosphere(nu, radius=nu / 32.0)
    V = len(xyz)
    ob = np.zeros(V, np.uint8)
    F = feab.Feab(real=64)
    F.set_mesh(xyz, tri, ob)
    D = O.Oracle(xyz, tri, ob, 0, real=64)
    for X in (F, D):
        X.clear()
        X.set_equation(1)
        X.set_values(_seeded(V))
    for s in range(2):
        F.step(1)
        D.step(1)
        e = [rel_l2(F.get_values(), D.get_values()), rel_l2(F.get_state(0), D.get_state(0)), rel_l2(F.get_state(1), D.get_state(1))]
        assert max(e) <= 1e-8, (s, e)
        assert abs(F.stats()["iterations"][0] - int(D.last_stats()[0][0])) <= 3


@pytest.mark.timeout(1500)
def test_c5_gray_scott_16m_vertices_one_step_against_the_oracle():
    """configs[4], the benchmark mesh itself (icosphere nu = 1265, 16 002 252 rows): one seeded Gray-Scott step through the
    production kernels (pcg_dual + pcg_persistent tail) against the fp64 oracle (about 80 s of CPU).  Bar: 1e-8 relative
    L2 for u, v and surface->values; iteration counts within 3 of the oracle's."""
    nu = 1265
    xyz, tri = meshes.icosphere(nu, radius=nu / 32.0)
    V = len(xyz)
    ob = np.zeros(V, np.uint8)
    F = feab.Feab(real=64)
    F.set_mesh(xyz, tri, ob)
    F.set_equation(3)
    F.clear()
    F.set_values(_seeded(V))
    F.step(1)
    st = F.stats()
    got = [F.get_values(), F.get_state(0), F.get_state(1)]
    assert st["shared_iterations"] > 0
    F.close()
    D = O.Oracle(xyz, tri, ob, 0, real=64)
    D.clear()
    D.set_equation(3)
    D.set_values(_seeded(V))
    D.step(1)
    e = [rel_l2(got[0], D.get_values()), rel_l2(got[1], D.get_state(0)), rel_l2(got[2], D.get_state(1))]
    assert max(e) <= 1e-8, e
    oit = D.last_stats()[0]
    assert abs(st["iterations"][0] - int(oit[0])) <= 3 and abs(st["iterations"][1] - int(oit[1])) <= 3, (st["iterations"], oit)
