"""CPU tests (no GPU): the oracle against the golden vectors the unmodified reference produced, against the
live compiled reference when oracle/_ref is present, and against closed-form finite-element facts."""
import numpy as np
import pytest

from fea_visualizer_b200 import meshes
from oracle import oracle as O
from tests.util import load_golden, rel_l2, run_steps, small_meshes

MATS = "MKA"


@pytest.fixture(scope="module")
def gold():
    return load_golden("ref_small.npz")


@pytest.mark.parametrize("name", ["ico4", "grid", "disk"])
@pytest.mark.parametrize("bc", [0, 1])
def test_oracle32_bit_exact_vs_reference_golden(gold, name, bc):
    """fp32 oracle == Eigen reference: idx_map, CSC pattern, values and 5-step fields, all bit-for-bit."""
    xyz, tri, ob = gold[f"{name}/xyz"], gold[f"{name}/tri"], gold[f"{name}/onb"]
    A = O.Oracle(xyz, tri, ob, bc, real=32)
    assert (A.idx_map() == gold[f"{name}/bc{bc}/idx_map"]).all()
    assert A.max_row_nonzeros() == int(gold[f"{name}/bc{bc}/max_row"])
    for w, mn in enumerate(MATS):
        outer, inner, val = A.csc(w)
        assert (outer == gold[f"{name}/bc{bc}/{mn}/outer"]).all()
        assert (inner == gold[f"{name}/bc{bc}/{mn}/inner"]).all()
        assert (val == gold[f"{name}/bc{bc}/{mn}/val"]).all(), mn
    for eq in range(4):
        got = run_steps(A, len(xyz), eq, 5)
        assert (got == gold[f"{name}/bc{bc}/eq{eq}/values5"]).all(), f"eq {eq}"
        assert (A.get_state(0) == gold[f"{name}/bc{bc}/eq{eq}/u"]).all()
        assert (A.get_state(1) == gold[f"{name}/bc{bc}/eq{eq}/v"]).all()


@pytest.mark.parametrize("name", ["ico4", "disk"])
def test_oracle64_close_to_reference_golden(gold, name):
    """fp64 instantiation of the same code agrees with the fp32 reference to fp32 rounding."""
    xyz, tri, ob = gold[f"{name}/xyz"], gold[f"{name}/tri"], gold[f"{name}/onb"]
    D = O.Oracle(xyz, tri, ob, 0, real=64)
    for w, mn in enumerate(MATS):
        outer, inner, val = D.csc(w)
        assert (inner == gold[f"{name}/bc0/{mn}/inner"]).all()
        ref = gold[f"{name}/bc0/{mn}/val"]
        assert np.abs(val - ref).max() <= 2e-6 * np.abs(ref).max()
    for eq in range(4):
        got = run_steps(D, len(xyz), eq, 5)
        assert rel_l2(got, gold[f"{name}/bc0/eq{eq}/values5"]) < 5e-6


def test_bundled_meshes_sizes_and_c1_heat():
    """SURVEY.md 8(c) structural invariants of the reference's bundled meshes + the C1 config (heat, 100 steps)."""
    g = load_golden("ref_assets.npz")
    expect = {("icosphere", 0): (10242, 7, 71682, 71682, 71682), ("tube", 0): (4512, 7, 31200, 31200, 31200),
              ("tube", 1): (4704, 7, 32544, 32544, 31200)}
    for (name, bc), sizes in expect.items():
        assert tuple(g[f"{name}/bc{bc}/sizes"]) == sizes
        xyz, tri, ob = g[f"{name}/xyz"], g[f"{name}/tri"].astype(np.uint32), g[f"{name}/onb"]
        A = O.Oracle(xyz, tri, ob, bc, real=32)
        got = (A.num_unknowns(), A.max_row_nonzeros()) + tuple(len(A.csc(w)[1]) for w in range(3))
        assert got == sizes
    # tube.obj is all quads: the triangle list in the fixture comes from this repo's fan split (tinyobjloader's rule,
    # Surface.cpp:63 `triangulate = true`, is not in the snapshot).  Sizes and the boundary count do not depend on the
    # diagonal chosen; the per-entry values of the tube fixtures do.
    assert int(g["tube/onb"].sum()) == 192
    xyz, tri, ob = g["icosphere/xyz"], g["icosphere/tri"].astype(np.uint32), g["icosphere/onb"]
    A = O.Oracle(xyz, tri, ob, 0, real=32)
    got = run_steps(A, len(xyz), 0, 100)
    assert (got == g["icosphere/bc0/eq0/values100"]).all()  # bit-exact after 100 Eigen PCG solves
    # mass matrix sums to the surface area (about 4 pi)
    assert abs(float(g["icosphere/bc0/sums"][0]) - 12.5626) < 1e-3


@pytest.mark.skipif(not O.reference_available(), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("bc", [0, 1])
def test_oracle32_bit_exact_vs_live_reference(bc):
    xyz, tri = meshes.icosphere(6, displace=True) if bc == 0 else meshes.disk(9)
    ob = meshes.boundary_flags(len(xyz), tri)
    R = O.Reference(xyz, tri, ob, bc)
    A = O.Oracle(xyz, tri, ob, bc, real=32)
    for X in (R, A):
        X.set_velocity([0.3, -1.0, 0.5])
        X.set_params(0, 0.02, 0.1)
        X.set_params(3, 0.002, 0.1, 0.05, 0.03, 0.055)
    for w in range(3):
        for a, b in zip(R.csc(w), A.csc(w)):
            assert (a == b).all()
    for eq in range(4):
        assert (run_steps(R, len(xyz), eq, 7) == run_steps(A, len(xyz), eq, 7)).all()
        assert R.unstable() == A.unstable()


def _dense(n, csc):
    return O.csc_to_dense(n, *csc)


def test_single_triangle_closed_form():
    """M = |J|/24 [2 1 1;1 2 1;1 1 2] (FEMContext.cpp:127-132); K from the cotangent formula."""
    xyz = np.array([[0, 0, 0], [2, 0, 0], [0, 0, 1]], dtype=np.float32)
    tri = np.array([[0, 2, 1]], dtype=np.uint32)
    D = O.Oracle(xyz, tri, np.zeros(3, np.uint8), 1, real=64)
    M, K = _dense(3, D.csc(0)), _dense(3, D.csc(1))
    area = 1.0
    assert np.allclose(M, 2 * area / 24 * np.array([[2, 1, 1], [1, 2, 1], [1, 1, 2]]), rtol=1e-14)
    p = xyz.astype(np.float64)
    G = np.zeros((3, 3))
    nrm = np.cross(p[2] - p[0], p[1] - p[0])
    nrm /= np.linalg.norm(nrm)
    for i in range(3):  # grad phi_i = n x (opposite edge) / (2 area) for the orientation (0, 2, 1)
        a, b = [0, 2, 1][(([0, 2, 1].index(i)) + 1) % 3], [0, 2, 1][(([0, 2, 1].index(i)) + 2) % 3]
        G[i] = np.cross(nrm, p[b] - p[a]) / (2 * area)
    Kx = area * G @ G.T
    assert np.allclose(K, Kx, atol=1e-14)


@pytest.mark.parametrize("name", ["ico4", "grid", "disk"])
def test_fem_identities(name):
    xyz, tri = small_meshes()[name]
    ob = meshes.boundary_flags(len(xyz), tri)
    D = O.Oracle(xyz, tri, ob, 1, real=64)  # Neumann: every node is an unknown
    n = D.num_unknowns()
    assert n == len(xyz)
    M, K = _dense(n, D.csc(0)), _dense(n, D.csc(1))
    assert np.abs(K.sum(axis=1)).max() < 1e-12 * np.abs(K).max()          # gradients of a partition of unity
    assert np.abs(K - K.T).max() == 0 and np.abs(M - M.T).max() == 0      # exactly symmetric
    p = xyz.astype(np.float64)
    area = 0.5 * np.linalg.norm(np.cross(p[tri[:, 1]] - p[tri[:, 0]], p[tri[:, 2]] - p[tri[:, 0]]), axis=1).sum()
    assert abs(M.sum() - area) < 1e-12 * area
    # pattern: nnz = V + 2E
    e = np.concatenate([tri[:, [0, 1]], tri[:, [1, 2]], tri[:, [2, 0]]]).astype(np.int64)
    e.sort(axis=1)
    E = len(np.unique(e[:, 0] * len(xyz) + e[:, 1]))
    assert len(D.csc(0)[1]) == n + 2 * E
    if ob.sum() == 0:  # closed mesh: advection rows sum to zero too
        A = _dense(n, D.csc(2))
        assert np.abs(A.sum(axis=1)).max() < 1e-12 * np.abs(A).max()


def test_explicit_zeros_are_kept():
    xyz, tri = meshes.grid(5, 4)
    D = O.Oracle(xyz, tri, meshes.boundary_flags(len(xyz), tri), 1, real=32)
    _, inner, val = D.csc(1)
    assert (val == 0).sum() > 0          # right-angle cotangents: structural entries whose value is exactly 0
    assert len(inner) == len(D.csc(0)[1])


def test_zero_rhs_short_circuit_and_gray_scott_feed():
    xyz, tri = meshes.icosphere(3)
    ob = meshes.boundary_flags(len(xyz), tri)
    D = O.Oracle(xyz, tri, ob, 0, real=64)
    for eq in (0, 1, 2):
        D.clear()
        D.set_equation(eq)
        D.step(3)
        assert (D.get_values() == 0).all() and D.last_stats()[0][0] == 0
    D.clear()
    D.set_equation(3)
    D.step(2)
    assert (D.get_state(0) > 0).all() and (D.get_values() == 0).all()   # u fed by f (1 - u); v stays 0


def test_dirichlet_nodes_forced_to_zero_and_instability_flag():
    xyz, tri = meshes.disk(5)
    ob = meshes.boundary_flags(len(xyz), tri)
    D = O.Oracle(xyz, tri, ob, 0, real=64)
    D.set_equation(0)
    D.set_values(np.full(len(xyz), 3.0, np.float32))
    D.step(1)
    v = D.get_values()
    assert (v[ob == 1] == 0).all() and (v[ob == 0] > 0).all()
    assert not D.unstable()
    D.set_values(np.full(len(xyz), 5e4, np.float32))
    D.step(1)
    assert D.unstable()
    D.set_values(np.full(len(xyz), -5e4, np.float32))   # the check is signed (CPUSolver.cpp:43)
    D.step(1)
    assert not D.unstable()


def test_adv_diff_parity_mode_is_cg_on_transpose_and_bicgstab_is_correct():
    """Fact 3 of SURVEY.md: Eigen's CG multiplies by A^T; its answer is not the solution of A x = b.
    The BiCGSTAB mode solves the true system."""
    xyz, tri = meshes.icosphere(6)
    ob = meshes.boundary_flags(len(xyz), tri)
    n = len(xyz)
    P = O.Oracle(xyz, tri, ob, 0, real=64)
    B = O.Oracle(xyz, tri, ob, 0, real=64)
    B.set_solver(mode=1)
    u0 = np.zeros(n, np.float32)
    u0[meshes.seed_nodes(n)] = 1.0
    M, K, A = (O.csc_to_dense(n, *P.csc(w)) for w in range(3))
    dt, c = np.float64(np.float32(0.001)), np.float64(np.float32(0.25))
    Aop = M / dt + c * K - A
    b = (M / dt) @ u0.astype(np.float64)
    for X in (P, B):
        X.set_equation(2)
        X.set_values(u0)
        X.step(1)
    xp, xb = P.get_state(0), B.get_state(0)
    res = lambda x: np.linalg.norm(b - Aop @ x) / np.linalg.norm(b)  # noqa: E731
    assert res(xb) < 1e-12
    assert res(xp) > 1e-4                      # reference behaviour: not a solution of the true system ...
    assert rel_l2(xp, xb) > 1e-4               # ... and visibly different from it


def test_unit_sphere_gray_scott_blows_up_in_the_reference_itself():
    """Why bench.py scales the icosphere to radius nu / 32 instead of SURVEY 8(d)'s unit sphere: on the
    unit sphere at ~1e5 vertices the UNMODIFIED reference's Gray-Scott step (reaction terms not mass-weighted,
    src/FEM/CPUSolver.cpp:134,137) leaves the seeded field growing without bound -- max|v| passes 1e3 after 5 steps and
    has_numerical_instability() (src/FEM/CPUSolver.cpp:41-48) fires on the 6th -- while the same mesh at the bundled
    icosphere.obj's element size stays bounded.  A unit-sphere benchmark would time Eigen's 2n-iteration NaN loop."""
    if not O.reference_available():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    for radius, blows in ((1.0, True), (100 / 32.0, False)):
        xyz, tri = meshes.icosphere(100, radius=radius)
        V = len(xyz)
        R = O.Reference(xyz, tri, np.zeros(V, np.uint8), 0)
        R.set_equation(3)
        seeds = meshes.seed_nodes(V)
        fired = None
        peak = 0.0
        for s in range(6):
            if s < 5:
                v = R.get_values()
                v[seeds] = 1.0
                R.set_values(v)
            R.step(1)
            peak = max(peak, float(np.nanmax(np.abs(R.get_values()))))
            if R.unstable():
                fired = s
                break
        if blows:
            assert fired is not None and fired <= 5 and peak > 1e3, (fired, peak)
        else:
            assert fired is None and peak <= 1.0, (fired, peak)
        R.close()


def _advdiff_system(D, u0):
    """A = M/dt + c K - Adv (src/FEM/CPUSolver.cpp:95) and b = (M/dt) u0 as fp64 CSR, from the oracle's own matrices."""
    import scipy.sparse as sp
    n = D.num_unknowns()
    mats = [sp.csc_matrix((v.astype(np.float64), i, o), shape=(n, n)) for o, i, v in (D.csc(w) for w in range(3))]
    dt, c = np.float64(np.float32(0.001)), np.float64(np.float32(0.25))
    Md = mats[0] / dt
    A = ((Md + c * mats[1]) - mats[2]).tocsr()
    A.sort_indices()
    return A, Md @ u0


@pytest.mark.parametrize("name,bc", [("ico8d", 0), ("disk14", 0), ("disk14", 1), ("grid", 0)])
def test_bicgstab_restatement_against_eigens_own_bicgstab(name, bc):
    """the non-symmetric mode is checked against Eigen's BiCGSTAB<SparseMatrix<double>> (Jacobi
    right-preconditioned, lib/eigen/Eigen/src/IterativeLinearSolvers/BiCGSTAB.h:32-119, compiled from the vendored tree into
    oracle/_ref/libeigen_bicgstab.so).  Pins oracle mode 1 (fea_oracle_impl.h FN(bicgstab)): same iterates up to the
    summation order of the dot products (Eigen reduces in SIMD packets), same iteration count within 1, both at eps."""
    if not O.eigen_bicgstab_available():
        pytest.skip("oracle/_ref/libeigen_bicgstab.so not built (needs /root/reference)")
    xyz, tri = {**small_meshes(), "ico8d": meshes.icosphere(8, displace=True), "disk14": meshes.disk(14)}[name]
    V = len(xyz)
    ob = meshes.boundary_flags(V, tri)
    D = O.Oracle(xyz, tri, ob, bc, real=64)
    D.set_solver(mode=1)
    D.set_equation(2)
    vals = np.zeros(V, np.float32)
    vals[meshes.seed_nodes(V)] = 1.0
    idx = D.idx_map()
    for step in range(3):
        u0 = np.zeros(D.num_unknowns())
        u0[idx[idx >= 0]] = vals[idx >= 0].astype(np.float64)
        A, b = _advdiff_system(D, u0)
        x, it, err, info = O.eigen_bicgstab(A.indptr, A.indices, A.data, b)
        D.set_values(vals)
        D.step(1)
        mine = D.get_state(0)
        oit, oerr = D.last_stats()
        assert err <= 2.3e-16 and oerr[0] <= 2.3e-16        # info() compares error() with the ABSOLUTE tolerance: not used
        assert rel_l2(mine, x) <= 1e-12, (step, rel_l2(mine, x))
        assert abs(int(oit[0]) - it) <= 1, (oit, it)
        assert np.linalg.norm(b - A @ x) / np.linalg.norm(b) <= 1e-13
        vals = D.get_values()
