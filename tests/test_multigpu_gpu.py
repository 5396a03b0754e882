"""Partitioned (multi-GPU) solve, SURVEY.md 8(e), exercised with VIRTUAL RANKS: P partitioned handles on one B200,
advanced in lock-step (feab_mg_group_step).  Same numbering, kernels, halo pushes and mailbox all-reduces as the
one-process-per-GPU run; only the waiting is replaced by stream order.  Bars: layout / pattern / index work bit-exact;
fields vs the single-GPU solver within reduction-order tolerance (1e-11), vs the fp64 oracle within 1e-8 per step.
"""
import numpy as np
import pytest

from fea_visualizer_b200 import feab, meshes
from fea_visualizer_b200.multigpu import VirtualRanks
from oracle import oracle as O
from tests.util import rel_l2, run_steps

pytestmark = pytest.mark.gpu


def mesh(name):
    if name == "ico12":
        xyz, tri = meshes.icosphere(12)
    elif name == "ico8d":
        xyz, tri = meshes.icosphere(8, displace=True)
    elif name == "disk14":
        xyz, tri = meshes.disk(14)
    else:
        xyz, tri = meshes.grid(15, 11)
    return xyz, tri, meshes.boundary_flags(len(xyz), tri)


def single(real, xyz, tri, ob, bc, **solver):
    F = feab.Feab(real=real)
    F.set_boundary_condition(bc)
    if solver:
        F.set_solver(**solver)
    F.set_mesh(xyz, tri, ob)
    return F


def virtual(world, real, xyz, tri, ob, bc, part=None, **solver):
    G = VirtualRanks(world, real=real, part=part, **solver)
    for h in G.hs:
        h.set_boundary_condition(bc)
    G.set_mesh(xyz, tri, ob)
    return G


@pytest.mark.parametrize("name,bc", [("ico12", 0), ("disk14", 0), ("disk14", 1), ("grid", 0)])
@pytest.mark.parametrize("world", [2, 3, 8])
def test_partition_layout_and_matrices(name, bc, world):
    xyz, tri, ob = mesh(name)
    F = single(64, xyz, tri, ob, bc, assembly=feab.ASSEMBLY_ORDERED)
    G = virtual(world, 64, xyz, tri, ob, bc, assembly=feab.ASSEMBLY_ORDERED)
    idx = F.idx_map()
    n = F.num_unknowns()
    part = feab.partition_rcb(xyz, world)
    seen = np.zeros(len(xyz), int)
    for r, (h, (row_node, col_unknown)) in enumerate(zip(G.hs, G.rows)):
        lay = h.mg_layout()
        assert (h.idx_map() == idx).all()
        seen[row_node] += 1
        assert (part[row_node] == r).all() and (idx[row_node] >= 0).all()            # rows = the unknowns this rank owns
        assert lay["n_own"] == len(row_node) and lay["n_cols"] == len(col_unknown)
        # interior rows first, each group in ascending global order
        g = idx[row_node]
        ni = lay["n_interior"]
        assert (np.diff(g[:ni]) > 0).all() and (np.diff(g[ni:]) > 0).all()
        # ghost columns belong to other ranks; interior rows never touch them
        ghosts = col_unknown[lay["n_own"]:]
        ghosts = ghosts[ghosts >= 0]
        assert len(ghosts) == lay["n_ghost"]
        node_of = np.flatnonzero(idx >= 0)
        assert (part[node_of[ghosts]] != r).all()
        rowptr, col, _ = h.csr(0)
        if ni:
            assert col[:rowptr[ni]].max() < lay["n_own"]
        assert (np.diff(np.concatenate([col[rowptr[i]:rowptr[i + 1]] for i in range(min(50, len(row_node)))])) != 0).all()
    assert (seen[idx >= 0] == 1).all() and (seen[idx < 0] == 0).all()               # every unknown owned exactly once
    # the global matrices assembled from the pieces are the single-GPU matrices, bit for bit
    for which in range(3):
        rowptr, col, val = F.csr(which)
        A = np.zeros((n, n))
        P = np.zeros((n, n), bool)
        rows = np.repeat(np.arange(n), np.diff(rowptr))
        A[rows, col] = val
        P[rows, col] = True
        GA, GP = G.global_csr_dense(which)
        assert (GP == P).all(), which
        assert (GA == A).all(), which


@pytest.mark.parametrize("name,bc", [("ico12", 0), ("disk14", 0), ("disk14", 1)])
@pytest.mark.parametrize("world", [2, 4, 7])
@pytest.mark.parametrize("eq", [0, 1, 2, 3])
def test_virtual_ranks_step_parity(name, bc, world, eq):
    xyz, tri, ob = mesh(name)
    V = len(xyz)
    F = single(64, xyz, tri, ob, bc)
    G = virtual(world, 64, xyz, tri, ob, bc)
    D = O.Oracle(xyz, tri, ob, bc, real=64)
    ref = []
    run_steps(F, V, eq, 5, per_step=lambda s, X: ref.append((X.get_values(), X.get_state(0), X.get_state(1), X.stats()["iterations"])))
    orc = []
    run_steps(D, V, eq, 5, per_step=lambda s, X: orc.append((X.get_values(), X.get_state(0), X.get_state(1))))
    worst = [0.0, 0.0]

    def check(s, X):
        v, u, vv, it = ref[s]
        worst[0] = max(worst[0], rel_l2(X.get_values(), v), rel_l2(X.get_state(0), u), rel_l2(X.get_state(1), vv))
        worst[1] = max(worst[1], rel_l2(X.get_values(), orc[s][0]), rel_l2(X.get_state(0), orc[s][1]), rel_l2(X.get_state(1), orc[s][2]))
        git = X.stats()["iterations"]
        assert abs(git[0] - it[0]) <= 2 and abs(git[1] - it[1]) <= 2, (git, it)

    run_steps(G, V, eq, 5, per_step=check)
    assert worst[0] <= 1e-11, worst      # vs single GPU: summation order only
    assert worst[1] <= 1e-8, worst       # vs the fp64 oracle: the north star's per-step bar
    assert G.unstable() == F.unstable()


def test_worst_case_partition_every_row_is_boundary():
    """part = node % P: no interior rows at all, every neighbour pair exchanges halos."""
    xyz, tri, ob = mesh("ico8d")
    V = len(xyz)
    part = (np.arange(V) % 4).astype(np.int32)
    G = virtual(4, 64, xyz, tri, ob, 0, part=part)
    assert all(h.mg_layout()["n_interior"] == 0 for h in G.hs)
    F = single(64, xyz, tri, ob, 0)
    for eq in (0, 3):
        a = run_steps(F, V, eq, 4)
        b = run_steps(G, V, eq, 4)
        assert rel_l2(b, a) <= 1e-6          # values are float32
        assert rel_l2(G.get_state(0), F.get_state(0)) <= 1e-11


def test_virtual_ranks_fp32_and_multi_step_call():
    xyz, tri, ob = mesh("ico12")
    V = len(xyz)
    F = single(32, xyz, tri, ob, 0)
    G = virtual(3, 32, xyz, tri, ob, 0)
    a = run_steps(F, V, 3, 5)
    b = run_steps(G, V, 3, 5)
    assert rel_l2(b, a) <= 5e-6
    # several steps inside one call == the same steps one by one (state halos are pushed between steps)
    v0 = np.zeros(V, np.float32)
    v0[meshes.seed_nodes(V)] = 1.0
    ordered = dict(assembly=feab.ASSEMBLY_ORDERED)     # scatter assembly sums the diagonal in atomic arrival order
    A, B = virtual(2, 64, xyz, tri, ob, 0, **ordered), virtual(2, 64, xyz, tri, ob, 0, **ordered)
    for X in (A, B):
        X.set_equation(1)
        X.set_values(v0)
    A.step(4)
    for _ in range(4):
        B.step(1)
    assert (A.get_state(0) == B.get_state(0)).all() and (A.get_state(1) == B.get_state(1)).all()


def test_unconnected_partition_and_lockstep_bicgstab_are_refused():
    """BiCGSTAB on a partition runs through the persistent kernel only (bicg_persistent.cuh has the halo exchange); the
    lock-step group driver and the WHILE-graph kernels refuse it."""
    xyz, tri, ob = mesh("grid")
    h = feab.Feab(real=64)
    h.mg_config(0, 2)
    h.set_mesh(xyz, tri, ob)
    with pytest.raises(feab.FeabError):
        h.step(1)
    G = virtual(2, 64, xyz, tri, ob, 0, solver=feab.SOLVER_BICGSTAB)
    G.set_equation(2)
    with pytest.raises(feab.FeabError):
        G.step(1)
    T = threaded(2, 64, xyz, tri, ob, 0, solver=feab.SOLVER_BICGSTAB, loop_mode=feab.LOOP_GRAPH_WHILE)
    T.set_equation(2)
    with pytest.raises(feab.FeabError):
        T.step(1)


def test_one_process_per_gpu_partitioned_solve():
    """Real ranks: torchrun with one process per visible GPU (2..8) runs tests/mp_partitioned.py -- cudaIpc peer
    mappings, halo pushes and mailbox all-reduces over NVLink, WHILE-graph and host-batched loops, parity against the
    fp64 oracle.  Skipped on a single-GPU box (the virtual-rank tests above cover the data path there)."""
    import os
    import socket
    import subprocess
    import sys
    import torch
    ngpu = torch.cuda.device_count()
    if ngpu < 2:
        pytest.skip("needs >= 2 GPUs")
    world = min(ngpu, 8)
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(root, "tests", "mp_partitioned.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=root)
    assert out.returncode == 0 and "MP_PARTITIONED_OK" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]


# ---- the PRODUCTION multi-GPU kernels on one GPU: one host thread per virtual rank, ordinary feab_step ----------------
def threaded(world, real, xyz, tri, ob, bc, part=None, **solver):
    from fea_visualizer_b200.multigpu import ThreadedRanks
    sm = feab.Feab(real=64).device_info()["sm_count"]
    G = ThreadedRanks(world, real=real, part=part, sm_count=sm, **solver)
    for h in G.hs:
        h.set_boundary_condition(bc)
    G.set_mesh(xyz, tri, ob)
    return G


@pytest.mark.parametrize("name,bc", [("ico12", 0), ("disk14", 0), ("disk14", 1)])
@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("eq", [0, 1, 2, 3])
def test_threaded_ranks_persistent_kernels_step_parity(name, bc, world, eq):
    """Same bars as the lock-step test above, but through feab_step on every rank concurrently: pcg_dual / pcg_persistent
    with world > 1 (grid syncs merged with the mailbox all-reduce, halo stores fused into the direction phase and the RHS
    kernel, state pushes fused into the epilogue) -- the kernels bench.py times at N > 1."""
    xyz, tri, ob = mesh(name)
    V = len(xyz)
    F = single(64, xyz, tri, ob, bc)
    G = threaded(world, 64, xyz, tri, ob, bc)
    D = O.Oracle(xyz, tri, ob, bc, real=64)
    ref = []
    run_steps(F, V, eq, 5, per_step=lambda s, X: ref.append((X.get_values(), X.get_state(0), X.get_state(1), X.stats()["iterations"])))
    orc = []
    run_steps(D, V, eq, 5, per_step=lambda s, X: orc.append((X.get_values(), X.get_state(0), X.get_state(1))))
    worst = [0.0, 0.0]

    def check(s, X):
        v, u, vv, it = ref[s]
        worst[0] = max(worst[0], rel_l2(X.get_values(), v), rel_l2(X.get_state(0), u), rel_l2(X.get_state(1), vv))
        worst[1] = max(worst[1], rel_l2(X.get_values(), orc[s][0]), rel_l2(X.get_state(0), orc[s][1]), rel_l2(X.get_state(1), orc[s][2]))
        git = X.stats()["iterations"]
        assert abs(git[0] - it[0]) <= 2 and abs(git[1] - it[1]) <= 2, (git, it)
        if eq == 3:
            assert X.stats()["shared_iterations"] > 0     # the lock-step kernel ran on the partition

    run_steps(G, V, eq, 5, per_step=check)
    assert worst[0] <= 1e-11, worst      # vs single GPU: summation order only
    assert worst[1] <= 1e-8, worst       # vs the fp64 oracle: the north star's per-step bar
    assert G.unstable() == F.unstable()


@pytest.mark.parametrize("world", [2, 8])
def test_threaded_ranks_many_tiles_per_cta_and_device_resident_steps(world):
    """~92 k rows: with the grid capped at 296 / world CTAs every CTA walks several tiles and the halo wait sits in the middle
    of a tile loop; 6 steps in ONE call (no gather between steps: the RHS kernel re-reads the solver's own vector through
    float) against the fp64 oracle stepping one by one."""
    xyz, tri = meshes.icosphere(96)
    ob = meshes.boundary_flags(len(xyz), tri)
    V = len(xyz)
    G = threaded(world, 64, xyz, tri, ob, 0)
    D = O.Oracle(xyz, tri, ob, 0, real=64)
    v0 = np.zeros(V, np.float32)
    v0[meshes.seed_nodes(V)] = 1.0
    for eq in (3, 1):
        for X in (G, D):
            X.clear()
            X.set_equation(eq)
            X.set_values(v0)
        G.step(6)
        for _ in range(6):
            D.step(1)
        assert rel_l2(G.get_values(), D.get_values()) <= 1e-6      # values are float32
        assert rel_l2(G.get_state(0), D.get_state(0)) <= 1e-8 and rel_l2(G.get_state(1), D.get_state(1)) <= 1e-8


@pytest.mark.parametrize("world", [2, 4])
def test_threaded_ranks_rows_above_the_tile_staging_capacity(world):
    """The un-staged tile path (rows with 13 entries in whole tiles, meshes.fan_refined) of the partitioned persistent kernels:
    Gray-Scott through pcg_dual and heat through pcg_persistent on 2 / 4 concurrent virtual ranks vs the fp64 oracle."""
    bx, bt = meshes.icosphere(28, radius=28 / 32.0)
    xyz, tri = meshes.fan_refined(bx, bt)
    V = len(xyz)
    ob = meshes.boundary_flags(V, tri)
    G = threaded(world, 64, xyz, tri, ob, 0)
    assert all(h.counts()[3] >= 12 for h in G.hs)
    D = O.Oracle(xyz, tri, ob, 0, real=64)
    for eq in (3, 0):
        orc = []
        run_steps(D, V, eq, 4, per_step=lambda s, X: orc.append((X.get_values(), X.get_state(0), X.get_state(1))))
        worst = [0.0]

        def check(s, X):
            worst[0] = max(worst[0], rel_l2(X.get_values(), orc[s][0]), rel_l2(X.get_state(0), orc[s][1]), rel_l2(X.get_state(1), orc[s][2]))

        run_steps(G, V, eq, 4, per_step=check)
        assert worst[0] <= 1e-8, (eq, worst)


@pytest.mark.parametrize("name,bc", [("disk14", 0), ("disk14", 1), ("ico12", 0)])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_threaded_ranks_bicgstab_partitioned(name, bc, world):
    """BASELINE config 4 ("non-symmetric solve, 1 and 2 B200"): the persistent BiCGSTAB kernel on a row partition -- y and z
    halos pushed by the phases that produce them, five merged syncs per trip -- against the single-GPU run (summation order
    only), the fp64 oracle in BiCGSTAB mode (1e-8 per step) and Eigen's own BiCGSTAB when it travelled with the snapshot."""
    xyz, tri, ob = mesh(name)
    V = len(xyz)
    F = single(64, xyz, tri, ob, bc, solver=feab.SOLVER_BICGSTAB)
    G = threaded(world, 64, xyz, tri, ob, bc, solver=feab.SOLVER_BICGSTAB)
    D = O.Oracle(xyz, tri, ob, bc, real=64)
    D.set_solver(mode=1)
    ref, orc = [], []
    run_steps(F, V, 2, 4, per_step=lambda s, X: ref.append((X.get_values(), X.get_state(0), X.stats()["iterations"][0])))
    run_steps(D, V, 2, 4, per_step=lambda s, X: orc.append((X.get_values(), X.get_state(0))))
    worst = [0.0, 0.0]

    def check(s, X):
        worst[0] = max(worst[0], rel_l2(X.get_state(0), ref[s][1]))
        worst[1] = max(worst[1], rel_l2(X.get_state(0), orc[s][1]), rel_l2(X.get_values(), orc[s][0]))
        assert abs(X.stats()["iterations"][0] - ref[s][2]) <= 2

    run_steps(G, V, 2, 4, per_step=check)
    assert worst[0] <= 1e-10, worst      # vs one GPU: summation order of the dot products (BiCGSTAB amplifies it more than CG)
    assert worst[1] <= 1e-8, worst       # vs the fp64 oracle


@pytest.mark.parametrize("name", ["ico12", "disk14", "grid", "ico40r", "fan"])
def test_device_rcb_equals_host_rcb(name):
    """The default partition is computed on the device (csrc/partition.cuh: bounding box, radix select of the split coordinate,
    ties by vertex index) and must be the array feab_partition_rcb (host, std::nth_element) returns, element for element, for
    every world size -- grids and disks have thousands of equal coordinates (the tie path), the icospheres have +-0."""
    if name == "ico40r":
        xyz, tri = meshes.icosphere(40, radius=40 / 32.0)
    elif name == "fan":
        xyz, tri = meshes.fan_refined(*meshes.grid(40, 30, 1.25, 0.9375))
    else:
        xyz, tri, _ = mesh(name)
    ob = meshes.boundary_flags(len(xyz), tri)
    for world in (2, 3, 4, 5, 7, 8):
        h = feab.Feab(real=64)
        h.mg_config(world - 1, world)
        h.set_mesh(xyz, tri, ob)
        got = h.mg_partition()
        want = feab.partition_rcb(xyz, world)
        assert (got == want).all(), (name, world, int((got != want).sum()))
        h.close()


def test_set_mesh_rejects_out_of_range_triangle_indices():
    xyz, tri, ob = mesh("grid")
    bad = tri.copy()
    bad[3, 1] = len(xyz)
    h = feab.Feab(real=64)
    with pytest.raises(feab.FeabError):
        h.set_mesh(xyz, bad, ob)
    h.set_mesh(xyz, tri, ob)      # the handle is still usable
    h.step(1)
