"""GPU tests of the error paths on ONE GPU: a grid sync of a persistent solve kernel that times out must
surface as FEAB_ECOMM, leave surface->values untouched and leave the handle usable; feab_unstable must report an
excursion in ANY step of a multi-step call (the reference tests after every advance_time, src/Application.cpp:532-533).

The fault is injected with FEAB_FAULT_GSYNC=k (feab.cu step()): before the k-th stepping call one arrival is taken away
from the grid-sync counter, so the first sync of that call waits its full bounded time (10 s, comm.cuh kSpinTimeoutNs).
"""
import numpy as np
import pytest

from fea_visualizer_b200 import feab, meshes
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def seeded(V, seed=3):
    v = np.zeros(V, np.float32)
    v[meshes.seed_nodes(V, 40)] = 1.0
    return v


@pytest.mark.parametrize("eq", [0, 3])      # heat: pcg_persistent; Gray-Scott: pcg_dual + pcg_persistent
def test_timed_out_grid_sync_is_reported_and_the_handle_recovers(eq, monkeypatch):
    monkeypatch.setenv("FEAB_CLUSTER", "0")          # the grid-wide kernels, not the one-cluster kernel
    monkeypatch.setenv("FEAB_FAULT_GSYNC", "2")      # the 2nd stepping call loses an arrival
    xyz, tri = meshes.icosphere(24)
    ob = meshes.boundary_flags(len(xyz), tri)
    V = len(xyz)
    F = feab.Feab(real=64)
    F.set_mesh(xyz, tri, ob)
    F.set_equation(eq)
    F.clear()
    F.set_values(seeded(V))
    F.step(2)
    before = F.get_values()
    with pytest.raises(feab.FeabError) as e:
        F.step(3)
    assert e.value.code == -5 and "grid sync" in str(e.value), str(e.value)      # FEAB_ECOMM, kind 6
    # fail closed: nothing of the abandoned call reached surface->values
    assert np.array_equal(F.get_values(), before)
    # ... and after feab_clear (what the message asks for) the handle works again and agrees with the oracle
    D = O.Oracle(xyz, tri, ob, 0, real=64)
    D.set_equation(eq)
    for X in (F, D):
        X.clear()
        X.set_values(seeded(V))
        X.step(5)
    for which in ((0, 1) if eq == 3 else (0,)):
        got, want = F.get_state(which), D.get_state(which)
        assert np.linalg.norm(got - want) <= 1e-8 * np.linalg.norm(want), which
    ns = 2 if eq == 3 else 1
    assert [int(i) for i in F.stats()["iterations"][:ns]] == [int(i) for i in D.last_stats()[0][:ns]]
    F.close()


def test_unstable_reports_an_excursion_in_any_step_of_the_call(monkeypatch):
    """Heat on a Dirichlet disk started at 2e4: the first steps are above 1e4, diffusion + the zero boundary bring the
    field below it later.  One call covering both phases must still say `unstable`; a later call that stays below must not."""
    for cluster in ("1", "0"):
        monkeypatch.setenv("FEAB_CLUSTER", cluster)
        xyz, tri = meshes.disk(10)
        V = len(xyz)
        ob = meshes.boundary_flags(V, tri)
        F = feab.Feab(real=64)
        F.set_mesh(xyz, tri, ob)
        F.set_equation(0)
        F.set_params(0, dt=0.05, c0=1.0)
        F.clear()
        F.set_values(np.full(V, 2e4, np.float32))
        n = 0
        while F.get_state(0).max() > 1e4 or n == 0:       # single steps until the field is back under the threshold
            F.step(1)
            n += 1
            assert n < 400
        assert n >= 2 and not F.unstable()                # the LAST step was fine ...
        F.clear()
        F.set_values(np.full(V, 2e4, np.float32))
        F.step(n)                                         # ... the same n steps in ONE call: the early ones were not
        assert F.get_state(0).max() <= 1e4 and F.unstable(), cluster
        F.step(2)
        assert not F.unstable(), cluster
        F.close()


@pytest.mark.parametrize("real", [64, 32])
def test_measurement_hook_runs_every_kernel(real, monkeypatch):
    """feab_time_kernel (bench.py's per-kernel timings) launches the SpMV, the two PCG vector kernels and the RHS build of
    every equation on their own: all of them read the handle's error flag now, so they need a valid communicator."""
    monkeypatch.setenv("FEAB_CLUSTER", "0")
    xyz, tri = meshes.icosphere(24)
    ob = meshes.boundary_flags(len(xyz), tri)
    F = feab.Feab(real=real)
    F.set_mesh(xyz, tri, ob)
    for eq in (0, 1, 2, 3):
        F.set_equation(eq)
        F.clear()
        F.set_values(seeded(len(xyz)))
        F.step(1)
        for kernel in range(4):
            ms, nbytes = F.time_kernel(kernel, 3)
            assert ms > 0 and nbytes > 0, (eq, kernel)
        F.step(1)
    assert np.isfinite(F.get_values()).all()
    F.close()


def test_registered_host_values_give_the_same_steps_as_pageable_ones():
    """feab_register_host_values page-locks the caller's surface->values array (cudaHostRegister); the numbers must not
    depend on it, re-registration and un-registration must be clean, and a handle can be destroyed while registered."""
    xyz, tri = meshes.icosphere(24)
    ob = meshes.boundary_flags(len(xyz), tri)
    V = len(xyz)
    outs = []
    for registered in (False, True, True):
        F = feab.Feab(real=64)
        F.set_mesh(xyz, tri, ob)
        F.set_equation(3)
        F.clear()
        host = seeded(V).copy()
        other = np.zeros(V, np.float32)
        if registered:
            F.register_host_values(other)
            F.register_host_values(host)          # replaces the first registration
        for _ in range(4):
            F.step(1, values=host, sync_values=True)
        outs.append(host.copy())
        if registered and len(outs) == 2:
            F.register_host_values(None)          # explicit un-registration; the third round leaves it to feab_destroy
            F.step(1, values=host, sync_values=True)
        with pytest.raises(feab.FeabError):
            F.lib.feab_register_host_values.argtypes = None
            F._ck(F.lib.feab_register_host_values(F.h, host.ctypes.data_as(__import__("ctypes").c_void_p), __import__("ctypes").c_int64(0)))
        F.close()
    assert np.array_equal(outs[0], outs[1]) and np.array_equal(outs[0], outs[2])
    assert np.isfinite(outs[0]).all() and outs[0].any()
