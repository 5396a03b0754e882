"""Small-mesh runs of the kernels smoke() does not reach, for tools/sanitize.sh (compute-sanitizer): bicg_persistent on one
GPU, and -- argument `ranks` -- the partitioned kernels as threaded virtual ranks (world 2 and 4) on one device.  Every case
is checked against the fp64 oracle so that a sanitizer-clean but wrong run cannot pass."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")
from fea_visualizer_b200 import feab, meshes  # noqa: E402
from fea_visualizer_b200.multigpu import ThreadedRanks  # noqa: E402
from oracle import oracle as O  # noqa: E402


def check(F, D, V, eq, tag, steps=2):
    v0 = np.zeros(V, np.float32)
    v0[meshes.seed_nodes(V)] = 1.0
    for X in (F, D):
        X.clear()
        X.set_equation(eq)
        X.set_values(v0)
        X.step(steps)
    a, b = np.asarray(F.get_values(), np.float64), D.get_values().astype(np.float64)
    err = np.linalg.norm(a - b) / np.linalg.norm(b)
    assert err <= 1e-6, (tag, err)
    print(f"{tag}: rel l2 vs oracle {err:.2e}", flush=True)


def main():
    what = sys.argv[1] if len(sys.argv) > 1 else "single"
    xyz, tri = meshes.icosphere(8)
    ob = meshes.boundary_flags(len(xyz), tri)
    V = len(xyz)
    os.environ["FEAB_CLUSTER"] = "0"
    if what == "single":
        F = feab.Feab(real=64)
        F.set_solver(solver=feab.SOLVER_BICGSTAB)
        F.set_mesh(xyz, tri, ob)
        D = O.Oracle(xyz, tri, ob, 0, real=64)
        D.set_solver(mode=1)
        check(F, D, V, 2, "bicg_persistent (advection-diffusion, one GPU)")
        F.close()
        return
    for world in (2, 4):
        for eq, solver, tag in ((0, feab.SOLVER_PARITY_CG, "pcg_persistent"), (3, feab.SOLVER_PARITY_CG, "pcg_dual"),
                                (2, feab.SOLVER_BICGSTAB, "bicg_persistent")):
            F = ThreadedRanks(world, real=64, solver=solver)
            F.set_mesh(xyz, tri, ob)
            D = O.Oracle(xyz, tri, ob, 0, real=64)
            D.set_solver(mode=1 if solver == feab.SOLVER_BICGSTAB else 0)
            check(F, D, V, eq, f"{tag} world {world} (threaded virtual ranks)")
            F.close()


if __name__ == "__main__":
    main()
