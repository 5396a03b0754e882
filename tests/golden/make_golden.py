"""Generates the committed golden fixtures by running the UNMODIFIED reference (oracle/_ref/libfea_ref.so =
/root/reference/src/FEM/{FEMContext,CPUSolver}.cpp + Eigen, fp32) in the build container.

    python tests/golden/make_golden.py        (needs /root/reference and `make -C oracle`)

Outputs (tests/golden/):
  ref_small.npz   synthetic meshes (icosphere nu=4, 12x9 grid, 6-ring disk): mesh arrays, idx_map, the three CSC
                  matrices for both boundary conditions, and surface->values after 5 steps of each equation
  ref_assets.npz  the reference's bundled meshes assets/fem_meshes/{icosphere,tube}.obj as arrays (so the GPU box,
                  which has no /root/reference, can run the C1 config), their matrix sizes, and values after
                  100 heat steps (icosphere: the BASELINE.json C1 config) / 20 steps of each equation (tube)
The brush protocol is SURVEY.md 8(d): 50 seed nodes (k * 2654435761) mod V set to 1.0 before step 0; Gray-Scott
re-applies the seed before each of the first 5 steps.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from fea_visualizer_b200 import meshes  # noqa: E402
from oracle import oracle as O  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def run_steps(X, V, eq, nsteps):
    seeds = meshes.seed_nodes(V)
    X.clear()
    X.set_equation(eq)
    v = np.zeros(V, np.float32)
    v[seeds] = 1.0
    X.set_values(v)
    for s in range(nsteps):
        if eq == 3 and 0 < s < 5:
            v = X.get_values()
            v[seeds] = 1.0
            X.set_values(v)
        X.step(1)
    return X.get_values()


def small():
    out = {}
    cases = {"ico4": meshes.icosphere(4), "grid": meshes.grid(12, 9), "disk": meshes.disk(6)}
    for name, (xyz, tri) in cases.items():
        ob = meshes.boundary_flags(len(xyz), tri)
        out[f"{name}/xyz"], out[f"{name}/tri"], out[f"{name}/onb"] = xyz, tri, ob
        for bc in (0, 1):
            R = O.Reference(xyz, tri, ob, bc)
            out[f"{name}/bc{bc}/idx_map"] = R.idx_map()
            out[f"{name}/bc{bc}/max_row"] = np.int32(R.max_row_nonzeros())
            for w, mn in enumerate("MKA"):
                outer, inner, val = R.csc(w)
                out[f"{name}/bc{bc}/{mn}/outer"], out[f"{name}/bc{bc}/{mn}/inner"], out[f"{name}/bc{bc}/{mn}/val"] = outer, inner, val
            for eq in range(4):
                out[f"{name}/bc{bc}/eq{eq}/values5"] = run_steps(R, len(xyz), eq, 5)
                out[f"{name}/bc{bc}/eq{eq}/u"] = R.get_state(0)
                out[f"{name}/bc{bc}/eq{eq}/v"] = R.get_state(1)
            R.close()
    np.savez_compressed(os.path.join(OUT, "ref_small.npz"), **out)
    print("ref_small.npz", len(out), "arrays")


def assets():
    out = {}
    ref = "/root/reference/assets/fem_meshes"
    for name, steps in (("icosphere", 100), ("tube", 20)):
        # tube.obj is all quads; meshes.load_obj splits them as a fan (0,1,2),(0,2,3).  tinyobjloader (empty submodule in
        # the snapshot) is NOT available, so the tube/* arrays pin the reference's FEM code on this repo's split rule only.
        xyz, tri = meshes.load_obj(os.path.join(ref, name + ".obj"))
        ob = meshes.boundary_flags(len(xyz), tri)
        out[f"{name}/xyz"] = xyz
        out[f"{name}/tri"] = tri.astype(np.uint16)
        out[f"{name}/onb"] = ob
        for bc in (0, 1):
            R = O.Reference(xyz, tri, ob, bc)
            sizes = [R.num_unknowns(), R.max_row_nonzeros()] + [len(R.csc(w)[1]) for w in range(3)]
            out[f"{name}/bc{bc}/sizes"] = np.array(sizes, dtype=np.int64)  # n, max_row, nnz M, K, Adv
            sums = [float(np.sum(R.csc(w)[2].astype(np.float64))) for w in range(3)]
            out[f"{name}/bc{bc}/sums"] = np.array(sums)
            if name == "icosphere" and bc == 1:
                R.close()
                continue
            eqs = (0,) if name == "icosphere" else (0, 1, 2, 3)
            for eq in eqs:
                out[f"{name}/bc{bc}/eq{eq}/values{steps}"] = run_steps(R, len(xyz), eq, steps)
            R.close()
    np.savez_compressed(os.path.join(OUT, "ref_assets.npz"), **out)
    print("ref_assets.npz", len(out), "arrays")


if __name__ == "__main__":
    assert O.reference_available(), "build oracle/_ref first (make -C oracle)"
    small()
    assets()
