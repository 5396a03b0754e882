"""Generates tests/golden/ref_surface.npz by running the UNMODIFIED reference Surface.cpp / ColorMap.cpp
(oracle/_ref/libfea_surface_ref.so) in the build container.

    python tests/golden/make_golden_surface.py      (needs /root/reference and `make -C oracle`)

Contents
  <mesh>_xyz/_nrm/_tri/_val   inputs of every case (so the GPU box needs nothing else)
  <mesh>_onb                  on_boundary as Surface::init_from_obj computes it (the mesh is written to a temporary
                              OBJ and read back through the reference's own ingest)
  ply_<mesh>_<type>_<k>       bytes of the .ply Surface::export_to_ply writes, for mesh_type in {0,1,2} and the
                              (vertex_extrusion, threshold) pairs in CASES
  cm_<name>_t/_rgb            ColorMap::get_color samples for three of the reference's maps
  pick_<mesh>_*               200 seeded rays through the reference BVH (depth 10, include/Application.hpp:43): hit triangle,
                              point, distance (BVH::ray_triangle_intersection) and the node Application::brush would paint
  pslg_*                      a 48-gon through the reference's own PSLG hand-off (Surface::init_from_PSLG -> Triangle "Qeqpza0.004",
                              src/Utils/Surface.cpp:25-51, :431-486): vertices, triangles (winding reversed), point-marker
                              boundary flags, and surface->values after 20 advection-diffusion / heat steps of the unmodified
                              CPUSolver (oracle/_ref/libfea_ref.so) on that mesh
  obj_icosphere_*             the bundled assets/fem_meshes/icosphere.obj after init_from_obj: counts and a checksum
"""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from fea_visualizer_b200 import meshes  # noqa: E402
from oracle import surface_oracle as S  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
CASES = [(0.25, 0.0), (0.25, 0.1), (0.0, 0.3), (-0.5, 0.05), (1.5, -0.2)]   # (vertex_extrusion, threshold)


def surface_cases():
    out = {}
    xyz, tri = meshes.icosphere(6, displace=True)
    nrm = (xyz / np.linalg.norm(xyz, axis=1, keepdims=True)).astype(np.float32)
    val = (np.sin(3 * xyz[:, 0]) * np.cos(2 * xyz[:, 1]) + 0.2 * xyz[:, 2]).astype(np.float32)
    val[::7] = np.float32(0.1)          # exactly on a threshold: (v - thr) >= 1e-9 is false
    val[::11] = np.float32(0.0)
    out["ico6d"] = (xyz, nrm, tri, val)
    xyz, tri = meshes.grid(9, 7)
    nrm = np.tile(np.array([0, 1, 0], np.float32), (len(xyz), 1))
    val = (np.sin(5 * xyz[:, 0]) * np.sin(4 * xyz[:, 2])).astype(np.float32)
    out["grid"] = (xyz, nrm, tri, val)
    xyz, tri = meshes.disk(5)
    nrm = np.tile(np.array([0, 1, 0], np.float32), (len(xyz), 1))
    val = (1.0 - 2.0 * np.hypot(xyz[:, 0], xyz[:, 2])).astype(np.float32)
    out["disk"] = (xyz, nrm, tri, val)
    return out


def pick_rays(xyz, n=200, seed=7):
    """Rays from a shell around the mesh towards jittered points of it (most hit), plus a few that miss."""
    rng = np.random.default_rng(seed)
    c = xyz.mean(0)
    r = float(np.linalg.norm(xyz - c, axis=1).max())
    o = rng.normal(size=(n, 3))
    o = (c + 3.0 * r * o / np.linalg.norm(o, axis=1, keepdims=True)).astype(np.float32)
    tgt = xyz[rng.integers(0, len(xyz), n)] + rng.normal(scale=0.05 * r, size=(n, 3))
    tgt[::10] += 5.0 * r                                    # misses
    d = tgt - o
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    return o, d


def write_obj(path, xyz, nrm, tri):
    with open(path, "w") as f:
        for p in xyz:
            f.write("v %.9g %.9g %.9g\n" % tuple(p))
        for n in nrm:
            f.write("vn %.9g %.9g %.9g\n" % tuple(n))
        for t in tri:
            f.write("f %d//%d %d//%d %d//%d\n" % (t[0] + 1, t[0] + 1, t[1] + 1, t[1] + 1, t[2] + 1, t[2] + 1))


def main():
    assert S.available(), "build oracle/_ref first (make -C oracle)"
    data = {}
    tmp = tempfile.mkdtemp()
    for name, (xyz, nrm, tri, val) in surface_cases().items():
        data[f"{name}_xyz"], data[f"{name}_nrm"], data[f"{name}_tri"], data[f"{name}_val"] = xyz, nrm, tri.astype(np.uint32), val
        # boundary flags through the reference's own OBJ ingest
        R = S.RefSurface()
        obj = os.path.join(tmp, name + ".obj")
        write_obj(obj, xyz, nrm, tri)
        R.init_from_obj(obj)
        x2, n2, t2, ob = R.get_mesh()
        assert (x2 == xyz).all() and (t2 == tri).all()
        data[f"{name}_onb"] = ob
        data[f"{name}_objnrm"] = n2
        R2 = S.RefSurface()
        R2.set_mesh(xyz, nrm, tri)
        R2.set_values(val)
        for mt in (0, 1, 2):
            for k, (ve, thr) in enumerate(CASES):
                p = os.path.join(tmp, "o.ply")
                R2.export_to_ply(p, ve, thr, mt)
                data[f"ply_{name}_{mt}_{k}"] = np.frombuffer(open(p, "rb").read(), dtype=np.uint8)
    for name in ("ico6d", "disk"):
        xyz, nrm, tri, val = surface_cases()[name]
        R = S.RefSurface()
        R.set_mesh(xyz, nrm, tri)
        R.bvh_build(10)
        o, d = pick_rays(xyz)
        res = [R.pick(o[i], d[i]) for i in range(len(o))]
        data[f"pick_{name}_o"], data[f"pick_{name}_d"] = o, d
        data[f"pick_{name}_tri"] = np.array([r[0] for r in res], np.int32)
        data[f"pick_{name}_point"] = np.array([r[1] for r in res], np.float32)
        data[f"pick_{name}_dist"] = np.array([r[2] for r in res], np.float32)
        data[f"pick_{name}_node"] = np.array([R.brush(o[i], d[i], 1.0 + i) for i in range(len(o))], np.int32)
        data[f"pick_{name}_values"] = R.get_values()
    from oracle import oracle as O
    th = np.linspace(0, 2 * np.pi, 48, endpoint=False)
    R = S.RefSurface()
    R.init_from_pslg(np.stack([1.3 * np.cos(th), 1.3 * np.sin(th)], 1), 0.004)
    xyz, _, tri, ob = R.get_mesh()
    data["pslg_xyz"], data["pslg_tri"], data["pslg_onb"] = xyz, tri, ob
    seeds = meshes.seed_nodes(len(xyz))
    for eq in (0, 2):
        X = O.Reference(xyz, tri, ob, 0)
        X.clear()
        X.set_equation(eq)
        v = np.zeros(len(xyz), np.float32)
        v[seeds] = 1.0
        X.set_values(v)
        X.step(20)
        data[f"pslg_values_eq{eq}"] = X.get_values()
    t = np.concatenate([np.linspace(-0.25, 1.25, 301), [0.0, 1.0, 0.5, 1e-20, -0.0]]).astype(np.float32)
    for cm, coeffs in S.COLOR_MAPS.items():
        R = S.RefSurface()
        R.set_colormap(coeffs)
        data[f"cm_{cm}_t"], data[f"cm_{cm}_rgb"] = t, R.get_color(t)
    R = S.RefSurface()
    R.init_from_obj("/root/reference/assets/fem_meshes/icosphere.obj")
    V, T, nb, closed = R.counts()
    _, nrm, _, _ = R.get_mesh()
    data["obj_icosphere_counts"] = np.array([V, T, nb, int(closed)], np.int64)
    data["obj_icosphere_nrm"] = nrm
    np.savez_compressed(os.path.join(OUT, "ref_surface.npz"), **data)
    print("wrote ref_surface.npz", sum(v.nbytes for v in data.values()) // 1024, "KiB raw")


if __name__ == "__main__":
    main()
