#!/usr/bin/env python
"""bench_surface.py -- the rows either side of the hot path on the metric's 16 M-vertex
icosphere: .ply build (clip + de-dup + colour) and full export, boundary-edge detection, smooth normals; next to the
UNMODIFIED reference Surface.cpp (oracle/_ref/libfea_surface_ref.so) on a bounded sample.  One JSON line per row.

    python bench_surface.py [--nu 1265] [--cpu-nu 316] [--reps 5] [--no-cpu] [--export /tmp/out.ply]

Not the driver's bench (that is bench.py); same timing rules: warm-up, CUDA events on the launching stream
(feas_last_ms), inputs far larger than L2.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def field(xyz, radius):
    u = xyz / np.float32(radius)
    return (0.5 + 0.5 * np.sin(6 * u[:, 0]) * np.cos(5 * u[:, 1]) + 0.1 * u[:, 2]).astype(np.float32)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nu", type=int, default=1265)
    ap.add_argument("--cpu-nu", type=int, default=316)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--export", default="")
    args = ap.parse_args()
    from fea_visualizer_b200 import meshes
    from fea_visualizer_b200.surface import MESH_CLOSED, MESH_OPEN, DeviceSurface

    radius = args.nu / 32.0
    xyz, tri = meshes.icosphere(args.nu, radius=radius)
    V, T = len(xyz), len(tri)
    nrm = (xyz / np.float32(radius)).astype(np.float32)
    val = field(xyz, radius)
    D = DeviceSurface()
    t0 = time.perf_counter()
    D.set_mesh(xyz, nrm, tri)
    D.set_values(val)
    upload_s = time.perf_counter() - t0
    workload = f"icosphere nu={args.nu} ({V} vertices, {T} triangles)"
    rows = []

    def timed(fn):
        fn()
        ms = []
        for _ in range(args.reps):
            fn()
            ms.append(D.last_ms()[0])
        return float(np.median(ms))

    # N2 boundary detection: 3T edge inserts into a (key u64, count i32) table of >= 6T slots + one pass over the table
    ms = timed(lambda: D.detect_boundary())
    slots = 1 << int(np.ceil(np.log2(max(6 * T, 1024))))
    byts = 12 * T + 3 * T * 12 + slots * 12 + slots * 12 + V
    rows.append({"row": "N2 detect_boundary", "ms": ms, "edges_per_s": 3 * T / ms * 1e3, "algorithmic_bytes": byts, "GBps": byts / ms / 1e6,
                 "note": "bytes = tri read + 3T table updates (12 B) + table clear + table scan + flags"})
    # N3 smooth normals: read tri + gathered xyz/nrm/val once, 9 float atomics per triangle
    ms = timed(lambda: D.calculate_normals(0.25))
    byts = 12 * T + 28 * V + 12 * V + 9 * 4 * T
    rows.append({"row": "N3 smooth_normals", "ms": ms, "triangles_per_s": T / ms * 1e3, "algorithmic_bytes": byts, "GBps": byts / ms / 1e6,
                 "note": "events bracket the clear + the kernel; the D2H of the result is outside"})
    # N4 picking: every triangle tested against one ray (read tri + gathered xyz)
    org = np.array([0.3 * radius, 0.2 * radius, 3.0 * radius], np.float32)
    dr = np.array([-0.05, 0.02, -1.0], np.float32)
    dr /= np.linalg.norm(dr)
    hit = {}

    def pk():
        hit["t"] = D.pick(org, dr)
    ms = timed(pk)
    byts = 12 * T + 12 * V
    rows.append({"row": "N4 pick (brute force over all triangles)", "ms": ms, "triangles_per_s": T / ms * 1e3, "algorithmic_bytes": byts,
                 "GBps": byts / ms / 1e6, "hit_triangle": hit["t"][0]})
    # N1 ply build (device only)
    for mt, thr in ((MESH_OPEN, 0.5), (MESH_CLOSED, 0.5), (MESH_OPEN, -10.0)):
        res = {}

        def go():
            res["n"] = D.ply_build(0.25, thr, mt, fetch=False)
        ms = timed(go)
        nv, nt = res["n"]
        rows.append({"row": f"N1 ply_build mesh_type={mt} threshold={thr}", "ms": ms, "triangles_per_s": T / ms * 1e3, "out_vertices": nv,
                     "out_triangles": nt})
    if args.export:
        t0 = time.perf_counter()
        D.export_to_ply(args.export, 0.25, 0.5, MESH_OPEN)
        rows.append({"row": "N1 export_to_ply (build + D2H + ASCII write)", "s": time.perf_counter() - t0,
                     "file_bytes": os.path.getsize(args.export), "threads": os.cpu_count()})
    cpu = None
    if not args.no_cpu:
        from oracle import surface_oracle as S
        if S.available():
            r2 = args.cpu_nu / 32.0
            x2, t2 = meshes.icosphere(args.cpu_nu, radius=r2)
            n2 = (x2 / np.float32(r2)).astype(np.float32)
            R = S.RefSurface()
            R.set_mesh(x2, n2, t2)
            R.set_values(field(x2, r2))
            out = "/tmp/_ref_export.ply"
            t0 = time.perf_counter()
            R.export_to_ply(out, 0.25, 0.5, 0)
            s = time.perf_counter() - t0
            D2 = DeviceSurface()
            D2.set_mesh(x2, n2, t2)
            D2.set_values(field(x2, r2))
            out2 = "/tmp/_dev_export.ply"
            t0 = time.perf_counter()
            D2.export_to_ply(out2, 0.25, 0.5, 0)
            s2 = time.perf_counter() - t0
            same = open(out, "rb").read() == open(out2, "rb").read()
            cpu = {"kind": "reference", "cores": 1, "sample": f"Surface::export_to_ply on icosphere nu={args.cpu_nu} ({len(x2)} vertices), Open, threshold 0.5",
                   "reference_s": s, "ours_s_same_mesh": s2, "byte_identical": same, "file_bytes": os.path.getsize(out)}
            os.remove(out)
            os.remove(out2)
    for r in rows:
        r.update({"workload": workload, "upload_s": upload_s})
        print(json.dumps(r))
    if cpu:
        print(json.dumps({"row": "cpu_baseline", **cpu}))
    print(json.dumps({"row": "launches", "kernel_launches": D.last_ms()[1]}))


if __name__ == "__main__":
    main()
