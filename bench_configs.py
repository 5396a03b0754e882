#!/usr/bin/env python
"""bench_configs.py -- BASELINE.json configs C1-C4 on one B200 next to the reference's Eigen CPU path (C5 is bench.py).

For every config: the same mesh, parameters and seed protocol through (i) libfeab.so, fp64, Eigen's stopping
rule, device-resident steps timed with CUDA events, and (ii) the UNMODIFIED reference (oracle/_ref/libfea_ref_fast.so,
fp32, OpenMP on all host cores) for a bounded number of steps; plus the relative L2 distance between the two fields after
the reference's last step (fp64 device arithmetic vs fp32 Eigen: expected ~1e-6).  One JSON line per config.

    python bench_configs.py [--only C1,C3] [--no-cpu]
    torchrun --nproc-per-node 2 ... bench_configs.py --only C4,C4b,C4p,C4pb     config 4 on a row partition ("1 and 2 B200")

Advection-diffusion lines carry `us_per_iteration` and, for BiCGSTAB, the byte rate of one trip (2 SpMV + 25 n reals:
376 B/row in fp64) against the measured HBM peak; `--oracle-steps K` adds the relative L2 distance to the fp64 oracle (in the
same solver mode) after K steps.
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

EQ_NAME = {0: "heat", 1: "wave", 2: "advection-diffusion", 3: "gray-scott"}


def rel_l2(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    nb = np.linalg.norm(b)
    return float(np.linalg.norm(a - b) / nb) if nb > 0 else float(np.linalg.norm(a))


def drive(X, V, eq, nsteps, seeds):
    """Application::run's order: brush, then advance_time (the Gray-Scott brush is held for 5 steps)."""
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


def pslg_disk(radius=57.3, sides=4096, area=0.004):
    """a regular 4096-gon through the reference's own PSLG hand-off (Surface::init_from_PSLG ->
    Triangle "Qeqpza<area>", src/Utils/Surface.cpp:25-51, :431-486) when the compiled reference travelled with the snapshot
    (oracle/_ref/libfea_surface_ref.so; mesh generation is input preparation, not the measured path).  area = 0.004 prints in
    fixed notation, so the reference's std::format string is parsed by Triangle as intended.  ~4.0 M triangles."""
    from oracle import surface_oracle as S
    if not S.available():
        return None
    th = np.linspace(0, 2 * np.pi, sides, endpoint=False)
    R = S.RefSurface()
    R.init_from_pslg(np.stack([radius * np.cos(th), radius * np.sin(th)], 1), area)
    xyz, _, tri, ob = R.get_mesh()
    return xyz, tri, ob


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--oracle-steps", type=int, default=0, help="advection-diffusion configs: compare with the fp64 oracle after this many steps")
    args = ap.parse_args()
    from fea_visualizer_b200 import feab, meshes
    from oracle import oracle as O
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        peak = 6650.0

    g = np.load(os.path.join(ROOT, "tests", "golden", "ref_assets.npz"))
    configs = []
    # C1: heat on the reference's bundled assets/fem_meshes/icosphere.obj (arrays travel in the golden fixture), 100 steps
    configs.append(("C1", "bundled icosphere.obj", lambda: (g["icosphere/xyz"], g["icosphere/tri"]), 0, 100, 100, {}))
    # C2: Gray-Scott, 1000 steps, bunny surrogate (the Stanford bunny is not in the reference repo)
    configs.append(("C2", "bunny surrogate: displaced icosphere nu=60, radius 60/32", lambda: meshes.icosphere(60, displace=True, radius=60 / 32.0),
                    3, 1000, 1000, {}))
    # C3: wave on an icosphere with ~1 M vertices
    configs.append(("C3", "icosphere nu=316, radius 316/32", lambda: meshes.icosphere(316, radius=316 / 32.0), 1, 100, 20, {}))
    # C4: advection-diffusion on a ~4 M-triangle disk, Dirichlet; reference-parity mode (PCG on A^T) and BiCGSTAB
    configs.append(("C4", "disk, 816 rings (~4.0 M triangles)", lambda: meshes.disk(816), 2, 20, 2, {}))
    configs.append(("C4b", "disk, 816 rings (~4.0 M triangles), BiCGSTAB mode", lambda: meshes.disk(816), 2, 20, 0, {"solver": 1}))
    # C4p: the same equation on the mesh SURVEY.md 8(d) specifies -- Triangle's unstructured output, boundary flags from its point markers
    configs.append(("C4p", "PSLG disk: 4096-gon R=57.3 through the reference's Triangle hand-off, area 0.004", pslg_disk, 2, 20, 2, {}))
    configs.append(("C4pb", "PSLG disk: 4096-gon R=57.3 through the reference's Triangle hand-off, area 0.004, BiCGSTAB mode", pslg_disk, 2, 20, 0,
                    {"solver": 1}))
    only = set(x for x in args.only.split(",") if x)
    for name, desc, mesh_fn, eq, gpu_steps, cpu_steps, solver in configs:
        if only and name not in only:
            continue
        if world > 1 and eq != 2:
            continue                                 # the partitioned Gray-Scott line is bench.py; config 4 is the one quoted on "1 and 2 B200"
        mesh = mesh_fn()
        if mesh is None:
            print(json.dumps({"config": name, "skipped": "oracle/_ref/libfea_surface_ref.so not present (mesh generator)"}), flush=True)
            continue
        xyz = np.ascontiguousarray(mesh[0], np.float32)
        tri = np.ascontiguousarray(mesh[1], np.uint32)
        V = len(xyz)
        ob = np.ascontiguousarray(mesh[2], np.uint8) if len(mesh) > 2 else meshes.boundary_flags(V, tri)
        seeds = meshes.seed_nodes(V)
        F = feab.Feab(real=64, device=local)
        if world > 1:
            F.mg_config(rank, world)
        if solver:
            F.set_solver(**solver)
        t0 = time.perf_counter()
        F.set_mesh(xyz, tri, ob)
        if world > 1:
            from fea_visualizer_b200.multigpu import connect_ranks
            connect_ranks(F, dist)
        setup_s = time.perf_counter() - t0
        _, n, nnz, _ = F.counts()
        if world > 1:
            import torch
            tot = torch.tensor([n, nnz], dtype=torch.float64, device="cuda")
            dist.all_reduce(tot)
            n, nnz = int(tot[0].item()), int(tot[1].item())
        # parity leg first: the same number of steps as the CPU run, then the timed device-resident run
        line = {"config": name, "equation": EQ_NAME[eq], "mesh": desc, "vertices": V, "triangles": len(tri), "unknowns": n, "nnz": nnz,
                "dtype": "f64", "solver": "bicgstab" if solver else "jacobi-pcg (Eigen recurrence, tol = eps)", "setup_s": setup_s,
                "n_gpus": world}
        ref_vals = None
        if cpu_steps and not args.no_cpu and world == 1 and (O.reference_available(fast=True) or O.reference_available()):
            fast = O.reference_available(fast=True)
            R = O.Reference(xyz, tri, ob, 0, fast=fast)
            cores = R.set_threads(os.cpu_count() or 1) if fast else 1
            t0 = time.perf_counter()
            drive(R, V, eq, cpu_steps, seeds)
            cpu_s = time.perf_counter() - t0
            ref_vals = R.get_values()
            line["cpu_baseline"] = {"kind": "reference", "cores": cores, "steps": cpu_steps, "steps_per_s": cpu_steps / cpu_s,
                                    "build": "-O3 -march=x86-64-v3 -fopenmp" if fast else "-O2"}
            drive(F, V, eq, cpu_steps, seeds)
            line["rel_l2_vs_reference_fp32_after_cpu_steps"] = rel_l2(F.get_values(), ref_vals)
            R.close()
        if args.oracle_steps > 0 and eq == 2:
            # parity in the same run: the fp64 oracle in the same solver mode (BiCGSTAB: Eigen's own recurrence, see tests)
            D = O.Oracle(xyz, tri, ob, 0, real=64)
            D.set_solver(mode=1 if solver else 0)
            t0 = time.perf_counter()
            drive(D, V, eq, args.oracle_steps, seeds)
            line["oracle_s_per_step"] = (time.perf_counter() - t0) / args.oracle_steps
            drive(F, V, eq, args.oracle_steps, seeds)
            idx = F.idx_map()
            row_node, _ = F.mg_maps()
            mine = F.get_state(0).astype(np.float64)
            ref = D.get_state(0)
            num, den = float(np.sum((mine - ref[idx[row_node]]) ** 2)), float(np.sum(ref ** 2))
            if world > 1:
                import torch
                tnum = torch.tensor([num], dtype=torch.float64, device="cuda")
                dist.all_reduce(tnum)
                num = float(tnum.item())
            line["rel_l2_vs_oracle_fp64"] = {"steps": args.oracle_steps, "value": float(np.sqrt(num / den)) if den > 0 else float(np.sqrt(num)),
                                            "oracle_iterations": [int(x) for x in D.last_stats()[0]], "device_iterations": F.stats()["iterations"]}
            D.close()
        # timed: device-resident
        F.clear()
        F.set_equation(eq)
        v = np.zeros(V, np.float32)
        v[seeds] = 1.0
        F.set_values(v)
        warm = 5 if eq == 3 else 3
        for s in range(warm):                       # warm-up (+ the held brush of Gray-Scott)
            if eq == 3 and s > 0:
                v = F.get_values()
                v[seeds] = 1.0
                F.set_values(v)
            F.step(1)
        F.step(gpu_steps - warm if gpu_steps > warm else gpu_steps)
        st = F.stats()
        steps_timed = st["last_call_steps"]
        ms_call = st["last_call_ms"]
        if world > 1:
            import torch
            tms = torch.tensor([ms_call], dtype=torch.float64, device="cuda")
            dist.all_reduce(tms, op=dist.ReduceOp.MAX)
            ms_call = float(tms.item())
        line.update({"gpu_steps_timed": steps_timed, "ms_per_step": ms_call / steps_timed,
                     "steps_per_s": steps_timed / ms_call * 1e3, "iterations_last_step": st["iterations"],
                     "rel_residual_last_step": st["rel_residual"], "unstable": bool(F.unstable())})
        if eq == 2 and st["solve_ms"] > 0 and st["iterations"][0] > 0:
            trips = st["spmv_passes"][0] / 2 if solver else st["spmv_passes"][0]
            line["us_per_iteration"] = st["solve_ms"] * 1e3 / max(trips, 1)
            # one BiCGSTAB trip: 2 SpMV (nnz (R + 4) + 4 n each) + 25 n R of vector traffic; one PCG pass: 192 B/row at nnz = 7 n
            per_trip = (2 * (nnz * 12 + 4 * n) + 25 * n * 8) if solver else (nnz * 12 + 4 * n + 13 * n * 8)
            gbps = per_trip * max(trips, 1) / st["solve_ms"] / 1e6 / world
            line["solve_byte_rate"] = {"algorithmic_bytes_per_iteration": per_trip, "GBps_per_gpu": gbps, "frac_of_measured_peak": gbps / peak,
                                       "peak_GBps": peak}
        prof = F.pcg_profile()
        if prof["iterations"]:   # FEAB_PCG_PROF=1: where CTA 0 of the persistent / cluster kernel spent its time
            line["phase_profile_us_per_iteration"] = {k: v / prof["iterations"] / 1e3 for k, v in prof["ns"].items() if v}
            line["phase_profile_iterations"] = prof["iterations"]
        if "cpu_baseline" in line:
            line["speedup_vs_reference_cpu"] = line["steps_per_s"] / line["cpu_baseline"]["steps_per_s"]
        if rank == 0:
            print(json.dumps(line), flush=True)
        F.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
