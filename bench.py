#!/usr/bin/env python
"""bench.py -- Gray-Scott timesteps/s on a synthetic icosphere (BASELINE.json metric) + the HBM roofline of the
dominant PCG kernel, next to the reference's Eigen CPU path.

    python bench.py --gpus 1 --steps K --warmup W            our arm (libfeab.so through the C ABI)
    python bench.py --impl reference --steps K --warmup W    the reference's CPU implementation (oracle/_ref)
    torchrun ... bench.py --gpus N ...                       N ranks, one mesh partition... see DESIGN.md (e)

A "step" is one Gray-Scott time step (two Jacobi-PCG solves run to Eigen's stopping rule) of the whole mesh.
`value` keeps the field device-resident (feab_step(NULL, K, 0)); `e2e` drives THE SAME K steps (state restored from a
snapshot taken after the warm-up) through the reference-facing call with a HOST values buffer (a pageable array page-locked once via feab_register_host_values), one host->device
and one device->host copy per step.  Before anything is timed the same handle type / loop driver / world size runs the
SURVEY 8(d) protocol on small meshes against the fp64 CPU oracle (`parity`, bar 1e-8 relative L2 per step); a parity
failure makes the run exit non-zero.  Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GS = dict(dt=0.001, Du=0.08, Dv=0.04, f=0.035, k=0.06)   # include/FEM/EquationParameters.hpp:43-49
EQ_GS = 3


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def ncu_traffic(rows, dtype):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the committed ncu --set full
    capture of the same workload (profiles/r01_spmv_traffic.json); None when the workload differs."""
    p = os.path.join(ROOT, "profiles", "r01_spmv_traffic.json")
    try:
        t = json.load(open(p))
        if t["rows"] == rows and t["dtype"] == dtype:
            return t["traffic_bytes_per_launch"]
    except Exception:
        pass
    return None


def ncu_solve_traffic(rows, dtype, passes, shared):
    """DRAM bytes of one step's persistent PCG kernels from the committed ncu captures (profiles/r02_solve_traffic.json:
    dram__bytes_read.sum + dram__bytes_write.sum of pcg_dual / pcg_persistent launches with the solve capped at several
    iteration counts -> bytes per lock-step pass, per single-system pass and per launch), scaled to THIS run's pass counts.
    None when the workload (rows, dtype) differs from the capture's."""
    p = os.path.join(ROOT, "profiles", "r02_solve_traffic.json")
    try:
        t = json.load(open(p))
        if t["rows"] != rows or t["dtype"] != dtype:
            return None, None
        single = max(passes) - shared if shared else passes[0] + passes[1]
        by = 0.0
        if shared:
            by += t["dual"]["fixed_bytes"] + t["dual"]["bytes_per_pass"] * shared
        if single:
            by += t["single"]["fixed_bytes"] * (1 if shared else 2) + t["single"]["bytes_per_pass"] * single
        return by, ("profiles/r02_solve_traffic.json (ncu dram bytes per lock-step pass / single pass / launch, "
                    f"scaled to this run: {shared} lock-step + {single} single passes)")
    except Exception:
        return None, None


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index=0):
        self.rows = []
        self.proc = None
        self.index = index

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for nm, val in zip(names, r[3:7]):
                    if val.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                continue
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def build_mesh(nu):
    """Icosphere with the element size of the reference's bundled icosphere.obj (radius = nu / 32): on the unit
    sphere the reference's own Gray-Scott step (reaction terms not mass-weighted, CPUSolver.cpp:134,137) reports
    has_numerical_instability() within 5 steps from ~1e5 vertices on."""
    from fea_visualizer_b200 import meshes
    xyz, tri = meshes.icosphere(nu, radius=nu / 32.0)
    ob = np.zeros(len(xyz), dtype=np.uint8)   # closed surface: no boundary nodes (checked by tests/test_meshes.py)
    return xyz, tri, ob


def seed_protocol(F, V, nsteps):
    """50 seed nodes = 1.0 before step 0, re-applied before each of the first 5 steps."""
    from fea_visualizer_b200 import meshes
    seeds = meshes.seed_nodes(V)
    for s in range(nsteps):
        if s < 5:
            F.brush(seeds, 1.0)
        F.step(1)


# ------------------------------------------------------------------------------------------------ parity, same run
def parity_meshes(world=1):
    """Meshes of the parity leg: (name, xyz, tri, on_boundary).  On a partitioned run (world > 1) the nu = 316 icosphere
    (998 562 rows, the mesh of BASELINE configs[2]) comes first, so that every rank certifies >= 125 k rows with interior
    and boundary tiles both live -- the three small meshes leave 12.5 k rows per rank at N = 8."""
    from fea_visualizer_b200 import meshes
    out = []
    if world > 1:
        xyz, tri = meshes.icosphere(316, radius=316 / 32.0)
        out.append(("icosphere nu=316 radius=9.875 (998562 vertices)", xyz, tri, np.zeros(len(xyz), np.uint8)))
    xyz, tri = meshes.icosphere(100, radius=100 / 32.0)      # 100 002 rows, the benchmark's element size
    out.append(("icosphere nu=100 radius=3.125 (100002 vertices)", xyz, tri, np.zeros(len(xyz), np.uint8)))
    g = os.path.join(ROOT, "tests", "golden", "ref_surface.npz")
    if os.path.exists(g):                                    # Triangle mesh from the reference's own PSLG hand-off, rows up to 10
        d = np.load(g)
        out.append(("reference PSLG 48-gon (1080 vertices, Dirichlet, max row 10)", d["pslg_xyz"], d["pslg_tri"], d["pslg_onb"]))
    bx, bt = meshes.icosphere(20, radius=20 / 32.0)
    xyz, tri = meshes.fan_refined(bx, bt)                    # 13 entries per row in the first tiles: the un-staged tile path
    out.append(("fan-refined icosphere nu=20 (12002 vertices, max row 13)", xyz, tri, np.zeros(len(xyz), np.uint8)))
    return out


def parity_leg(make_handle, torch, dist, rank, world, real, nsteps=5):
    """SURVEY.md 8(d) 'parity check performed in the same run': Gray-Scott, reference default parameters, seed protocol,
    `nsteps` device-resident steps through the handle type being benchmarked; after EVERY step the solver state u, v (own
    rows of this rank) and surface->values (own nodes) are compared with the fp64 CPU oracle (oracle/fea_oracle.c -- the
    checker, never the thing measured).  Relative L2 over the whole mesh (squared sums all-reduced over the ranks)."""
    from fea_visualizer_b200 import meshes
    from oracle import oracle as O
    bar = 1e-8 if real == 64 else 5e-6
    worst = 0.0
    per_mesh = {}
    for name, xyz, tri, ob in parity_meshes(world):
        V = len(xyz)
        seeds = meshes.seed_nodes(V)
        # the kernels the timed region runs: meshes this small would otherwise take the one-cluster kernel (pcg_cluster.cuh)
        os.environ["FEAB_CLUSTER"] = "0"
        try:
            F = make_handle(xyz, tri, ob)
        finally:
            del os.environ["FEAB_CLUSTER"]
        D = O.Oracle(xyz, tri, ob, 0, real=64)
        idx_map = F.idx_map()
        row_node, _ = F.mg_maps()
        grow = idx_map[row_node]
        for X in (F, D):
            X.clear()
            X.set_equation(EQ_GS)
            X.set_params(EQ_GS, GS["dt"], GS["Du"], GS["Dv"], GS["f"], GS["k"])
        mesh_worst = 0.0
        its = None
        for s in range(nsteps):
            F.brush(seeds, 1.0)
            dv = D.get_values()
            dv[seeds] = 1.0
            D.set_values(dv)
            F.step(1)
            D.step(1)
            ref = [D.get_state(0), D.get_state(1), D.get_values().astype(np.float64)]
            got = [F.get_state(0).astype(np.float64), F.get_state(1).astype(np.float64), F.get_values().astype(np.float64)[row_node]]
            num = [float(np.sum((got[0] - ref[0][grow]) ** 2)), float(np.sum((got[1] - ref[1][grow]) ** 2)),
                   float(np.sum((got[2] - ref[2][row_node]) ** 2))]
            bad = float(not all(np.isfinite(g).all() for g in got))
            tnum = torch.tensor(num + [bad], dtype=torch.float64, device="cuda")
            if world > 1:
                dist.all_reduce(tnum)
            num = tnum.tolist()
            den = [float(np.sum(r ** 2)) for r in ref]
            e = [float(np.sqrt(a / b)) if b > 0 else float(np.sqrt(a)) for a, b in zip(num[:3], den)]
            if num[3] > 0:
                e = [float("inf")] * 3
            mesh_worst = max(mesh_worst, *e)
            its = F.stats()["iterations"]
        oit, _ = D.last_stats()
        per_mesh[name] = {"rel_l2_max": mesh_worst, "rows": int((idx_map >= 0).sum()), "max_row": F.counts()[3],
                          "pcg_iterations_last_step": its, "oracle_iterations_last_step": [int(x) for x in oit],
                          "lock_step_passes": F.stats()["shared_iterations"]}
        worst = max(worst, mesh_worst)
        F.close()
        D.close()
    return {"rel_l2_max": worst, "bar": bar, "ok": bool(worst <= bar), "steps": nsteps, "equation": "gray-scott",
            "compared": "u, v (solver state) and surface->values after every step vs the fp64 CPU oracle, relative L2 over the whole mesh",
            "meshes": list(per_mesh.keys()), "per_mesh": per_mesh}


# ------------------------------------------------------------------------------------------------ CPU reference
def cpu_reference(nu, steps, warmup, threads, single_thread_steps=0, mesh=None):
    """Times the UNMODIFIED reference (FEMContext.cpp + CPUSolver.cpp, Eigen fp32 Jacobi-PCG; oracle/_ref/libfea_ref_fast.so:
    -O3 -march=x86-64-v3 -fopenmp) on the icosphere of frequency `nu` with the protocol of the GPU arm: assembly
    (FEMContext::init_from_surface, src/FEM/FEMContext.cpp:21-27, :57-64) timed separately, then `warmup` + `steps`
    Gray-Scott steps (CPUSolver::advance_time, src/FEM/CPUSolver.cpp:127-145) with all host threads (Eigen parallelises
    only the row-major SpMV), then `single_thread_steps` more steps on ONE thread -- what the reference ships
    (CMakeLists.txt:106 has no OpenMP)."""
    from fea_visualizer_b200 import meshes
    from oracle import oracle as O
    fast = O.reference_available(fast=True)
    if not (fast or O.reference_available()):
        return None
    xyz, tri, ob = mesh if mesh is not None else build_mesh(nu)
    V = len(xyz)
    t0 = time.perf_counter()
    R = O.Reference(xyz, tri, ob, 0, fast=fast)
    t_asm = time.perf_counter() - t0
    used = R.set_threads(threads) if fast else 1
    R.set_equation(EQ_GS)
    seeds = meshes.seed_nodes(V)
    times = []
    for s in range(warmup + steps):
        if s < 5:
            vals = R.get_values()
            vals[seeds] = 1.0
            R.set_values(vals)
        t0 = time.perf_counter()
        R.step(1)
        dt = time.perf_counter() - t0
        if s >= warmup:
            times.append(dt)
    sps = len(times) / sum(times)
    single = None
    if single_thread_steps > 0 and fast and used > 1:
        R.set_threads(1)
        t1 = []
        for _ in range(single_thread_steps):
            t0 = time.perf_counter()
            R.step(1)
            t1.append(time.perf_counter() - t0)
        single = len(t1) / sum(t1)
    unstable = bool(R.unstable())
    R.close()
    return {"steps_per_s": sps, "steps_per_s_1_thread": single, "rows": V, "assembly_s": t_asm, "cores": used, "kind": "reference",
            "timed_steps": len(times), "warmup_steps": warmup, "single_thread_steps": single_thread_steps if single else 0,
            "unstable": unstable, "build": "fast(-O3 -march=x86-64-v3 -fopenmp)" if fast else "pin(-O2)"}


def cpu_baseline_record(res, nu, V_target):
    """cpu_baseline object of the JSON line.  Measured on the metric's own mesh unless --cpu-nu asked for a smaller
    sample, in which case the figure is converted by the row ratio and labelled as such."""
    same = res["rows"] == V_target
    ratio = V_target / res["rows"]
    out = {"value": res["steps_per_s"] / ratio, "unit": "steps/s", "cores": res["cores"], "kind": res["kind"],
           "same_config": same, "timed_on": f"icosphere nu={nu} ({res['rows']} vertices)", "timed_steps": res["timed_steps"],
           "warmup_steps": res["warmup_steps"], "assembly_s": res["assembly_s"],
           "sample": (f"{res['timed_steps']} Gray-Scott steps after {res['warmup_steps']} warm-up on icosphere nu={nu} ({res['rows']} rows): "
                      f"{res['steps_per_s']:.4f} steps/s, fp32 Eigen PCG tol 1.19e-7, {res['cores']} threads, {res['build']}; "
                      f"assembly (init_from_surface) {res['assembly_s']:.1f} s not included"
                      + ("" if same else f"; CONVERTED to the {V_target}-row mesh by the row ratio {ratio:.3f}"))}
    if res.get("steps_per_s_1_thread"):
        out["single_thread_value"] = res["steps_per_s_1_thread"] / ratio      # the reference as shipped (no OpenMP)
        out["single_thread_steps"] = res["single_thread_steps"]
    return out


# ------------------------------------------------------------------------------------------------ main
_REAL_STDOUT = None


def claim_stdout():
    """stdout carries exactly ONE JSON line: native libraries (NCCL prints "NCCL version ..." with printf) get fd 2 instead."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--nu", type=int, default=env_int("FEAB_BENCH_NU", 1265), help="icosphere frequency; 1265 = 16 002 252 vertices")
    ap.add_argument("--real", type=int, default=64, choices=[32, 64])
    ap.add_argument("--cpu-nu", type=int, default=env_int("FEAB_BENCH_CPU_NU", 0), help="mesh of the CPU reference legs; 0 = the metric's own mesh (--nu)")
    ap.add_argument("--cpu-steps", type=int, default=3, help="timed steps of the CPU reference (SURVEY 8d: 3 at 16 M)")
    ap.add_argument("--cpu-warmup", type=int, default=2)
    ap.add_argument("--no-parity", action="store_true", help="skip the parity leg (profiling runs only)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--loop", default=os.environ.get("FEAB_BENCH_LOOP", "auto"), choices=["auto", "persistent", "graph", "host"],
                    help="solve-loop driver: auto = the library default; persistent whole-solve kernel, device-side WHILE graph, or host-batched")
    ap.add_argument("--tol", type=float, default=-1.0, help="<0: epsilon of the arithmetic type (Eigen default)")
    args = ap.parse_args()
    if args.cpu_nu <= 0:
        args.cpu_nu = args.nu

    rank = env_int("RANK", 0)
    world = env_int("WORLD_SIZE", 1)
    local_rank = env_int("LOCAL_RANK", 0)
    V_target = 10 * args.nu * args.nu + 2
    workload = (f"gray-scott icosphere nu={args.nu} radius={args.nu / 32.0:g} ({V_target} vertices, {20 * args.nu * args.nu} triangles), "
                "reference default parameters")
    ncores = os.cpu_count() or 1

    if args.impl == "reference":
        if rank != 0:
            return 0
        # the reference's own CPU path on the metric's mesh (a step is seconds at 16 M rows: the driver's K / W are clamped
        # to what fits a few minutes, and the line prints the counts really run)
        k = max(1, min(args.steps, args.cpu_steps))
        w = max(0, min(args.warmup, args.cpu_warmup))
        res = cpu_reference(args.cpu_nu, k, w, ncores, single_thread_steps=1)
        if res is None:
            emit({"impl": "reference", "unavailable": "oracle/_ref not built"})
            return 0
        cb = cpu_baseline_record(res, args.cpu_nu, V_target)
        val = cb["value"]
        line = {"metric": "gray_scott_timesteps_per_s", "value": val, "unit": "steps/s", "n_gpus": args.gpus, "steps": k,
                "warmup": w, "requested": {"steps": args.steps, "warmup": args.warmup}, "ms_per_step": 1e3 / val,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic", "impl": "reference",
                "config": {"workload": workload, "timed_on": cb["timed_on"], "same_config": cb["same_config"],
                           "assembly_s": res["assembly_s"], "unstable": res["unstable"]},
                "cpu_baseline": cb,
                "e2e": {"value": val, "unit": "steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        emit(line)
        return 0

    import torch
    import torch.distributed as dist
    from fea_visualizer_b200 import feab

    if not torch.cuda.is_available():
        emit({"error": "no CUDA device: libfeab has no CPU fallback"})
        return 1
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    torch.cuda.set_device(local_rank)
    loop_mode = None if args.loop == "auto" else {"host": feab.LOOP_HOST_BATCHED, "graph": feab.LOOP_GRAPH_WHILE, "persistent": feab.LOOP_PERSISTENT}[args.loop]

    def make_handle(xyz, tri, ob):
        """One handle of the configuration being benchmarked (arithmetic, loop driver, world size) on the given mesh."""
        F = feab.Feab(real=args.real, device=local_rank)
        if world > 1:
            F.mg_config(rank, world)
        if args.tol >= 0:
            F.set_solver(tolerance=args.tol)
        if loop_mode is not None:
            F.set_solver(loop_mode=loop_mode)
        F.set_mesh(xyz, tri, ob)
        if world > 1:
            from fea_visualizer_b200.multigpu import connect_ranks
            connect_ranks(F, dist)
        return F

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # --- parity against the fp64 CPU oracle, same run, same kernels (device-resident steps, production loop driver) ---
    parity = None
    if not args.no_parity:
        parity = parity_leg(make_handle, torch, dist, rank, world, args.real)
        if rank == 0:
            sys.stderr.write(f"[parity] {json.dumps(parity)}\n")

    # N > 1: ONE mesh, rows partitioned over the ranks (RCB), halos and dot products through peer memory inside the
    # solver kernels -> strong scaling of the metric's 16 M-vertex mesh
    xyz, tri, ob = build_mesh(args.nu)
    V = len(xyz)
    t0 = time.perf_counter()
    F = make_handle(xyz, tri, ob)
    setup_s = time.perf_counter() - t0
    try:
        # device memory in use on this rank's GPU once the handle exists (cudaMemGetInfo: CUDA context + mesh + matrices + vectors)
        free_b, total_b = torch.cuda.mem_get_info(local_rank)
        mem_used_gb = (total_b - free_b) / 1e9
    except Exception:
        mem_used_gb = -1.0
    _, n, nnz, max_row = F.counts()           # this rank's rows / entries
    lay = F.mg_layout()
    F.set_equation(EQ_GS)
    F.set_params(EQ_GS, GS["dt"], GS["Du"], GS["Dv"], GS["f"], GS["k"])
    F.clear()

    # warm-up (includes the seed protocol of the first 5 steps)
    warm = max(args.warmup, 3)
    seed_protocol(F, V, warm)
    # snapshot: both timed legs advance the SAME steps (warm .. warm + K - 1) from the same state
    snap_u, snap_v, snap_vals = F.get_state(0), F.get_state(1), F.get_values()
    # surface->values on the host: an ordinary (pageable) array, like the reference's std::vector<float>; the documented binding
    # page-locks it once per surface load through the C ABI (feab_register_host_values -> cudaHostRegister), further down
    host_np = np.zeros(V, dtype=np.float32)
    host = torch.from_numpy(host_np)
    if world > 1:
        # surface->values complete on rank 0's host: every owner contributes its nodes
        row_node, _ = F.mg_maps()
        full = torch.zeros(V, dtype=torch.float32, device="cuda")
        idx = torch.from_numpy(row_node.astype(np.int64)).cuda()
        full[idx] = torch.from_numpy(snap_vals).cuda()[idx]
        dist.all_reduce(full)
        host.copy_(full.cpu())
        del full, idx
    else:
        host.copy_(torch.from_numpy(snap_vals))
    launches0 = F.stats()["kernel_launches"]

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    F.pcg_profile(reset=True)
    # --- device-resident timed region ---
    barrier()
    F.step(args.steps)                       # K steps in one call; CUDA events on the solver stream bracket it inside
    barrier()
    st = F.stats()
    ms = st["last_call_ms"]
    its = st["iterations"]
    passes = st["spmv_passes"]
    launches = st["kernel_launches"] - launches0
    solve_ms = st.get("solve_ms", 0.0)       # persistent loop: device time of the last step's PCG kernel(s)
    solve_kernel = F.last_solve_kernel()     # "pcg_dual + pcg_persistent" (lock-step), "pcg_pair" (staggered), ...
    prof = F.pcg_profile(reset=True)          # FEAB_PCG_PROF=1: where CTA 0 of the persistent kernel spent its time
    step_ms = {"value_leg_last_step": prof["step_ms"]} if any(prof["step_ms"].values()) else None   # FEAB_STEP_PROF=1
    if prof["iterations"]:
        prof = {"us_per_iteration": {k: v / prof["iterations"] / 1e3 for k, v in prof["ns"].items()}, "iterations": prof["iterations"]}
        sys.stderr.write(f"[rank {rank}] pcg phase profile: {json.dumps(prof)}\n")
    else:
        prof = None
    # --- the same call from the UNREGISTERED pageable array, once, for the record (one GPU; 2 steps, not part of any timed leg) ---
    sync_mode = feab.SYNC_ROOT if world > 1 else feab.SYNC_FULL
    pageable_sps = None
    if world == 1:
        keep = host_np.copy()
        F.set_state(0, snap_u)
        F.set_state(1, snap_v)
        F.set_values(snap_vals)
        F.step_ptr(host_np.ctypes.data, 1, sync_mode)
        t0 = time.perf_counter()
        for _ in range(2):
            F.step_ptr(host_np.ctypes.data, 1, sync_mode)
        torch.cuda.synchronize()
        pageable_sps = 2.0 / (time.perf_counter() - t0)
        host_np[:] = keep
    F.register_host_values(host_np)
    # --- end-to-end through the reference-facing call with HOST buffers: the same K steps again ---
    F.set_state(0, snap_u)
    F.set_state(1, snap_v)
    F.set_values(snap_vals)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        # one GPU: H2D of values, one step, D2H of values, synchronised.  Partitioned: surface->values lives on rank 0's host
        # (FEAB_SYNC_ROOT): rank 0 uploads, NVLink forwards to the peers, only rank 0 downloads the result
        F.step_ptr(host_np.ctypes.data, 1, sync_mode)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    barrier()
    F.register_host_values(None)             # the array is about to go out of use: undo the page-lock first
    its_e2e = F.stats()["iterations"]
    if step_ms is not None:
        step_ms["e2e_leg_last_step"] = F.pcg_profile(reset=False)["step_ms"]
        sys.stderr.write(f"[rank {rank}] step phases (ms): {json.dumps(step_ms)}\n")
    clocks = sampler.stop() if rank == 0 else None

    t = torch.tensor([ms, e2e_s * 1e3, mem_used_gb], dtype=torch.float64, device="cuda")
    tot = torch.tensor([n, nnz, lay["n_ghost"], launches], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot)
    ms_max, e2e_ms_max, mem_used_gb_max = t.tolist()
    n_all, nnz_all, ghosts_all, launches_all = (int(x) for x in tot.tolist())
    rb = 8 if args.real == 64 else 4
    # algorithmic bytes of one Gray-Scott step: fused RHS + (it_u + it_v) PCG iterations
    it_bytes_row = nnz_all / n_all * (rb + 4) + 4 + 13 * rb
    # iterations the u- and the v-solve ran in lock-step (pcg_dual.cuh) read rowptr + columns once for both systems
    shared = min(st.get("shared_iterations", 0), its[0], its[1])
    solve_bytes = (its[0] + its[1]) * it_bytes_row * n_all - shared * (4 * nnz_all + 4 * n_all)
    step_bytes = nnz_all * (rb + 4) + 4 * n_all + 8 * n_all * rb + solve_bytes

    # --- roofline of the dominant kernel (SpMV + dot), measured live with CUDA events on the solver stream ---
    roof = None
    kernels = {}
    peak, peak_src = measured_peak()
    if rank == 0 and world == 1:
        for kid, name in ((0, "spmv_dot"), (1, "cg_update"), (2, "cg_direction"), (3, "rhs_init")):
            kms, kbytes = F.time_kernel(kid, 20)
            kernels[name] = {"ms": kms, "algorithmic_bytes": kbytes, "GBps": kbytes / kms / 1e6}
        k0 = kernels["spmv_dot"]
        it_ms = kernels["spmv_dot"]["ms"] + kernels["cg_update"]["ms"] + kernels["cg_direction"]["ms"]
        it_bytes = sum(kernels[k]["algorithmic_bytes"] for k in ("spmv_dot", "cg_update", "cg_direction"))
        dt_name = "f64" if args.real == 64 else "f32"
        spmv_alone = {"kernel": "spmv_dot (the standalone SpMV + dot kernel, timed alone; the persistent kernels run the same tile loop)",
                      "achieved": k0["GBps"], "frac": k0["GBps"] / peak, "frac_of_nominal_8TBps": k0["GBps"] / 8000.0,
                      "algorithmic_bytes": k0["algorithmic_bytes"], "traffic": ncu_traffic(n, dt_name),
                      "traffic_source": "profiles/r01_spmv_traffic.json (ncu --set full, per launch)"}
        if solve_ms:
            # the dominant kernels of the timed region: the persistent PCG solves (pcg_dual for the lock-step passes, pcg_persistent
            # for the longer solve's tail; pcg_persistent alone with FEAB_DUAL=0), events on the solver stream around the last step's
            traffic, traffic_src = ncu_solve_traffic(n, dt_name, passes, st.get("shared_iterations", 0))
            head = {"kernel": "persistent PCG solve kernels of one step (" + (solve_kernel or "pcg_persistent x2") + ")",
                    "achieved": solve_bytes / solve_ms / 1e6, "algorithmic_bytes": solve_bytes, "ms": solve_ms,
                    "traffic": traffic, "traffic_source": traffic_src}
        else:
            head = {"kernel": "spmv_dot", "achieved": k0["GBps"], "algorithmic_bytes": k0["algorithmic_bytes"], "ms": k0["ms"],
                    "traffic": spmv_alone["traffic"], "traffic_source": spmv_alone["traffic_source"]}
        roof = {"bound": "hbm", "kernel": head["kernel"], "achieved": head["achieved"], "peak": peak, "unit": "GB/s", "frac": head["achieved"] / peak,
                "traffic": head["traffic"], "algorithmic_bytes": head["algorithmic_bytes"], "ms_per_launch_group": head["ms"],
                "traffic_source": head["traffic_source"], "peak_source": peak_src, "frac_of_nominal_8TBps": head["achieved"] / 8000.0,
                "spmv_dot_alone": spmv_alone,
                "pcg_iteration": {"ms": it_ms, "algorithmic_bytes": it_bytes, "GBps": it_bytes / it_ms / 1e6,
                                  "frac": it_bytes / it_ms / 1e6 / peak},
                "pcg_persistent_last_step": None if not solve_ms else {
                    "ms": solve_ms, "iterations": its, "spmv_passes": passes, "shared_iterations": shared, "algorithmic_bytes": solve_bytes,
                    "GBps": solve_bytes / solve_ms / 1e6, "frac": solve_bytes / solve_ms / 1e6 / peak},
                "whole_step": {"algorithmic_bytes": step_bytes, "GBps": step_bytes / (ms_max / args.steps) / 1e6,
                               "frac": step_bytes / (ms_max / args.steps) / 1e6 / peak},
                "kernels": kernels}
    elif rank == 0:
        # partitioned run: the solver kernels wait for their peers, so no kernel is timed alone; the figure is the
        # whole step (RHS + all PCG iterations, halo waits and all-reduce latency included) per GPU
        gbps = step_bytes / (ms_max / args.steps) / 1e6 / world
        roof = {"bound": "hbm", "kernel": "whole Gray-Scott step (rhs_init + PCG iterations), per GPU", "achieved": gbps, "peak": peak,
                "unit": "GB/s", "frac": gbps / peak, "traffic": None, "algorithmic_bytes": step_bytes / world, "peak_source": peak_src,
                "frac_of_nominal_8TBps": gbps / 8000.0,
                "pcg_persistent_last_step": None if not solve_ms else {
                    "ms": solve_ms, "iterations": its, "per_gpu_GBps": (its[0] + its[1]) * it_bytes_row * n_all / solve_ms / 1e6 / world},
                "halo": {"ghost_columns_total": ghosts_all, "bytes_per_iteration_total": ghosts_all * rb,
                         "allreduces_per_iteration": 2, "collective": "P2P stores + mailbox flags over NVLink inside the solver kernels"}}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        # the unmodified reference on the host cores of this box, on the metric's own mesh (already built above)
        F.close()
        res = cpu_reference(args.cpu_nu, args.cpu_steps, min(args.cpu_warmup, 1), ncores, single_thread_steps=1,
                            mesh=(xyz, tri, ob) if args.cpu_nu == args.nu else None)
        if res is not None:
            cpu = cpu_baseline_record(res, args.cpu_nu, V)

    rc = 0
    if rank == 0:
        total_steps = args.steps              # one mesh: N ranks advance the same time step together
        value = total_steps / (ms_max / 1e3)
        e2e_value = total_steps / (e2e_ms_max / 1e3)
        line = {"metric": "gray_scott_timesteps_per_s", "value": value, "unit": "steps/s", "n_gpus": world, "steps": args.steps,
                "warmup": warm, "ms_per_step": ms_max / args.steps, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f64" if args.real == 64 else "f32", "data": "synthetic",
                "config": {"workload": workload, "rows": n_all, "nnz": nnz_all, "max_row": max_row,
                           "parallelism": "single GPU" if world == 1 else f"rows partitioned x{world} (RCB), P2P halo + mailbox all-reduce over NVLink",
                           "tolerance": "eps of dtype (Eigen default)" if args.tol < 0 else args.tol,
                           "pcg_iterations_last_step": its, "spmv_passes_last_step": passes, "loop": args.loop, "solve_kernel": solve_kernel,
                           "l2_policy": "inputs larger than L2 (matrix + vectors >> 126 MB)", "setup_s": setup_s,
                           "device_mem_used_gb_max_over_ranks": round(mem_used_gb_max, 3) if mem_used_gb_max >= 0 else None,
                           "timed_steps": f"steps {warm}..{warm + args.steps - 1} of the run, in both legs (state restored from a snapshot in between)"},
                "us_per_pcg_iteration": (solve_ms * 1e3 / max(1, its[0] + its[1])) if solve_ms else None,
                "e2e": {"value": e2e_value, "unit": "steps/s", "h2d_bytes_per_step": 4 * V, "d2h_bytes_per_step": 4 * V,
                        "mode": "host buffers on rank 0 (FEAB_SYNC_ROOT), peers fed over NVLink" if world > 1 else "host buffers, full sync",
                        "host_memory": "pageable numpy array, page-locked once through the C ABI (feab_register_host_values -> cudaHostRegister), as the documented binding does with surface->values",
                        "pageable_unregistered_value": pageable_sps,
                        "pcg_iterations_last_step": its_e2e, "same_steps_as_value": list(its_e2e) == list(its), "le_value": e2e_value <= value},
                "gpu_launches": launches_all, "pcg_phase_profile": prof, "step_phase_ms": step_ms, "parity": parity, "roofline": roof, "cpu_baseline": cpu, "clocks": clocks}
        emit(line)
        if parity is not None and not parity["ok"]:
            sys.stderr.write(f"PARITY FAILED: {json.dumps(parity)}\n")
            rc = 3
    if world > 1:
        rct = torch.tensor([rc], dtype=torch.int32, device="cuda")
        dist.broadcast(rct, 0)
        rc = int(rct.item())
        dist.destroy_process_group()
    return rc


if __name__ == "__main__":
    sys.exit(main())
