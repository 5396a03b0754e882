"""DRAM traffic of the persistent PCG kernels per pass, for bench.py's roofline.traffic (profiles/r02_solve_traffic.json).

Run the driver under ncu on one GPU, then fit:

    ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
        --kernel-name regex:pcg_ --csv --log-file gpurun_out/traffic.csv python tools/ncu_traffic.py run gpurun_out/traffic_plan.json
    python tools/ncu_traffic.py fit gpurun_out/traffic.csv gpurun_out/traffic_plan.json profiles/r02_solve_traffic.json

`run` advances the metric's workload (Gray-Scott, icosphere nu = 1265, fp64) by single steps whose solves are capped at
4 / 8 / 12 / 16 iterations, first with the lock-step kernel (pcg_dual), then with separate solves (FEAB_DUAL=0,
pcg_persistent), and writes the sequence of solve launches it caused.  `fit` pairs that sequence with ncu's launch list
and fits bytes = fixed + per_pass * passes for either kernel (least squares)."""
import csv
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
CAPS = (4, 8, 12, 16)


def run(plan_path, nu=1265):
    from fea_visualizer_b200 import feab, meshes
    xyz, tri = meshes.icosphere(nu, radius=nu / 32.0)
    V = len(xyz)
    ob = np.zeros(V, np.uint8)
    seeds = meshes.seed_nodes(V)
    plan = {"rows": None, "dtype": "f64", "launches": []}
    for dual in (1, 0):
        os.environ["FEAB_DUAL"] = str(dual)
        F = feab.Feab(real=64)
        F.set_mesh(xyz, tri, ob)
        plan["rows"] = F.counts()[1]
        F.set_equation(3)
        F.clear()
        for cap in CAPS:
            F.set_solver(max_iter=cap)
            F.brush(seeds, 1.0)
            F.step(1)
            st = F.stats()
            p = st["spmv_passes"]
            if dual:
                plan["launches"].append({"kernel": "pcg_dual", "passes": st["shared_iterations"]})
                plan["launches"].append({"kernel": "pcg_persistent", "passes": max(p) - st["shared_iterations"]})
            else:
                plan["launches"].append({"kernel": "pcg_persistent", "passes": p[0]})
                plan["launches"].append({"kernel": "pcg_persistent", "passes": p[1]})
        F.close()
    json.dump(plan, open(plan_path, "w"), indent=1)
    print("plan written:", plan_path, plan["launches"])


def fit(csv_path, plan_path, out_path):
    plan = json.load(open(plan_path))
    rows = [r for r in csv.reader(l for l in open(csv_path) if not l.startswith("=="))]
    hdr = rows[0]
    ik, im, iv = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value")
    iid = hdr.index("ID")
    per = {}
    order = []
    for r in rows[1:]:
        if len(r) <= iv:
            continue
        key = int(r[iid])
        if key not in per:
            per[key] = {"kernel": r[ik]}
            order.append(key)
        per[key][r[im]] = float(r[iv].replace(",", ""))
    unit = {}
    iu = hdr.index("Metric Unit")
    for r in rows[1:]:
        if len(r) > iv:
            unit[r[im]] = r[iu]
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    launches = [per[k] for k in order if "pcg_" in per[k]["kernel"]]
    assert len(launches) == len(plan["launches"]), (len(launches), len(plan["launches"]))
    pts = {"pcg_dual": [], "pcg_persistent": []}
    for got, want in zip(launches, plan["launches"]):
        assert want["kernel"] in got["kernel"], (got["kernel"], want)
        by = got["dram__bytes_read.sum"] * scale[unit["dram__bytes_read.sum"]] + got["dram__bytes_write.sum"] * scale[unit["dram__bytes_write.sum"]]
        if want["passes"] > 0:
            pts[want["kernel"]].append((want["passes"], by, got["gpu__time_duration.sum"]))
    out = {"rows": plan["rows"], "dtype": plan["dtype"], "caps": list(CAPS)}
    for name, key in (("dual", "pcg_dual"), ("single", "pcg_persistent")):
        x = np.array([p[0] for p in pts[key]], float)
        y = np.array([p[1] for p in pts[key]], float)
        A = np.stack([np.ones_like(x), x], 1)
        (c0, c1), res, _, _ = np.linalg.lstsq(A, y, rcond=None)
        out[name] = {"fixed_bytes": float(c0), "bytes_per_pass": float(c1), "bytes_per_row_per_pass": float(c1 / plan["rows"]),
                     "points": [[int(a), float(b)] for a, b, _ in pts[key]],
                     "max_fit_error": float(np.max(np.abs(A @ np.array([c0, c1]) - y) / y))}
    out["source"] = ("ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none of tools/ncu_traffic.py run "
                     "(solves capped at 4/8/12/16 iterations); least-squares fit bytes = fixed + per_pass x passes")
    json.dump(out, open(out_path, "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    if sys.argv[1] == "run":
        run(sys.argv[2], *(int(a) for a in sys.argv[3:]))
    else:
        fit(*sys.argv[2:5])
