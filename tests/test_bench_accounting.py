"""CPU tests of bench.py's bookkeeping (no GPU): the committed ncu traffic files are the ones the roofline object reads,
they belong to the metric's workload, and they are consistent with the algorithmic byte model of DESIGN.md section 4."""
import json
import os

import bench

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ROWS_16M = 16002252          # icosphere nu = 1265 (BASELINE.json metric)
NNZ_16M = 7 * ROWS_16M - 12


def test_spmv_traffic_file_matches_the_metric_workload():
    t = bench.ncu_traffic(ROWS_16M, "f64")
    assert t is not None and bench.ncu_traffic(ROWS_16M, "f32") is None and bench.ncu_traffic(ROWS_16M + 1, "f64") is None
    algorithmic = NNZ_16M * 12 + (ROWS_16M + 1) * 4 + 2 * ROWS_16M * 8          # 104 B/row
    assert 0.95 * algorithmic <= t <= 1.05 * algorithmic                          # nothing is re-read


def test_solve_traffic_file_matches_the_metric_workload():
    """profiles/r02_solve_traffic.json holds ncu DRAM bytes per lock-step pass / per single-system pass (fit over solves capped
    at 4 / 8 / 12 / 16 iterations); bench.py scales them to the run's own pass counts, so the driver's run gets a figure
    whatever its iteration counts are."""
    raw = json.load(open(os.path.join(ROOT, "profiles", "r02_solve_traffic.json")))
    assert raw["rows"] == ROWS_16M and raw["dual"]["max_fit_error"] < 0.01 and raw["single"]["max_fit_error"] < 0.01
    it_bytes_row = NNZ_16M / ROWS_16M * 12 + 4 + 13 * 8                           # 192 B/row: bench.py's model of one iteration
    shared_saving = (4 * NNZ_16M + 4 * ROWS_16M) / ROWS_16M                      # index stream read once per lock-step pass
    assert 0.93 <= raw["single"]["bytes_per_row_per_pass"] / it_bytes_row <= 1.02   # x += alpha p deferred: one vector read less
    assert 0.93 <= raw["dual"]["bytes_per_row_per_pass"] / (2 * it_bytes_row - shared_saving) <= 1.02
    for its, shared in (([15, 22], 16), ([16, 23], 17), ([16, 23], 0)):
        passes = [its[0] + 1, its[1] + 1]                                        # converged solves: one SpMV pass more than iterations()
        t, src = bench.ncu_solve_traffic(ROWS_16M, "f64", passes, shared)
        assert t is not None and "r02_solve_traffic" in src
        model = (its[0] + its[1]) * it_bytes_row * ROWS_16M - min(shared, *its) * (4 * NNZ_16M + 4 * ROWS_16M)
        assert 0.95 <= t / model <= 1.08, (its, shared, t / model)                # nothing is re-read
    assert bench.ncu_solve_traffic(ROWS_16M + 1, "f64", [16, 23], 16)[0] is None   # another workload: no figure
    assert bench.ncu_solve_traffic(ROWS_16M, "f32", [16, 23], 16)[0] is None


def test_measured_peak_has_a_source():
    peak, src = bench.measured_peak()
    assert 5000.0 < peak < 8000.0 and ("measured" in src or "fallback" in src)


def test_final_build_ncu_capture_agrees_with_the_traffic_model():
    """profiles/r02_v22_ncu_full.json: `ncu --set full` of one pcg_dual and one pcg_persistent launch of the FINAL round-2 build.
    Their DRAM bytes must be a whole number of passes of the per-pass figures bench.py scales (profiles/r02_solve_traffic.json,
    captured on an earlier build): the kernels changed (row order of phase C, one-GPU syncs), the traffic did not."""
    cap = json.load(open(os.path.join(ROOT, "profiles", "r02_v22_ncu_full.json")))
    raw = json.load(open(os.path.join(ROOT, "profiles", "r02_solve_traffic.json")))
    assert cap["rows"] == raw["rows"] == ROWS_16M
    passes = {}
    for kernel, fit in (("pcg_dual", raw["dual"]), ("pcg_persistent", raw["single"])):
        k = cap["kernels"][kernel]
        total = k["dram_bytes_read"] + k["dram_bytes_write"]
        p = (total - fit["fixed_bytes"]) / fit["bytes_per_pass"]
        passes[kernel] = round(p)
        assert abs(p - round(p)) < 0.15, (kernel, p)               # within 1 % of a whole number of passes
        gbps = total / k["duration_ms"] / 1e6
        assert gbps > 5800.0, (kernel, gbps)                       # 6 228 / 6 030 GB/s: 0.95 / 0.92 of the 6 551.7 GB/s copy peak of that pool
        assert k["registers"] == 64 and k["grid"] == 296 and k["block"] == 512   # 2 resident CTAs per SM x 148 SMs
    # the captured step is one of the seed-protocol steps: the u solve stops first, the tail finishes the v solve
    assert 15 <= passes["pcg_dual"] <= 19 and 5 <= passes["pcg_persistent"] <= 9
