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
    t = bench.ncu_solve_traffic(ROWS_16M, "f64", [16, 23], 16)
    assert t is not None
    assert bench.ncu_solve_traffic(ROWS_16M, "f64", [16, 23], 0) is None          # separate solves: other kernels, other traffic
    assert bench.ncu_solve_traffic(ROWS_16M, "f64", [17, 23], 16) is None
    raw = json.load(open(os.path.join(ROOT, "profiles", "r01_v7_solve_traffic.json")))
    assert raw["pcg_dual_dram_bytes_per_launch"] + raw["pcg_persistent_tail_dram_bytes_per_step"] == t
    # bench.py's model: (it_u + it_v) x 192 B/row minus the shared index stream (4 nnz + 4 n per lock-step pass)
    it_bytes_row = NNZ_16M / ROWS_16M * 12 + 4 + 13 * 8
    model = (16 + 23) * it_bytes_row * ROWS_16M - 16 * (4 * NNZ_16M + 4 * ROWS_16M)
    # ncu also sees the SpMV pass per solve that Eigen's iteration count leaves out ("Iteration accounting")
    assert 1.0 <= t / model <= 1.10


def test_measured_peak_has_a_source():
    peak, src = bench.measured_peak()
    assert 5000.0 < peak < 8000.0 and ("measured" in src or "fallback" in src)
