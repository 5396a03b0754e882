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
