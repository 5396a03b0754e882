#!/bin/bash
O=gpurun_out
(timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "dual or loop_modes or step_parity" > $O/r01_s27_pytest_dual.log 2>&1; echo "pytest exit $?" >> $O/r01_s27_pytest_dual.log); tail -15 $O/r01_s27_pytest_dual.log
FEAB_DUAL=0 timeout 300 python bench.py --steps 10 --warmup 5 --no-cpu > $O/r01_s27_bench_dual0.json 2> $O/r01_s27_bench_dual0.err; echo "bench0 exit $?"; cat $O/r01_s27_bench_dual0.json
FEAB_DUAL=1 timeout 300 python bench.py --steps 10 --warmup 5 --no-cpu > $O/r01_s27_bench_dual1.json 2> $O/r01_s27_bench_dual1.err; echo "bench1 exit $?"; cat $O/r01_s27_bench_dual1.json
tail -5 $O/r01_s27_bench_dual1.err
