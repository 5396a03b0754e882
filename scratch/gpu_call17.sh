#!/bin/bash
# one GPU box visit: full GPU test suite, plain bench, ncu launch list + per-kernel DRAM metrics, full captures, 64M bench
O=gpurun_out
(timeout 900 python -m pytest tests -m gpu -x -q > $O/r01_s17_pytest.log 2>&1; echo "pytest exit $?" >> $O/r01_s17_pytest.log)
tail -4 $O/r01_s17_pytest.log
timeout 300 python bench.py --steps 10 --warmup 5 > $O/r01_s17_bench.json 2> $O/r01_s17_bench.err; echo "bench exit $?"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv \
  --log-file $O/r01_s17_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > $O/r01_s17_ncu_launches.log 2>&1; echo "ncu launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pcg_persistent -s 4 -c 2 -o $O/r01_s17_pcg_full \
  python bench.py --steps 2 --warmup 3 --no-cpu > $O/r01_s17_ncu_pcg.log 2>&1; echo "ncu pcg exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'rhs_init|cg_update|cg_direction|gather_values|finish_step|assemble_scatter|build_operator' -c 30 -o $O/r01_s17_small_full \
  python bench.py --steps 2 --warmup 3 --no-cpu > $O/r01_s17_ncu_small.log 2>&1; echo "ncu small exit $?"
timeout 600 python bench.py --nu 2530 --steps 5 --warmup 3 --no-cpu > $O/r01_s17_bench_64m.json 2> $O/r01_s17_bench_64m.err; echo "bench 64m exit $?"
ls -la $O | tail -15
