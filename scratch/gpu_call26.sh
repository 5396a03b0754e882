#!/bin/bash
O=gpurun_out
(timeout 900 python -m pytest tests -m gpu -x -q > $O/r01_s26_pytest.log 2>&1; echo "pytest exit $?" >> $O/r01_s26_pytest.log); tail -4 $O/r01_s26_pytest.log
timeout 300 python bench.py --impl reference --steps 5 --warmup 3 > $O/r01_s26_bench_reference.json 2> $O/r01_s26_bench_reference.err; echo "ref exit $?"
timeout 300 python bench.py --steps 10 --warmup 5 > $O/r01_s26_bench.json 2> $O/r01_s26_bench.err; echo "bench exit $?"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv \
  --log-file $O/r01_s26_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > $O/r01_s26_ncu_launches.log 2>&1; echo "ncu launches exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'rhs_init_tma|spmv_dot_tma' -c 6 -o $O/r01_s26_rhs_spmv_full \
  python bench.py --steps 2 --warmup 3 --no-cpu > $O/r01_s26_ncu_rhs.log 2>&1; echo "ncu rhs exit $?"
python -c "import __graft_entry__ as g; g.smoke()"
