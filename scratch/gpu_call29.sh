#!/bin/bash
O=gpurun_out
(timeout 900 python -m pytest tests -m gpu -x -q > $O/r01_s29_pytest.log 2>&1; echo "pytest exit $?" >> $O/r01_s29_pytest.log); tail -4 $O/r01_s29_pytest.log
timeout 300 python bench.py --steps 10 --warmup 5 > $O/r01_s29_bench.json 2> $O/r01_s29_bench.err; echo "bench exit $?"; cut -c1-300 $O/r01_s29_bench.json
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv \
  --log-file $O/r01_s29_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > $O/r01_s29_ncu_launches.log 2>&1; echo "ncu launches exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'pcg_dual' -s 3 -c 1 -o $O/r01_s29_pcg_dual_full \
  python bench.py --steps 2 --warmup 3 --no-cpu > $O/r01_s29_ncu_dual.log 2>&1; echo "ncu dual exit $?"
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
