#!/bin/bash
O=gpurun_out
(timeout 900 python -m pytest tests -m gpu -x -q > $O/r01_s33_pytest.log 2>&1; echo "pytest exit $?" >> $O/r01_s33_pytest.log); tail -3 $O/r01_s33_pytest.log
timeout 300 python bench.py --impl reference --steps 5 --warmup 3 > $O/r01_s33_bench_reference.json 2> $O/r01_s33_bench_reference.err; echo "ref exit $?"
timeout 300 python bench.py --steps 10 --warmup 5 > $O/r01_s33_bench.json 2> $O/r01_s33_bench.err; echo "bench exit $?"; cut -c1-260 $O/r01_s33_bench.json
timeout 300 python bench_configs.py --only C1,C2 > $O/r01_s33_configs.jsonl 2> $O/r01_s33_configs.err; echo "configs exit $?"; cut -c1-600 $O/r01_s33_configs.jsonl
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
