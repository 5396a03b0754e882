#!/bin/bash
O=gpurun_out
(timeout 100 python -m pytest tests/test_parity_gpu.py tests/test_abi.py -m gpu -x -q -k "dual or abi or struct" > $O/r01_s36_pytest.log 2>&1; echo "pytest exit $?" >> $O/r01_s36_pytest.log); tail -3 $O/r01_s36_pytest.log
timeout 100 python bench.py --steps 5 --warmup 3 --no-cpu > $O/r01_s36_bench.json 2> $O/r01_s36_bench.err; echo "bench exit $?"; python -c "
import json; j=json.loads(open('$O/r01_s36_bench.json').read().strip().splitlines()[-1]); r=j['roofline']; print(j['value'], r['kernel'], r['achieved'], r['frac'], r['traffic'], r['algorithmic_bytes'], r['spmv_dot_alone']['frac'])"
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
