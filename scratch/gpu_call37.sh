#!/bin/bash
O=gpurun_out
(timeout 200 python -m pytest tests/test_parity_gpu.py tests/test_configs_gpu.py -m gpu -x -q > $O/r01_s37_pytest.log 2>&1; echo "pytest exit $?" >> $O/r01_s37_pytest.log); tail -3 $O/r01_s37_pytest.log
timeout 100 python bench.py --steps 10 --warmup 5 --no-cpu > $O/r01_s37_bench.json 2> $O/r01_s37_bench.err; echo "bench exit $?"; python -c "
import json; j=json.loads(open('$O/r01_s37_bench.json').read().strip().splitlines()[-1]); r=j['roofline']; print(j['value'], j['e2e']['value'], r['frac'], j['clocks'])"
timeout 120 python bench_configs.py --only C1,C2 --no-cpu > $O/r01_s37_configs.jsonl 2> $O/r01_s37_configs.err; python -c "
import json
for l in open('$O/r01_s37_configs.jsonl'):
    j=json.loads(l); print(j['config'], j['steps_per_s'], j['ms_per_step'], j['iterations_last_step'])"
