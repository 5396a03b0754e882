#!/bin/bash
# One GPU-box visit: GPU test suite, then the driver's bench command (both arms optional).  usage: tools/gpu_check.sh <tag> [pytest args]
TAG=${1:-run}; shift
O=gpurun_out
mkdir -p $O
(timeout 1700 python -m pytest tests -m gpu -x -q "$@" > $O/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest.log); tail -15 $O/${TAG}_pytest.log
timeout 900 python bench.py --steps 10 --warmup 5 ${BENCH_ARGS:---no-cpu} > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench exit $?"; tail -5 $O/${TAG}_bench.err
python - <<PY
import json
try:
    j = json.loads(open('$O/${TAG}_bench.json').read().strip().splitlines()[-1]); r = j['roofline']
    print('value', j['value'], 'e2e', j['e2e'], 'frac', r['frac'], 'traffic', r['traffic'], 'its', j['config'].get('pcg_iterations_last_step'), j['clocks'])
    print('parity', j['parity']['rel_l2_max'], j['parity']['ok'], 'cpu', j['cpu_baseline'])
except Exception as e:
    print('no bench line', e)
PY
