#!/bin/bash
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517"
(timeout 400 $TR tests/mp_partitioned.py > $O/r01_s28_mp2.log 2>&1; echo "mp exit $?" >> $O/r01_s28_mp2.log); tail -6 $O/r01_s28_mp2.log
FEAB_DUAL=0 timeout 300 $TR bench.py --gpus 2 --steps 10 --warmup 5 > $O/r01_s28_bench_n2_dual0.json 2> $O/r01_s28_bench_n2_dual0.err; echo "bench0 exit $?"; cut -c1-420 $O/r01_s28_bench_n2_dual0.json
FEAB_DUAL=1 timeout 300 $TR bench.py --gpus 2 --steps 10 --warmup 5 > $O/r01_s28_bench_n2_dual1.json 2> $O/r01_s28_bench_n2_dual1.err; echo "bench1 exit $?"; cut -c1-420 $O/r01_s28_bench_n2_dual1.json
tail -3 $O/r01_s28_bench_n2_dual1.err
