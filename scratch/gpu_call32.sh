#!/bin/bash
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29525"
(timeout 300 $TR tests/mp_partitioned.py > $O/r01_s32_mp2.log 2>&1; echo "mp exit $?" >> $O/r01_s32_mp2.log); tail -2 $O/r01_s32_mp2.log
export FEAB_BENCH_NU=632
FEAB_DUAL=1 FEAB_PCG_PROF=2 timeout 200 $TR bench.py --gpus 2 --steps 20 --warmup 5 > $O/r01_s32_n2_small_dual1.json 2> $O/r01_s32_n2_small_dual1.err; echo "exit $?"; cut -c1-200 $O/r01_s32_n2_small_dual1.json; grep "phase profile" $O/r01_s32_n2_small_dual1.err
FEAB_DUAL=1 timeout 200 $TR bench.py --gpus 2 --steps 20 --warmup 5 > $O/r01_s32_n2_small_dual1_noprof.json 2> /dev/null; cut -c1-200 $O/r01_s32_n2_small_dual1_noprof.json
