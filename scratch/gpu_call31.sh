#!/bin/bash
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29523"
export FEAB_BENCH_NU=632
FEAB_DUAL=0 FEAB_PCG_PROF=1 timeout 200 $TR bench.py --gpus 2 --steps 20 --warmup 5 > $O/r01_s31_n2_small_dual0.json 2> $O/r01_s31_n2_small_dual0.err; echo "exit $?"; cut -c1-200 $O/r01_s31_n2_small_dual0.json; grep "phase profile" $O/r01_s31_n2_small_dual0.err
FEAB_DUAL=1 FEAB_PCG_PROF=2 timeout 200 $TR bench.py --gpus 2 --steps 20 --warmup 5 > $O/r01_s31_n2_small_dual1.json 2> $O/r01_s31_n2_small_dual1.err; echo "exit $?"; cut -c1-200 $O/r01_s31_n2_small_dual1.json; grep "phase profile" $O/r01_s31_n2_small_dual1.err
FEAB_DUAL=1 timeout 200 $TR bench.py --gpus 2 --steps 20 --warmup 5 > $O/r01_s31_n2_small_dual1_noprof.json 2> /dev/null; cut -c1-200 $O/r01_s31_n2_small_dual1_noprof.json
FEAB_DUAL=0 timeout 200 $TR bench.py --gpus 2 --steps 20 --warmup 5 > $O/r01_s31_n2_small_dual0_noprof.json 2> /dev/null; cut -c1-200 $O/r01_s31_n2_small_dual0_noprof.json
