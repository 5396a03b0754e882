#!/bin/bash
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519"
timeout 300 $TR bench.py --gpus 8 --steps 10 --warmup 5 > $O/r01_s30_bench_n8.json 2> $O/r01_s30_bench_n8.err; echo "bench8 exit $?"; cut -c1-330 $O/r01_s30_bench_n8.json
TR4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29521"
timeout 300 $TR4 bench.py --gpus 4 --steps 10 --warmup 5 > $O/r01_s30_bench_n4.json 2> $O/r01_s30_bench_n4.err; echo "bench4 exit $?"; cut -c1-330 $O/r01_s30_bench_n4.json
