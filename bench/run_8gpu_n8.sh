#!/bin/bash
mkdir -p gpurun_out
timeout 150 python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 8 --master-port 29801 bench.py --gpus 8 --steps 200 --warmup 10 > gpurun_out/bench_n8.log 2>&1
echo "bench_n8 exit=$? $(grep '^{' gpurun_out/bench_n8.log | tail -n 1 | cut -c1-200)"
