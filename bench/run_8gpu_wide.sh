#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 240 $TR --nproc-per-node 8 --master-port 29871 bench/ps_bench.py --model wide --mode async --placement flat --batch 4096 --hidden 8192 --steps 10 --warmup 3 --lr 0.0001 > gpurun_out/cfg4_wide_async.log 2>&1
echo "wide async exit=$? $(grep '^{' gpurun_out/cfg4_wide_async.log | tail -n 1 | cut -c1-200)"
timeout 240 $TR --nproc-per-node 8 --master-port 29881 bench/ps_bench.py --model wide --mode sync --placement flat --batch 4096 --hidden 8192 --steps 10 --warmup 3 --lr 0.0001 > gpurun_out/cfg4_wide_sync.log 2>&1
echo "wide sync exit=$? $(grep '^{' gpurun_out/cfg4_wide_sync.log | tail -n 1 | cut -c1-200)"
