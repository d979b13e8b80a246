#!/bin/bash
# 8-GPU session (round 1c): headline bench at N=8/4/2 + wide MLP async after the step-latency work.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
run() { # name timeout nproc port cmd...
  local name=$1 to=$2 np=$3 port=$4; shift 4
  timeout $to $TR --nproc-per-node $np --master-port $port "$@" > gpurun_out/$name.log 2>&1
  echo "$name exit=$? $(grep '^{' gpurun_out/$name.log | tail -n 1 | cut -c1-200)"
}
run bench_n8 150 8 29801 bench.py --gpus 8 --steps 200 --warmup 10
run bench_n4 150 4 29811 bench.py --gpus 4 --steps 200 --warmup 10
run bench_n2 150 2 29821 bench.py --gpus 2 --steps 200 --warmup 10
run cfg4_wide_async 240 8 29871 bench/ps_bench.py --model wide --mode async --placement flat --batch 4096 --hidden 8192 --steps 10 --warmup 3 --lr 0.0001
run torch_nccl_n8 150 8 29881 bench/nccl_ps_baseline.py --steps 100 --warmup 10
