#!/bin/bash
# 8-GPU session after the owner-applies push and the CTA-pair GEMM: headline bench at N=8 (both push
# modes) and N=4, config 2 (1 PS-only task + 7 workers), config 4 (wide MLP) async + sync.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
run() { # name timeout nproc port cmd...
  local name=$1 to=$2 np=$3 port=$4; shift 4
  timeout $to $TR --nproc-per-node $np --master-port $port "$@" > gpurun_out/$name.log 2>&1
  echo "$name exit=$? $(grep '^{' gpurun_out/$name.log | tail -n 1 | cut -c1-330)"
}
run bench_n8 150 8 29801 bench.py --gpus 8 --steps 200 --warmup 10
run bench_n8_rmw 150 8 29806 bench.py --gpus 8 --steps 200 --warmup 10 --push-mode rmw
run bench_n4 150 4 29811 bench.py --gpus 4 --steps 200 --warmup 10
run cfg2_mlp_ps1_async 150 8 29841 bench/ps_bench.py --model mlp --topology ps1 --mode async --steps 200 --warmup 10
run cfg4_wide_async 240 8 29871 bench/ps_bench.py --model wide --mode async --placement flat --batch 4096 --hidden 8192 --steps 10 --warmup 3 --lr 0.0001
run cfg4_wide_async_rmw 240 8 29876 bench/ps_bench.py --model wide --mode async --placement flat --batch 4096 --hidden 8192 --steps 10 --warmup 3 --lr 0.0001 --push-mode rmw
run cfg4_wide_sync 240 8 29881 bench/ps_bench.py --model wide --mode sync --placement flat --batch 4096 --hidden 8192 --steps 10 --warmup 3 --lr 0.0001
