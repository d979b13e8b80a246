#!/bin/bash
# One 8-GPU session: headline bench at N=8/4/2, the BASELINE.json configs 2-4, the push/pull sweep.
# Every command has its own short timeout; logs go to gpurun_out/.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
run() { # name timeout nproc port cmd...
  local name=$1 to=$2 np=$3 port=$4; shift 4
  timeout $to $TR --nproc-per-node $np --master-port $port "$@" > gpurun_out/$name.log 2>&1
  echo "$name exit=$? $(grep '^{' gpurun_out/$name.log | tail -n 1 | cut -c1-420)"
}
nvidia-smi topo -m > gpurun_out/topo_8gpu.txt 2>&1
run bench_n8 120 8 29801 bench.py --gpus 8 --steps 200 --warmup 10
run bench_n4 120 4 29811 bench.py --gpus 4 --steps 200 --warmup 10
run bench_n2 120 2 29821 bench.py --gpus 2 --steps 200 --warmup 10
run bench_n8_rr 120 8 29831 bench.py --gpus 8 --steps 100 --warmup 10 --placement round_robin
run cfg2_mlp_ps1_async 120 8 29841 bench/ps_bench.py --model mlp --topology ps1 --mode async --steps 200 --warmup 10
run cfg3_mlp_sync_rr 120 8 29851 bench/ps_bench.py --model mlp --mode sync --placement round_robin --steps 200 --warmup 10 --ckpt-dir gpurun_out/ckpt_cfg3
run cnn_sync_flat 120 8 29861 bench/ps_bench.py --model cnn --mode sync --placement flat --batch 128 --steps 100 --warmup 10 --lr 0.0005
run cfg4_wide_async 240 8 29871 bench/ps_bench.py --model wide --mode async --placement flat --batch 4096 --hidden 8192 --steps 10 --warmup 3 --lr 0.0001
run cfg4_wide_sync 240 8 29881 bench/ps_bench.py --model wide --mode sync --placement flat --batch 4096 --hidden 8192 --steps 10 --warmup 3 --lr 0.0001
run sweep_n8 240 8 29891 bench/pushpull_sweep.py --out gpurun_out/sweep_n8.json
ls gpurun_out/ckpt_cfg3 2>/dev/null | head -12
rm -rf gpurun_out/ckpt_cfg3
