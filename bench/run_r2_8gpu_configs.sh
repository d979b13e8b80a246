cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
export DTF_B200_HANG_DUMP=150
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
one() { tag=$1; shift; timeout 200 $TR --nproc-per-node 8 --master-port 29871 "$@" > gpurun_out/$tag.log 2>&1; echo "$tag exit=$? $(grep '^{' gpurun_out/$tag.log | tail -n 1 | cut -c1-330)"; grep -q '^{' gpurun_out/$tag.log || tail -25 gpurun_out/$tag.log; }
one r2_cfg4_wide_async bench/ps_bench.py --model wide --mode async --placement flat --batch 4096 --hidden 8192 --steps 100 --warmup 5 --lr 0.0001
one r2_cfg4_wide_async_fusedpush bench/ps_bench.py --model wide --mode async --placement flat --batch 4096 --hidden 8192 --steps 100 --warmup 5 --lr 0.0001 --opt fused_push=1
one r2_cfg4_wide_sync bench/ps_bench.py --model wide --mode sync --placement flat --batch 4096 --hidden 8192 --steps 100 --warmup 5 --lr 0.0001
one r2_cfg2_mlp_ps1_async bench/ps_bench.py --model mlp --topology ps1 --steps 200 --warmup 10
one r2_cfg3_mlp_sync_rr bench/ps_bench.py --model mlp --mode sync --placement round_robin --steps 200 --warmup 10 --ckpt-dir /tmp/r2_ckpt
one r2_bench_n8 bench.py --gpus 8 --steps 1000 --warmup 10
timeout 200 python bench/gpu_cli_baseline.py --steps 1000 --workers 4 --ps 4 --gpus 8 --out gpurun_out/gpu_cli_4ps_4w.json 2>&1 | tail -3 | cut -c1-600
timeout 200 python bench/gpu_cli_baseline.py --steps 1000 --workers 2 --ps 1 --gpus 3 --out gpurun_out/gpu_cli_1ps_2w.json 2>&1 | tail -3 | cut -c1-600
