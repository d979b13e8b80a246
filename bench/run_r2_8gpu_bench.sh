cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
export DTF_B200_HANG_DUMP=150
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
one() { tag=$1; shift; timeout 200 $TR --nproc-per-node 8 --master-port 29871 "$@" > gpurun_out/$tag.log 2>&1; echo "$tag exit=$? $(grep '^{' gpurun_out/$tag.log | tail -n 1 | cut -c1-200)"; grep -q '^{' gpurun_out/$tag.log || tail -25 gpurun_out/$tag.log; }
one r2f_bench_n8_20 bench.py --gpus 8 --steps 20 --warmup 3
one r2f_bench_n8_200 bench.py --gpus 8 --steps 200 --warmup 10 --timeline gpurun_out/timeline_cnn_step_r2_n8.json
one r2f_bench_n8_1000 bench.py --gpus 8 --steps 1000 --warmup 10
one r2f_cfg4_wide_async bench/ps_bench.py --model wide --mode async --placement flat --batch 4096 --hidden 8192 --steps 100 --warmup 5 --lr 0.0001
