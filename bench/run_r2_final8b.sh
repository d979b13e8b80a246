cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
export DTF_B200_HANG_DUMP=100
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29871 bench.py --gpus 8 --steps 200 --warmup 10 > gpurun_out/r2g_bench_n8_200.log 2>&1
echo "exit=$? $(grep '^{' gpurun_out/r2g_bench_n8_200.log | tail -n 1 | cut -c1-200)"
