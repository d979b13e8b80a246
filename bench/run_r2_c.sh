cd $GRAFT_REPO_ROOT
N=${1:-2}
mkdir -p gpurun_out
export DTF_B200_HANG_DUMP=100
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29700 bench.py --gpus $N --steps 50 --warmup 10 --timeline gpurun_out/timeline_n$N.json > gpurun_out/tl_n$N.log 2>&1
echo rc=$?; grep -c '^{' gpurun_out/tl_n$N.log
timeout 150 python bench.py --gpus 1 --steps 50 --warmup 10 --timeline gpurun_out/timeline_n1.json > gpurun_out/tl_n1.log 2>&1
echo rc=$?
timeout 150 python bench.py --gpus 1 --steps 50 --warmup 10 --push-mode mailbox --timeline gpurun_out/timeline_n1_mbox.json 2>&1 | grep '^{' | cut -c1-200
