cd $GRAFT_REPO_ROOT
N=${1:-2}
mkdir -p gpurun_out
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29720 bench/pushpull_sweep.py --max-mb 1024 --out gpurun_out/pushpull_sweep_r2_n$N.json > gpurun_out/sweep_n$N.log 2>&1
echo "rc=$?"; tail -4 gpurun_out/sweep_n$N.log | cut -c1-900
