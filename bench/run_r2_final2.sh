cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
export DTF_B200_HANG_DUMP=100
run() { tag=$1; shift; envs=$1; shift
env $envs timeout 100 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29871 bench.py --gpus 2 --steps 200 --warmup 5 "$@" > gpurun_out/bis_$tag.log 2>&1
echo "$tag $(grep '^{' gpurun_out/bis_$tag.log | tail -n 1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['back_to_back_ms_per_step'], d['e2e']['value'])")"; }
run c0 X=1 --opt drain_ctas=0
run c148 X=1 --opt drain_ctas=148
