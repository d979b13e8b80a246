cd $GRAFT_REPO_ROOT
N=${1:-2}
mkdir -p gpurun_out
export DTF_B200_HANG_DUMP=100


run() {
  tag=$1; shift
  timeout 170 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29700 bench.py --gpus $N --steps 200 --warmup 10 "$@" > gpurun_out/bench_n${N}_$tag.log 2>&1
  echo "rc=$? $tag"
  grep '^{' gpurun_out/bench_n${N}_$tag.log | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print('$tag', d['n_gpus'], d['value'], d['ms_per_step'], d['back_to_back_ms_per_step'], d['e2e']['value'], d['final_global_step'], d['final_loss'], d['e2e']['host_us_per_step_rank0'])
"
  grep -q '^{' gpurun_out/bench_n${N}_$tag.log || tail -40 gpurun_out/bench_n${N}_$tag.log
}
run dflt

