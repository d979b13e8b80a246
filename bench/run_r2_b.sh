# 2..8-GPU bench variants (each under its own timeout) + the multi-GPU tests
cd $GRAFT_REPO_ROOT
N=${1:-2}
mkdir -p gpurun_out
export DTF_B200_HANG_DUMP=100
run() {
  tag=$1; shift
  timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29700 bench.py --gpus $N --steps 200 --warmup 10 "$@" > gpurun_out/bench_n${N}_$tag.log 2>&1
  echo "rc=$? $tag"
  grep '^{' gpurun_out/bench_n${N}_$tag.log | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print('$tag', d['n_gpus'], d['value'], d['ms_per_step'], d['back_to_back_ms_per_step'], d['e2e']['value'], d['final_global_step'], d['final_loss'])
" || tail -30 gpurun_out/bench_n${N}_$tag.log
  grep -q '^{' gpurun_out/bench_n${N}_$tag.log || tail -40 gpurun_out/bench_n${N}_$tag.log
}
run mirror_start
run mirror_end --opt drain_at=end
run daemon_nomirror --opt mirror=0 --opt graph_drain=0
run mirror_daemon --opt graph_drain=0
if [ "$2" = "tests" ]; then
unset DTF_B200_HANG_DUMP
timeout 600 python -m pytest tests/test_multigpu.py -x -q 2>&1 | tail -15
fi
