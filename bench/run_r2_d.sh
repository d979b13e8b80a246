cd $GRAFT_REPO_ROOT
N=${1:-2}
mkdir -p gpurun_out
export DTF_B200_HANG_DUMP=100
timeout 200 python -m pytest tests/test_kernels_gpu.py -x -q -k "mailbox" 2>&1 | tail -3
run() {
  tag=$1; shift
  timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29700 bench.py --gpus $N --steps 200 --warmup 10 "$@" > gpurun_out/bench_n${N}_$tag.log 2>&1
  echo "rc=$? $tag"
  grep '^{' gpurun_out/bench_n${N}_$tag.log | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print('$tag', d['n_gpus'], d['value'], d['ms_per_step'], d['back_to_back_ms_per_step'], d['e2e']['value'], d['final_global_step'], d['final_loss'])
"
  grep -q '^{' gpurun_out/bench_n${N}_$tag.log || tail -40 gpurun_out/bench_n${N}_$tag.log
}
run c0 --opt drain_ctas=0
run c64 --opt drain_ctas=64
run c148 --opt drain_ctas=148
run c0_end --opt drain_ctas=0 --opt drain_at=end
run c148_tl --opt drain_ctas=148 --timeline gpurun_out/timeline_n${N}_c148.json
timeout 200 python bench/micro_mailbox.py 2>&1 | grep -v "cap= 32" | tail -40
