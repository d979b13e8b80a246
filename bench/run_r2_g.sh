cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
export DTF_B200_HANG_DUMP=100
run() {
  tag=$1; shift
  timeout 170 python bench.py --gpus 1 --steps 200 --warmup 10 "$@" > gpurun_out/bench_n1_$tag.log 2>&1
  echo "rc=$? $tag"
  grep '^{' gpurun_out/bench_n1_$tag.log | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print('$tag', d['n_gpus'], d['value'], d['ms_per_step'], d['back_to_back_ms_per_step'], d['e2e']['value'], d['final_global_step'], d['final_loss'])
"
  grep -q '^{' gpurun_out/bench_n1_$tag.log || tail -40 gpurun_out/bench_n1_$tag.log
}
run base
run ea0 --opt early_apply=1
run ea64 --opt early_apply=1 --opt early_apply_ctas=64
run ea148 --opt early_apply=1 --opt early_apply_ctas=148
run ea296 --opt early_apply=1 --opt early_apply_ctas=296
run mbox --push-mode mailbox
run mbox_end --push-mode mailbox --opt drain_at=end --opt drain_ctas=0
