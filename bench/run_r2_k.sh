cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py -x -q -k "division or apply or mailbox" 2>&1 | tail -6
run() {
timeout 170 python bench.py --gpus 1 --steps 300 --warmup 10 $2 2>&1 | grep '^{' | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print('$1', d['value'], d['ms_per_step'], d['back_to_back_ms_per_step'], d['e2e']['value'], d['final_loss'])
"
}
run new1
run new2
