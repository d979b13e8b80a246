#!/bin/bash
# "owner applies" push bring-up on 2 GPUs: kernel + executor tests on GPU 0, then the multi-GPU tests,
# then rmw vs mailbox bench lines. Logs under gpurun_out/.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py tests/test_executor_gpu.py -q -x -p no:cacheprovider -k "mailbox" > gpurun_out/tests_mbox.log 2>&1
echo "mailbox unit tests exit=$? $(tail -n 15 gpurun_out/tests_mbox.log | cut -c1-300)"
timeout 600 python -m pytest tests/test_multigpu.py -q -x -p no:cacheprovider > gpurun_out/tests_multi.log 2>&1
echo "multigpu tests exit=$? $(tail -n 15 gpurun_out/tests_multi.log | cut -c1-400)"
for mode in rmw mailbox; do
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29701 \
    bench.py --gpus 2 --steps 200 --warmup 10 --push-mode $mode > gpurun_out/bench_n2_$mode.log 2>&1
  echo "bench N=2 [$mode] exit=$? $(grep '^{' gpurun_out/bench_n2_$mode.log | tail -n 1 | python -c 'import sys,json
try:
    d=json.loads(sys.stdin.read()); print(d["value"], d["ms_per_step"], d["back_to_back_ms_per_step"], d["e2e"]["value"], d["launches_per_step"], d["final_global_step"])
except Exception as e: print("parse error", e)')"
  tail -n 5 gpurun_out/bench_n2_$mode.log | cut -c1-300 | grep -v '^{'
done
