#!/bin/bash
# Kernel + executor tests, then the bench with feature toggles. Logs under gpurun_out/.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_executor_gpu.py tests/test_zz_no_alloc.py -q -x -p no:cacheprovider > gpurun_out/tests.log 2>&1
echo "tests exit=$? $(tail -n 6 gpurun_out/tests.log | cut -c1-400)"
for variant in "" "--no-fused-apply" "--no-overlap" "--no-pdl"; do
  tag=$(echo "base$variant" | tr -d ' ')
  timeout 200 python bench.py --steps 200 --warmup 10 $variant > gpurun_out/bench_$tag.log 2>&1
  echo "bench[$variant] exit=$? $(tail -n 1 gpurun_out/bench_$tag.log | python -c 'import sys,json
try:
    d=json.loads(sys.stdin.read()); print(d["value"], d["ms_per_step"], d["back_to_back_ms_per_step"], d["e2e"]["value"], d["launches_per_step"])
except Exception as e: print("parse error", e)')"
done
