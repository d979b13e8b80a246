#!/bin/bash
# bench.py with executor options toggled (cold-L2 value, ms/step, back-to-back, e2e, launches)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_executor_gpu.py tests/test_kernels_gpu.py tests/test_zz_no_alloc.py -q -x -p no:cacheprovider 2>&1 | tail -2
i=0
while IFS= read -r variant; do
  i=$((i+1))
  timeout 200 python bench.py --steps 200 --warmup 10 $variant > gpurun_out/bench_v$i.log 2>&1
  echo "bench[$variant] exit=$? $(tail -n 1 gpurun_out/bench_v$i.log | python -c 'import sys,json
try:
    d=json.loads(sys.stdin.read()); print(d["value"], d["ms_per_step"], d["back_to_back_ms_per_step"], d["e2e"]["value"], d["launches_per_step"])
except Exception as e: print("parse error", e)')"
done < "${1:-bench/variants.txt}"
