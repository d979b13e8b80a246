#!/bin/bash
# bench.py with executor options toggled one at a time (cold-L2 value, ms/step, back-to-back, e2e, launches)
mkdir -p gpurun_out
for variant in "" "--opt early_apply=1" "--opt zero_copy_input=0" "--opt host_sink=0" "--opt fused_head=0" "--no-l2-flush"; do
  tag=$(echo "v$variant" | tr -d ' =-')
  timeout 200 python bench.py --steps 200 --warmup 10 $variant > gpurun_out/bench_$tag.log 2>&1
  echo "bench[$variant] exit=$? $(tail -n 1 gpurun_out/bench_$tag.log | python -c 'import sys,json
try:
    d=json.loads(sys.stdin.read()); print(d["value"], d["ms_per_step"], d["back_to_back_ms_per_step"], d["e2e"]["value"], d["launches_per_step"])
except Exception as e: print("parse error", e)')"
done
