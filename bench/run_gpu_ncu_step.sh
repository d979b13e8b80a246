#!/bin/bash
# Final single-GPU verification + ONE ncu --set full capture of the 13 kernels of one (eager) CNN step.
mkdir -p gpurun_out
timeout 600 python -m pytest tests -q -m gpu -p no:cacheprovider > gpurun_out/tests_gpu_all.log 2>&1
echo "pytest -m gpu exit=$? $(tail -n 2 gpurun_out/tests_gpu_all.log | cut -c1-200)"
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
echo "smoke exit=$? $(tail -n 1 gpurun_out/smoke.log | cut -c1-200)"
timeout 200 python bench.py --steps 200 --warmup 10 > gpurun_out/bench_base.log 2>&1
echo "bench exit=$? $(tail -n 1 gpurun_out/bench_base.log | cut -c1-160)"
timeout 200 python bench.py --steps 2 --warmup 3 --no-cuda-graph > gpurun_out/bench_eager.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:dtf --launch-skip 52 --launch-count 13 \
  -o gpurun_out/ncu_step_r1c -f python bench.py --steps 2 --warmup 3 --no-cuda-graph > gpurun_out/ncu_step.log 2>&1
echo "ncu exit=$?"
