#!/bin/bash
# One GPU round: executor tests, smoke, bench. Logs under gpurun_out/.
mkdir -p gpurun_out
echo "=== executor tests"
timeout 600 python -m pytest tests/test_executor_gpu.py -q -x -p no:cacheprovider > gpurun_out/executor.log 2>&1
echo "exit=$? $(tail -n 15 gpurun_out/executor.log)"
echo "=== smoke"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
echo "exit=$? $(tail -n 5 gpurun_out/smoke.log)"
echo "=== bench"
timeout 300 python bench.py --steps 200 --warmup 10 > gpurun_out/bench.log 2>&1
echo "exit=$? $(tail -n 5 gpurun_out/bench.log)"
