#!/bin/bash
# CTA-pair GEMM bring-up: its tests first (short timeout), then the kernel/executor suites, the
# kernel micro-benchmarks and the headline bench. Logs under gpurun_out/.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py -q -x -p no:cacheprovider -k "pair" > gpurun_out/tests_pair.log 2>&1
echo "pair tests exit=$? $(tail -n 12 gpurun_out/tests_pair.log | cut -c1-300)"
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_executor_gpu.py tests/test_zz_no_alloc.py -q -p no:cacheprovider -k "not pair" > gpurun_out/tests.log 2>&1
echo "tests exit=$? $(tail -n 6 gpurun_out/tests.log | cut -c1-400)"
timeout 300 python bench/kernels_bench.py --out gpurun_out/kernels_bench.json > gpurun_out/kernels_bench.log 2>&1
echo "kernels_bench exit=$?"; grep -E "gemm_umma|conv1" gpurun_out/kernels_bench.log | cut -c1-330
timeout 200 python bench.py --steps 200 --warmup 10 > gpurun_out/bench_base.log 2>&1
echo "bench exit=$? $(tail -n 1 gpurun_out/bench_base.log | cut -c1-600)"
