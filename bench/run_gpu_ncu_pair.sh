#!/bin/bash
# Executor tests + N=1 bench (host-path changes), then ONE ncu --set full capture of the CTA-pair GEMM.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_executor_gpu.py tests/test_kernels_gpu.py tests/test_zz_no_alloc.py -q -p no:cacheprovider > gpurun_out/tests.log 2>&1
echo "tests exit=$? $(tail -n 4 gpurun_out/tests.log | cut -c1-300)"
timeout 200 python bench.py --steps 200 --warmup 10 > gpurun_out/bench_base.log 2>&1
echo "bench exit=$? $(tail -n 1 gpurun_out/bench_base.log | python -c 'import sys,json
d=json.loads(sys.stdin.read()); print(d["value"], d["ms_per_step"], d["back_to_back_ms_per_step"], d["e2e"]["value"], d["launches_per_step"])')"
timeout 200 python bench.py --steps 200 --warmup 10 --push-mode mailbox > gpurun_out/bench_n1_mailbox.log 2>&1
echo "bench[mailbox] exit=$? $(tail -n 1 gpurun_out/bench_n1_mailbox.log | python -c 'import sys,json
d=json.loads(sys.stdin.read()); print(d["value"], d["ms_per_step"], d["back_to_back_ms_per_step"], d["e2e"]["value"], d["launches_per_step"], d["final_global_step"])')"
timeout 120 python bench/one_gemm.py 256 > gpurun_out/one_gemm.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_umma_pair -s 3 -c 1 \
  -o gpurun_out/ncu_pair_gemm -f python bench/one_gemm.py 256 > gpurun_out/ncu_pair.log 2>&1
echo "ncu exit=$? $(cat gpurun_out/one_gemm.log | tail -n 1)"
