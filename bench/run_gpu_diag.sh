#!/bin/bash
# Runs every GPU kernel test group in its own process (a trapped kernel kills only its group) and
# collects the logs under gpurun_out/diag/. Usage: bash bench/run_gpu_diag.sh [pytest -k filters...]
mkdir -p gpurun_out/diag
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/diag/gpu.txt 2>&1
GROUPS_=("$@")
if [ ${#GROUPS_[@]} -eq 0 ]; then
  GROUPS_=(test_gemm_layouts test_gemm_bias test_gemm_relu test_gemm_split test_softmax test_push_apply test_pull test_elementwise test_conv1 test_conv2_fwd test_conv2_backward test_pool2)
fi
for g in "${GROUPS_[@]}"; do
  echo "=== $g"
  timeout 300 python -m pytest tests/test_kernels_gpu.py -q -x -k "$g" -p no:cacheprovider > gpurun_out/diag/$g.log 2>&1
  echo "exit=$? $(tail -n 3 gpurun_out/diag/$g.log | tr '\n' ' ')"
done
