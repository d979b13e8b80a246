#!/bin/bash
mkdir -p gpurun_out
timeout 200 python bench/profile_host_path.py 2000 > gpurun_out/host_profile.log 2>&1
echo "host profile exit=$?"; grep -E "e2e per step" gpurun_out/host_profile.log
for extra in "" "--cuda-graph"; do
  timeout 200 python bench/nccl_ps_baseline.py --steps 200 --warmup 10 $extra > gpurun_out/torch_baseline_n1$extra.log 2>&1
  echo "torch baseline [$extra] exit=$? $(grep '^{' gpurun_out/torch_baseline_n1$extra.log | tail -n 1 | cut -c1-260)"
done
