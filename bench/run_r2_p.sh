cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_executor_gpu.py tests/test_kernels_gpu.py -x -q 2>&1 | tail -4
timeout 170 python bench.py --gpus 1 --steps 1000 --warmup 10 > gpurun_out/r2_bench_n1.log 2>&1; grep '^{' gpurun_out/r2_bench_n1.log | cut -c1-120
timeout 200 python bench.py --steps 3 --warmup 3 --no-cuda-graph > gpurun_out/bench_eager.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:kernel --launch-skip 39 --launch-count 26 \
  -o gpurun_out/ncu_step_r2 -f python bench.py --steps 3 --warmup 3 --no-cuda-graph > gpurun_out/ncu_step.log 2>&1
echo "ncu step exit=$?"; ls -la gpurun_out/ncu_step_r2.ncu-rep
