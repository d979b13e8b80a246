cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 500 python -m pytest tests/test_executor_gpu.py tests/test_kernels_gpu.py -x -q 2>&1 | tail -15
