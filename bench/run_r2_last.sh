cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 55 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/tests_gpu_last.log 2>&1
echo "pytest -m gpu exit=$? $(tail -n 1 gpurun_out/tests_gpu_last.log | cut -c1-200)"
