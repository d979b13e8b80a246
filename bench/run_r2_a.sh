set -x
cd $GRAFT_REPO_ROOT
python -m pytest tests/test_kernels_gpu.py -x -q -k "mailbox" 2>&1 | tail -15
python -m pytest tests/test_executor_gpu.py -x -q 2>&1 | tail -15
python bench.py --steps 50 --warmup 5 2>&1 | tail -2
