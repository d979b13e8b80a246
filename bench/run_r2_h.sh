cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 150 python -m pytest tests/test_executor_gpu.py -x -q -s -k "sync" > gpurun_out/sync_dbg.log 2>&1
echo rc=$?
grep -v "^\[sync\|^frame" gpurun_out/sync_dbg.log | tail -30
grep "^\[sync b200\] worker 2" gpurun_out/sync_dbg.log | tail -3
