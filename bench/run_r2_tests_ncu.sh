cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/tests_gpu_all.log 2>&1
echo "pytest -m gpu exit=$? $(tail -n 2 gpurun_out/tests_gpu_all.log | cut -c1-200)"
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
echo "smoke exit=$? $(tail -n 1 gpurun_out/smoke.log | cut -c1-200)"
timeout 200 python bench.py --steps 3 --warmup 3 --no-cuda-graph > gpurun_out/bench_eager.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:dtf --launch-skip 39 --launch-count 26 \
  -o gpurun_out/ncu_step_r2 -f python bench.py --steps 3 --warmup 3 --no-cuda-graph > gpurun_out/ncu_step.log 2>&1
echo "ncu step exit=$?"
N=1638400 timeout 200 python bench/micro_mailbox.py > gpurun_out/micro_plain.log 2>&1 && \
N=1638400 timeout 900 ncu --set full --clock-control none --import-source on -k regex:mailbox_drain\|mailbox_push --launch-skip 30 --launch-count 4 \
  -o gpurun_out/ncu_mailbox_r2 -f python bench/micro_mailbox.py > gpurun_out/ncu_mailbox.log 2>&1
echo "ncu mailbox exit=$?"
ls -la gpurun_out/*.ncu-rep
