cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_multigpu.py -x -q -k "killed" 2>&1 | tail -15
timeout 200 python bench/gpu_cli_baseline.py --steps 20000 --workers 2 --ps 1 --gpus 3 --out gpurun_out/gpu_cli_1ps_2w_20000.json 2>&1 | tail -2 | cut -c1-900
