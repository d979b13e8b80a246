cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 280 python bench/gpu_cli_baseline.py --steps 1000 --workers 2 --ps 1 --gpus 2 --out gpurun_out/gpu_cli_1ps_2w.json 2>&1 | tail -30
timeout 200 python bench/gpu_cli_baseline.py --steps 1000 --workers 2 --ps 2 --gpus 2 --out gpurun_out/gpu_cli_2ps_2w.json 2>&1 | tail -30
