mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
run() { local name=$1 to=$2 np=$3 port=$4; shift 4
  timeout $to $TR --nproc-per-node $np --master-port $port "$@" > gpurun_out/$name.log 2>&1
  echo "$name exit=$? $(grep '^{' gpurun_out/$name.log | tail -n 1 | cut -c1-400)"; [ $? ] && tail -n 4 gpurun_out/$name.log | cut -c1-300 | grep -v "^{" | tail -n 3; }
run t_cnn_sync 100 2 29901 bench/ps_bench.py --model cnn --mode sync --placement flat --batch 128 --steps 50 --warmup 5 --lr 0.0005
run t_wide_async 150 2 29911 bench/ps_bench.py --model wide --mode async --placement flat --batch 1024 --hidden 2048 --steps 10 --warmup 3 --lr 0.0001
run t_wide_sync 150 2 29921 bench/ps_bench.py --model wide --mode sync --placement flat --batch 1024 --hidden 2048 --steps 10 --warmup 3 --lr 0.0001
run t_sweep 150 2 29931 bench/pushpull_sweep.py --max-mb 64 --out gpurun_out/sweep_t.json
grep -o '"bytes": [0-9]*\|"push_sgd_red_gbs": [0-9.]*' gpurun_out/t_sweep.log | paste - - | tail -4
