cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
export DTF_B200_HANG_DUMP=150
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
one() { tag=$1; shift; timeout 200 $TR --nproc-per-node 2 --master-port 29871 "$@" > gpurun_out/$tag.log 2>&1; echo "$tag exit=$? $(grep '^{' gpurun_out/$tag.log | tail -n 1 | cut -c1-130)"; grep -q '^{' gpurun_out/$tag.log || tail -25 gpurun_out/$tag.log; }
one w2_c64 bench/ps_bench.py --model wide --mode async --placement flat --batch 4096 --hidden 8192 --steps 30 --warmup 5 --lr 0.0001 --timeline gpurun_out/tl_wide_n2_c64.json
one w2_c0 bench/ps_bench.py --model wide --mode async --placement flat --batch 4096 --hidden 8192 --steps 30 --warmup 5 --lr 0.0001 --opt drain_ctas=0 --timeline gpurun_out/tl_wide_n2_c0.json
one w2_daemon bench/ps_bench.py --model wide --mode async --placement flat --batch 4096 --hidden 8192 --steps 30 --warmup 5 --lr 0.0001 --opt graph_drain=0 --opt mirror=0
