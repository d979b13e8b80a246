#!/usr/bin/env python
"""Library baseline for the headline workload: the reference CNN (run_mnist_distributed.py:46-70,
batch 128 per worker, Adam 5e-4) written in stock PyTorch — cuDNN convolutions, cuBLAS GEMMs, bf16
autocast, autograd — with the parameter server restated on NCCL collectives
(distributedtensorflow_b200/parallel/nccl_baseline.py). Same timing rules as bench.py: CUDA events
on the launching stream, W warm-up steps, L2 flushed between timed steps, max over ranks.

    python bench/nccl_ps_baseline.py [--mode async|sync] [--steps 200] [--warmup 10] [--cuda-graph]
    torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench/nccl_ps_baseline.py ...

This is the comparator ("a solution that only calls NCCL ... is the baseline, not the product").
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import torch.nn.functional as F  # noqa: E402

from distributedtensorflow_b200.parallel.nccl_baseline import CollectivePs  # noqa: E402

SHAPES = [("conv1/kernel", (32, 1, 5, 5)), ("conv1/bias", (32,)), ("conv2/kernel", (64, 32, 5, 5)),
          ("conv2/bias", (64,)), ("dense/kernel", (1024, 3136)), ("dense/bias", (1024,)),
          ("dense_1/kernel", (10, 1024)), ("dense_1/bias", (10,))]


def forward(p, x_u8, y):
    x = x_u8.view(-1, 1, 28, 28).float().div_(255.0)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        h = F.max_pool2d(F.relu(F.conv2d(x, p["conv1/kernel"], p["conv1/bias"], padding=2)), 2)
        h = F.max_pool2d(F.relu(F.conv2d(h, p["conv2/kernel"], p["conv2/bias"], padding=2)), 2)
        h = F.relu(F.linear(h.flatten(1), p["dense/kernel"], p["dense/bias"]))
        logits = F.linear(h, p["dense_1/kernel"], p["dense_1/bias"])
    return F.cross_entropy(logits.float(), y)


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="async", choices=["async", "sync"])
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--cuda-graph", action="store_true", help="capture the step (N = 1 only)")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(0)
    ps = CollectivePs(SHAPES, dev, "adam", lr=5e-4)
    init = {n: (torch.randn(*s) * (1.0 / max(1, int(np.prod(s[1:]))) ** 0.5) if len(s) > 1 else torch.zeros(*s))
            for n, s in SHAPES}
    ps.init_from(init)
    rng = np.random.RandomState(3)
    x_pin = torch.from_numpy(rng.randint(0, 256, (args.batch, 784)).astype(np.uint8)).pin_memory()
    y_pin = torch.from_numpy(rng.randint(0, 10, (args.batch,)).astype(np.int64)).pin_memory()
    x_dev = torch.zeros_like(x_pin, device=dev)
    y_dev = torch.zeros_like(y_pin, device=dev)
    loss_dev = torch.zeros((), device=dev)
    stream = torch.cuda.Stream()

    def step():
        x_dev.copy_(x_pin, non_blocking=True)
        y_dev.copy_(y_pin, non_blocking=True)
        params = {n: t.detach().requires_grad_(True) for n, t in ps.pull().items()}
        loss = forward(params, x_dev, y_dev)
        grads = torch.autograd.grad(loss, list(params.values()))
        ps.set_grads(dict(zip(params.keys(), grads)))
        (ps.push_apply_async if args.mode == "async" else ps.push_apply_sync)()
        loss_dev.copy_(loss.detach())

    graph = None
    with torch.cuda.stream(stream):
        for _ in range(args.warmup):
            step()
        stream.synchronize()
        if args.cuda_graph and world == 1:
            graph = torch.cuda.CUDAGraph()
            b1p, b2p, gs = ps.b1p, ps.b2p, ps.global_step
            with torch.cuda.graph(graph, stream=stream):
                step()             # note: captures lr_t of this step (constant afterwards) — timing only
            ps.b1p, ps.b2p, ps.global_step = b1p, b2p, gs
        run = graph.replay if graph is not None else step
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        starts = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
        ends = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
        for i in range(args.steps):
            flush.fill_(i & 0xFF)
            starts[i].record(stream)
            run()
            ends[i].record(stream)
        stream.synchronize()
    dev_ms = sum(s.elapsed_time(e) for s, e in zip(starts, ends))
    t = torch.tensor([dev_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms = float(t.item())
    per_round = world if args.mode == "async" else 1
    if rank == 0:
        print(json.dumps({
            "impl": "torch+cudnn+nccl comparator", "metric": "global_steps_per_sec",
            "value": round(per_round * args.steps / (dev_ms / 1e3), 2), "n_gpus": world,
            "ms_per_step": round(dev_ms / args.steps, 5), "steps": args.steps, "warmup": args.warmup,
            "dtype": "bf16 autocast, fp32 master", "final_loss": round(float(loss_dev.item()), 5),
            "config": {"model": "mnist_cnn", "mode": args.mode, "batch_per_worker": args.batch,
                       "cuda_graph": graph is not None,
                       "ps": "flat shards; pull = all_gather, push = %s" % (
                           "all_to_all + W applies per owner" if args.mode == "async" else "reduce_scatter(AVG) + 1 apply")}}))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
