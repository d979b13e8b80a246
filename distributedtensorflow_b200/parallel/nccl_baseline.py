"""Parameter server on library collectives — the BASELINE COMPARATOR, not the product.

SURVEY 5.9 / BASELINE.json north star: "torch.distributed NCCL (GPU) and gloo (CPU) implementation
of the same PS API = the baseline comparator". The reference moves parameters and gradients with
per-tensor gRPC RecvTensor calls (run_mnist_distributed.py:150,161); the stock way to do the same
job on an NVLink box is collectives. This module restates the PS step with them so the peer-memory
kernels (``parallel/nvl.py``, ``csrc/kernels/{pull,apply,mailbox,sync_step}.cu``) have a library
baseline to be measured against on identical shapes:

    pull                         all_gather_into_tensor of the flat fp32 shards
    push, synchronous            reduce_scatter_tensor(AVG) of the flat gradient -> one apply per shard
    push, asynchronous-emulated  all_to_all_single of the gradient slices -> every owner applies the
                                 W gradients it received one after another (what the apply daemon does)

Every rank is one PS shard and one worker (the flat-sliced placement). Optimizer rules are TF1's
(``train/optimizers.py``). Works with backend "gloo" on CPU (tests) and "nccl" on GPUs (bench).
"""
from __future__ import annotations

import math
from typing import Dict, Sequence, Tuple

import torch
import torch.distributed as dist


class CollectivePs:
    def __init__(self, shapes: Sequence[Tuple[str, Sequence[int]]], device, optimizer: str = "adam",
                 lr: float = 5e-4, beta1: float = 0.9, beta2: float = 0.999, epsilon: float = 1e-8,
                 momentum: float = 0.9, group=None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.device = torch.device(device)
        self.shapes = [(n, tuple(int(d) for d in s)) for n, s in shapes]
        self.offsets: Dict[str, Tuple[int, int]] = {}
        o = 0
        for n, s in self.shapes:
            k = int(math.prod(s)) if len(s) else 1
            self.offsets[n] = (o, k)
            o += k
        self.n = o
        per = (o + self.world - 1) // self.world
        self.per = (per + 63) // 64 * 64                      # equal, aligned shards (tail is padding)
        self.padded = self.per * self.world
        self.kind, self.lr, self.b1, self.b2, self.eps, self.mom = optimizer, lr, beta1, beta2, epsilon, momentum
        f32 = dict(dtype=torch.float32, device=self.device)
        self.shard = torch.zeros(self.per, **f32)             # fp32 master slice owned by this rank
        self.slot0 = torch.full((self.per,), 0.1 if optimizer == "adagrad" else 0.0, **f32)
        self.slot1 = torch.zeros(self.per, **f32)
        self.full = torch.zeros(self.padded, **f32)           # the worker's pulled copy
        self.grad = torch.zeros(self.padded, **f32)           # the worker's flat gradient
        self.recv = torch.zeros(self.padded, **f32)           # all_to_all receive buffer (W slices)
        self.b1p, self.b2p = beta1, beta2                     # shared beta powers (TF1 Adam)
        self.global_step = 0

    # -- init / views ---------------------------------------------------------------------------------
    def init_from(self, values: Dict[str, torch.Tensor]) -> None:
        flat = torch.zeros(self.padded, dtype=torch.float32, device=self.device)
        for n, (o, k) in self.offsets.items():
            flat[o:o + k] = values[n].reshape(-1).to(self.device, torch.float32)
        self.shard.copy_(flat[self.rank * self.per:(self.rank + 1) * self.per])
        self.full.copy_(flat)

    def views(self, flat: torch.Tensor) -> Dict[str, torch.Tensor]:
        return {n: flat[o:o + k].view(s) for (n, s), (o, k) in zip(self.shapes, self.offsets.values())}

    # -- the PS API -------------------------------------------------------------------------------------
    def pull(self) -> Dict[str, torch.Tensor]:
        if self.world > 1:
            dist.all_gather_into_tensor(self.full, self.shard, group=self.group)
        else:
            self.full.copy_(self.shard)
        return self.views(self.full)

    def _apply(self, g: torch.Tensor) -> None:
        v, s0, s1 = self.shard, self.slot0, self.slot1
        if self.kind == "sgd":
            v.add_(g, alpha=-self.lr)
        elif self.kind == "momentum":
            s0.mul_(self.mom).add_(g)
            v.add_(s0, alpha=-self.lr)
        elif self.kind == "adagrad":
            s0.addcmul_(g, g)
            v.addcdiv_(g, s0.sqrt(), value=-self.lr)
        else:
            lr_t = self.lr * math.sqrt(1.0 - self.b2p) / (1.0 - self.b1p)
            s0.add_(g - s0, alpha=1.0 - self.b1)
            s1.add_(g * g - s1, alpha=1.0 - self.b2)
            v.addcdiv_(s0, s1.sqrt().add_(self.eps), value=-lr_t)
            self.b1p *= self.b1
            self.b2p *= self.b2
        self.global_step += 1

    def push_apply_sync(self) -> None:
        """SyncReplicasOptimizer round: mean of the W gradients, ONE apply per shard."""
        if self.world > 1:
            out = self.recv[:self.per]
            dist.reduce_scatter_tensor(out, self.grad, op=dist.ReduceOp.AVG if self.device.type == "cuda"
                                       else dist.ReduceOp.SUM, group=self.group)
            if self.device.type != "cuda":
                out.div_(self.world)
            self._apply(out)
        else:
            self._apply(self.grad[:self.per])

    def push_apply_async(self) -> None:
        """Async PS, emulated with a collective: every owner receives each worker's slice and applies
        the W gradients one after another (W applies, W global steps)."""
        if self.world > 1:
            dist.all_to_all_single(self.recv, self.grad, group=self.group)
            for w in range(self.world):
                self._apply(self.recv[w * self.per:(w + 1) * self.per])
        else:
            self._apply(self.grad[:self.per])

    def set_grads(self, grads: Dict[str, torch.Tensor]) -> None:
        for n, (o, k) in self.offsets.items():
            self.grad[o:o + k].copy_(grads[n].reshape(-1))

    def gather_full(self) -> Dict[str, torch.Tensor]:
        return {n: t.clone() for n, t in self.pull().items()}


def _selfcheck() -> int:
    """torchrun --nproc-per-node 2 -m distributedtensorflow_b200.parallel.nccl_baseline (gloo, CPU):
    the synchronous round must equal one Adam step on the mean gradient, the async-emulated push W
    sequential Adam steps, both computed locally on the full vector."""
    import os
    backend = "nccl" if torch.cuda.is_available() else "gloo"
    dist.init_process_group(backend)
    rank, world = dist.get_rank(), dist.get_world_size()
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0"))) if backend == "nccl" else torch.device("cpu")
    shapes = [("w", (37, 5)), ("b", (5,)), ("k", (3, 3, 2, 4))]
    torch.manual_seed(0)
    init = {n: torch.randn(*s) for n, s in shapes}
    ok = True
    for mode in ("sync", "async"):
        ps = CollectivePs(shapes, dev, "adam", lr=1e-2)
        ps.init_from(init)
        ref = torch.cat([init[n].reshape(-1) for n, _ in shapes]).to(dev)
        m, v, b1p, b2p = torch.zeros_like(ref), torch.zeros_like(ref), 0.9, 0.999
        for step in range(3):
            ps.pull()
            gens = [torch.Generator().manual_seed(100 * step + w) for w in range(world)]
            all_g = [torch.randn(ps.n, generator=g).to(dev) for g in gens]
            ps.grad.zero_()
            ps.grad[:ps.n].copy_(all_g[rank])
            applied = [sum(all_g) / world] if mode == "sync" else all_g
            (ps.push_apply_sync if mode == "sync" else ps.push_apply_async)()
            for g in applied:
                lr_t = 1e-2 * math.sqrt(1 - b2p) / (1 - b1p)
                m += (g - m) * 0.1
                v += (g * g - v) * 0.001
                ref -= lr_t * m / (v.sqrt() + 1e-8)
                b1p, b2p = b1p * 0.9, b2p * 0.999
        got = torch.cat([t.reshape(-1) for t in ps.gather_full().values()])
        err = float((got - ref).abs().max())
        steps = 3 * (1 if mode == "sync" else world)
        ok &= err < 1e-5 and ps.global_step == steps
        if rank == 0:
            print("nccl_baseline selfcheck %s world=%d max_err=%.2e global_step=%d" % (mode, world, err, ps.global_step))
    dist.destroy_process_group()
    if rank == 0:
        print("SELFCHECK_OK" if ok else "SELFCHECK_FAILED")
    return 0 if ok else 1


if __name__ == "__main__":
    raise SystemExit(_selfcheck())
