"""CPU executor: evaluates any graph with PyTorch ops and autograd.

This is the oracle for the B200 kernels and the compute path of the CPU transports (BASELINE
config 1: the reference loop of run_mnist_distributed.py:142-161 restated on CPU). Per training
step it does exactly what TF's partitioned graph does: pull every variable from the
PS, run forward + backward on the worker, push every gradient back to the owning PS where the
optimizer applies it, bump ``global_step``.
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional, Sequence

import numpy as np
import torch
import torch.nn.functional as F

from . import graph as G

_TORCH_DTYPES = {"float32": torch.float32, "int32": torch.int32, "int64": torch.int64,
                 "uint8": torch.uint8, "bool": torch.bool}


def _to_torch(x, dtype: Optional[str] = None) -> torch.Tensor:
    if isinstance(x, torch.Tensor):
        t = x
    else:
        t = torch.from_numpy(np.ascontiguousarray(x))
    if dtype is not None and dtype in _TORCH_DTYPES and t.dtype != _TORCH_DTYPES[dtype]:
        t = t.to(_TORCH_DTYPES[dtype])
    return t


def evaluate(fetches: Sequence[G.Tensor], feeds: Dict[G.Tensor, Any],
             var_values: Dict[str, torch.Tensor]) -> List[torch.Tensor]:
    """Evaluates ``fetches`` (memoised recursive walk). ``var_values`` maps variable name ->
    torch tensor (possibly requiring grad)."""
    cache: Dict[int, torch.Tensor] = {}

    def ev(t: G.Tensor) -> torch.Tensor:
        k = id(t)
        if k in cache:
            return cache[k]
        if t in feeds:
            r = _to_torch(feeds[t], t.dtype if t.dtype != "none" else None)
        else:
            r = _eval_node(t, ev, var_values)
        cache[k] = r
        return r

    return [ev(f) for f in fetches]


def _eval_node(t: G.Tensor, ev, var_values) -> torch.Tensor:
    op = t.op
    if op == "VariableV2":
        return var_values[t.name]
    if op == "Placeholder":
        raise ValueError(f"You must feed a value for placeholder tensor {t.name!r}")
    if op == "IteratorGetNext":
        raise ValueError(f"Iterator output {t.name!r} was not resolved (initialise the iterator)")
    if op == "Const":
        return _to_torch(np.asarray(t.attrs["value"]), t.dtype)
    x = [ev(i) for i in t.inputs]
    if op == "Reshape":
        return x[0].reshape(t.attrs["shape"])
    if op == "Relu":
        return torch.relu(x[0])
    if op == "Softmax":
        return torch.softmax(x[0], dim=-1)
    if op == "Log":
        return torch.log(x[0])
    if op == "ClipByValue":
        return torch.clamp(x[0], t.attrs["min"], t.attrs["max"])
    if op == "Mul":
        return x[0] * x[1]
    if op == "Add":
        return x[0] + x[1]
    if op == "Neg":
        return -x[0]
    if op == "Sum":
        return x[0].sum() if t.attrs["axis"] is None else x[0].sum(t.attrs["axis"])
    if op == "Mean":
        return x[0].mean() if t.attrs["axis"] is None else x[0].mean(t.attrs["axis"])
    if op == "MatMul":
        return x[0] @ x[1]
    if op == "BiasAdd":
        return x[0] + x[1]
    if op == "Cast":
        return x[0].to(_TORCH_DTYPES[t.attrs["dtype"]])
    if op == "ArgMax":
        return x[0].argmax(t.attrs["axis"])
    if op == "Equal":
        return x[0] == x[1]
    if op == "Conv2D":
        inp = x[0].permute(0, 3, 1, 2)                     # NHWC -> NCHW
        w = x[1].permute(3, 2, 0, 1)                       # [kh,kw,ci,co] -> [co,ci,kh,kw]
        if t.attrs["padding"] == "SAME":
            kh, kw = x[1].shape[0], x[1].shape[1]
            # TF SAME with stride 1: pad (k-1)//2 before, k//2 after
            inp = F.pad(inp, ((kw - 1) // 2, kw // 2, (kh - 1) // 2, kh // 2))
        out = F.conv2d(inp, w, None, stride=t.attrs["strides"])
        return out.permute(0, 2, 3, 1)
    if op == "MaxPool":
        out = F.max_pool2d(x[0].permute(0, 3, 1, 2), t.attrs["pool_size"], t.attrs["strides"])
        return out.permute(0, 2, 3, 1)
    if op == "SparseSoftmaxXent":
        return F.cross_entropy(x[0], x[1].long(), reduction="none")
    raise NotImplementedError(f"op {op} has no CPU implementation")


class CpuStepRunner:
    """Runs ``session.run`` requests against a host-memory ``PsClient``."""

    def __init__(self, graph: G.Graph, client, layout, worker_id: int = 0):
        self.graph = graph
        self.client = client
        self.layout = layout
        self.worker_id = worker_id
        self.last_pulled: Dict[str, np.ndarray] = {}

    def _needed_variables(self, fetches: Sequence[G.Tensor]) -> List[G.Variable]:
        seen, out, stack = set(), [], list(fetches)
        while stack:
            t = stack.pop()
            if id(t) in seen:
                continue
            seen.add(id(t))
            if isinstance(t, G.Variable):
                out.append(t)
            stack.extend(t.inputs)
        return out

    def run(self, fetches: Sequence[G.Tensor], feeds: Dict[G.Tensor, Any]) -> List[Any]:
        from ..train.optimizers import TrainOp
        from ..train.sync_replicas import run_sync_train_op
        train_ops = [f for f in fetches if isinstance(f, TrainOp)]
        results: Dict[int, Any] = {}
        plain = [f for f in fetches if not isinstance(f, TrainOp)]
        if train_ops:
            top = train_ops[0]
            wanted = [top.loss] + plain
            vars_needed = self._needed_variables(wanted)
            names = [v.name for v in vars_needed]
            pulled = self.client.pull_all(self.layout, names)
            self.last_pulled = pulled
            var_values: Dict[str, torch.Tensor] = {}
            train_names = {v.name for v in top.var_list}
            for n, a in pulled.items():
                tt = torch.from_numpy(a)
                if n in train_names:
                    tt = tt.clone().requires_grad_(True)
                var_values[n] = tt
            vals = evaluate(wanted, feeds, var_values)
            loss = vals[0]
            leafs = [var_values[v.name] for v in top.var_list]
            grads = torch.autograd.grad(loss, leafs, allow_unused=True)
            gdict = {v.name: (g if g is not None else torch.zeros_like(l)).detach().numpy()
                     for v, g, l in zip(top.var_list, grads, leafs)}
            gs_name = top.global_step.name if top.global_step is not None else None
            if top.sync is not None:
                new_step = run_sync_train_op(top, self.client, self.layout, gdict, self.worker_id)
            else:
                new_step = self.client.push_apply(self.layout, top.optimizer.spec(), gdict,
                                                  gs_name, self.worker_id)
            for f, v in zip(plain, vals[1:]):
                if isinstance(f, G.Variable) and top.global_step is not None and f is top.global_step:
                    results[id(f)] = np.asarray(new_step, dtype=np.int64 if f.dtype == "int64" else np.int32)
                else:
                    results[id(f)] = _to_numpy(v)
            for t_ in train_ops:
                results[id(t_)] = None
        elif plain:
            vars_needed = self._needed_variables(plain)
            pulled = self.client.pull_all(self.layout, [v.name for v in vars_needed])
            var_values = {n: torch.from_numpy(a) for n, a in pulled.items()}
            with torch.no_grad():
                vals = evaluate(plain, feeds, var_values)
            for f, v in zip(plain, vals):
                results[id(f)] = _to_numpy(v)
        return [results[id(f)] for f in fetches]


def _to_numpy(v):
    if isinstance(v, torch.Tensor):
        v = v.detach().cpu().numpy()
        if v.shape == ():
            return v[()]
    return v
