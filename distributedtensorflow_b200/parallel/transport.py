"""Worker-side view of the parameter servers: the transport interface and its host-memory
implementation.

``PsClient`` is the seam SURVEY.md §5.9 asks for — ``pull`` / ``push_apply`` /
``fetch_add_global_step`` plus the control-plane calls (create / init / ready / sync-replica
objects) — so the same session, hooks and optimizers run over

* ``LocalClient``  in-process shards (threads as tasks; the fake cluster used by the tests, like
  ``tf.train.Server.create_local_server``),
* ``TcpClient``    one OS process per task talking msgpack-RPC over TCP at the config.json
  addresses (the reference's gRPC arrangement; BASELINE config 1),
* ``NvlClient``    (parallel/nvl.py) the B200 product path over NVLink peer memory, which keeps
  this control plane but moves all tensors from inside CUDA kernels,
* ``NcclClient``   (parallel/nccl_baseline.py) the NCCL comparator.

One async step = pull all variables, compute gradients on the worker, push each gradient to the
shard that owns the variable where the optimizer is applied, then ``_finish`` (beta powers) and the
``global_step`` increment (run_mnist_distributed.py:150,161 -> SURVEY §3.3).
"""
from __future__ import annotations

import dataclasses
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from ..train.optimizers import OptimizerSpec
from .layout import VarLayout, powers_shard
from .ps_core import AbortedError, PsShard

SYNC_TOKEN_QUEUE = "sync_token_q"


def spec_to_dict(spec: OptimizerSpec) -> dict:
    return dataclasses.asdict(spec)


class PsClient:
    """Abstract worker-side client. ``call(shard, method, *args)`` is the only primitive the
    host-memory transports implement."""

    num_shards: int = 1

    def call(self, shard: int, method: str, *args):
        raise NotImplementedError

    def close(self) -> None:
        pass

    # ------------------------------------------------------------------------------------------
    # control plane
    # ------------------------------------------------------------------------------------------
    def create_variables(self, layout: Dict[str, VarLayout], spec: Optional[OptimizerSpec],
                         trainable_names: Sequence[str]) -> None:
        per_shard: Dict[int, List[dict]] = {s: [] for s in range(self.num_shards)}
        for vl in layout.values():
            for sl in vl.slices:
                if sl.n > 0:
                    per_shard[sl.shard].append(dict(name=vl.name, first=sl.first, n=sl.n,
                                                    dtype=vl.dtype, trainable=vl.trainable))
        pshard = powers_shard(layout, trainable_names) if (spec and trainable_names) else 0
        for s, defs in per_shard.items():
            sd = None
            if spec is not None:
                sd = spec_to_dict(spec)
                sd["_owns_powers"] = (s == pshard)
            self.call(s, "create", defs, sd)

    def init_variables(self, layout: Dict[str, VarLayout], values: Dict[str, np.ndarray]) -> None:
        for name, arr in values.items():
            flat = np.ascontiguousarray(arr).reshape(-1)
            for sl in layout[name].slices:
                if sl.n > 0:
                    self.call(sl.shard, "assign", name, flat[sl.first:sl.first + sl.n])

    def mark_initialized(self, value: bool = True) -> None:
        for s in range(self.num_shards):
            self.call(s, "mark_initialized", value)

    def reset_optimizer_state(self) -> None:
        for s in range(self.num_shards):
            self.call(s, "reset_slots")

    def uninitialized_variables(self, layout: Dict[str, VarLayout]) -> List[str]:
        per_shard: Dict[int, List[str]] = {}
        for vl in layout.values():
            for sl in vl.slices:
                if sl.n > 0:
                    per_shard.setdefault(sl.shard, []).append(vl.name)
        missing: List[str] = []
        for s, names in per_shard.items():
            missing.extend(self.call(s, "uninitialized", names))
        return sorted(set(missing))

    def request_shutdown(self) -> None:
        for s in range(self.num_shards):
            try:
                self.call(s, "request_shutdown")
            except Exception:
                pass

    # ------------------------------------------------------------------------------------------
    # data plane (host memory)
    # ------------------------------------------------------------------------------------------
    def pull_all(self, layout: Dict[str, VarLayout], names: Sequence[str]) -> Dict[str, np.ndarray]:
        per_shard: Dict[int, List[Tuple[str, int, int]]] = {}
        for n in names:
            for sl in layout[n].slices:
                if sl.n > 0:
                    per_shard.setdefault(sl.shard, []).append((n, sl.first, sl.n))
        out: Dict[str, np.ndarray] = {}
        for n in names:
            vl = layout[n]
            np_dtype = {"float32": np.float32, "int32": np.int32, "int64": np.int64}[vl.dtype]
            out[n] = np.empty(vl.num_elements, np_dtype)
        for s, req in per_shard.items():
            arrs = self.call(s, "pull", [r[0] for r in req])
            for (n, first, cnt), a in zip(req, arrs):
                out[n][first:first + cnt] = a
        return {n: out[n].reshape(layout[n].shape) for n in names}

    def push_apply(self, layout: Dict[str, VarLayout], spec: OptimizerSpec,
                   grads: Dict[str, np.ndarray], global_step_name: Optional[str],
                   worker_id: int = 0) -> Optional[int]:
        """Scatter ``grads`` to the owning shards, apply there, then _finish and bump
        ``global_step`` (returns its new value)."""
        sd = spec_to_dict(spec)
        names = list(grads.keys())
        pshard = powers_shard(layout, names) if names else 0
        powers = self.call(pshard, "read_powers") if spec.kind == "adam" else None
        per_shard: Dict[int, Dict[str, np.ndarray]] = {}
        for n, g in grads.items():
            flat = np.ascontiguousarray(g).reshape(-1)
            for sl in layout[n].slices:
                if sl.n > 0:
                    per_shard.setdefault(sl.shard, {})[n] = flat[sl.first:sl.first + sl.n]
        for s, gs in per_shard.items():
            self.call(s, "push_apply", sd, gs, powers)
        if spec.kind == "adam":
            self.call(pshard, "finish", sd)
        return self.add_global_step(layout, global_step_name, worker_id)

    def add_global_step(self, layout, global_step_name: Optional[str], worker_id: int = 0,
                        delta: int = 1) -> Optional[int]:
        if global_step_name is None:
            return None
        gshard = layout[global_step_name].home_shard
        step = self.call(gshard, "step_add", global_step_name, delta)
        self.call(gshard, "heartbeat", worker_id, step)
        return step

    def read_global_step(self, layout: Dict[str, VarLayout], global_step_name: str) -> int:
        return self.call(layout[global_step_name].home_shard, "read_scalar", global_step_name)

    # ------------------------------------------------------------------------------------------
    # checkpoint support
    # ------------------------------------------------------------------------------------------
    def snapshot_shard(self, shard: int) -> Dict[str, np.ndarray]:
        return self.call(shard, "snapshot")


class LocalCluster:
    """An in-process fake cluster: PS shards are plain objects, tasks are threads. Registered
    under a target string so ``Server``/``MonitoredTrainingSession(master=...)`` find it."""

    _registry: Dict[str, "LocalCluster"] = {}

    def __init__(self, num_ps: int = 1, name: Optional[str] = None):
        self.shards = [PsShard(i) for i in range(max(1, num_ps))]
        self.name = name or f"local://cluster-{id(self):x}"
        LocalCluster._registry[self.name] = self
        self.alive = [True] * len(self.shards)

    @classmethod
    def lookup(cls, target: str) -> Optional["LocalCluster"]:
        return cls._registry.get(target)

    def client(self) -> "LocalClient":
        return LocalClient(self)

    def kill_ps(self, shard: int) -> None:
        """Fault injection: the shard's state is lost and calls raise AbortedError."""
        self.alive[shard] = False

    def restart_ps(self, shard: int) -> None:
        self.shards[shard] = PsShard(shard)
        self.alive[shard] = True

    def close(self) -> None:
        LocalCluster._registry.pop(self.name, None)
        for sh in self.shards:
            if hasattr(sh, "mbox_stop"):
                sh.mbox_stop()


class LocalClient(PsClient):
    def __init__(self, cluster: LocalCluster):
        self.cluster = cluster
        self.num_shards = len(cluster.shards)

    def call(self, shard: int, method: str, *args):
        if not self.cluster.alive[shard]:
            raise AbortedError(f"/job:ps/task:{shard} is unavailable")
        return getattr(self.cluster.shards[shard], method)(*args)
