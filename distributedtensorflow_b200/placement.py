"""Variable placement onto PS shards: ``replica_device_setter`` and its strategies.

Parity target: ``tf.train.replica_device_setter`` as called at run_mnist_distributed.py:107,
templates/00_between_graph_replication_async_mnist.py:26 and (with ``ps_device="/job:ps/cpu:0"``)
templates/00_mnist_replica.py:134-137. Variables go to ``/job:ps/task:k`` with ``k`` chosen by a
strategy — round-robin in creation order by default, which reproduces TF's (pathological) layout
of the reference CNN: with 2 PS tasks all four kernels land on ps:0 and all four biases on ps:1
(SURVEY.md §2.5). Everything else stays on the worker. Optimizer slots and Adam's beta powers are
colocated with their variable instead of being round-robined.

B200 additions: ``GreedyLoadBalancingStrategy`` + ``byte_size_load_fn`` (TF contrib parity) and
``FlatSlicedStrategy`` — every variable above a size threshold is cut into equal row-aligned slices
across ALL shards, so one 12.8 MB tensor does not pin 98 % of the NVLink traffic on one GPU.
"""
from __future__ import annotations

import contextlib
import math
import threading
from dataclasses import dataclass
from typing import Callable, List, Optional, Sequence, Tuple, Union

from .cluster import ClusterSpec

STANDARD_PS_OPS = ("Variable", "VariableV2", "AutoReloadVariable", "MutableHashTable",
                   "MutableHashTableV2", "MutableHashTableOfTensors",
                   "MutableHashTableOfTensorsV2", "MutableDenseHashTable",
                   "MutableDenseHashTableV2", "VarHandleOp", "BoostedTreesEnsembleResourceHandleOp")


# ------------------------------------------------------------------------------------------------
# DeviceSpec ("/job:ps/task:1/device:CPU:0")
# ------------------------------------------------------------------------------------------------
@dataclass
class DeviceSpec:
    job: Optional[str] = None
    replica: Optional[int] = None
    task: Optional[int] = None
    device_type: Optional[str] = None
    device_index: Optional[int] = None

    @staticmethod
    def from_string(spec: Optional[str]) -> "DeviceSpec":
        d = DeviceSpec()
        if not spec:
            return d
        for part in spec.split("/"):
            if not part:
                continue
            kv = part.split(":")
            key = kv[0].lower()
            if key == "job" and len(kv) == 2:
                d.job = kv[1]
            elif key == "replica" and len(kv) == 2:
                d.replica = int(kv[1])
            elif key == "task" and len(kv) == 2:
                d.task = int(kv[1])
            elif key == "device" and len(kv) == 3:
                d.device_type, d.device_index = kv[1].upper(), (None if kv[2] == "*" else int(kv[2]))
            elif key in ("cpu", "gpu") and len(kv) == 2:
                d.device_type, d.device_index = key.upper(), (None if kv[1] == "*" else int(kv[1]))
            else:
                raise ValueError(f"Unknown attribute {part!r} in device spec {spec!r}")
        return d

    def merge_from(self, other: "DeviceSpec") -> "DeviceSpec":
        """Fields set in ``other`` override this spec (tf.DeviceSpec.merge_from)."""
        for f in ("job", "replica", "task", "device_type", "device_index"):
            v = getattr(other, f)
            if v is not None:
                setattr(self, f, v)
        return self

    def copy(self) -> "DeviceSpec":
        return DeviceSpec(self.job, self.replica, self.task, self.device_type, self.device_index)

    def to_string(self) -> str:
        s = ""
        if self.job is not None:
            s += f"/job:{self.job}"
        if self.replica is not None:
            s += f"/replica:{self.replica}"
        if self.task is not None:
            s += f"/task:{self.task}"
        if self.device_type is not None:
            idx = "*" if self.device_index is None else str(self.device_index)
            s += f"/device:{self.device_type}:{idx}"
        return s


# ------------------------------------------------------------------------------------------------
# "ops" as seen by a device function
# ------------------------------------------------------------------------------------------------
@dataclass
class OpInfo:
    """What a device function may inspect (the NodeDef surface TF passes to it)."""
    name: str
    type: str = "VariableV2"
    shape: Tuple[int, ...] = ()
    dtype_bytes: int = 4
    device: str = ""

    @property
    def num_elements(self) -> int:
        return int(math.prod(self.shape)) if self.shape else 1

    @property
    def byte_size(self) -> int:
        return self.num_elements * self.dtype_bytes


# ------------------------------------------------------------------------------------------------
# strategies
# ------------------------------------------------------------------------------------------------
class RoundRobinStrategy:
    """k-th placed variable -> task ``k mod num_tasks`` (TF's default ``_RoundRobinStrategy``)."""

    def __init__(self, num_tasks: int):
        self._num_tasks = num_tasks
        self._next_task = 0

    def __call__(self, unused_op: OpInfo) -> int:
        task = self._next_task
        self._next_task = (self._next_task + 1) % self._num_tasks
        return task


def byte_size_load_fn(op: OpInfo) -> int:
    """Load of a variable = its size in bytes (tf.contrib.training.byte_size_load_fn)."""
    return op.byte_size


class GreedyLoadBalancingStrategy:
    """Place each variable on the least-loaded task so far
    (tf.contrib.training.GreedyLoadBalancingStrategy)."""

    def __init__(self, num_tasks: int, load_fn: Callable[[OpInfo], int] = byte_size_load_fn):
        self._num_tasks = num_tasks
        self._load_fn = load_fn
        self._ps_loads = [0] * num_tasks

    def __call__(self, op: OpInfo) -> int:
        task = min(range(self._num_tasks), key=lambda t: (self._ps_loads[t], t))
        self._ps_loads[task] += self._load_fn(op)
        return task

    @property
    def loads(self) -> List[int]:
        return list(self._ps_loads)


SLICED = -1  # sentinel task index: the variable is sliced across all PS tasks


class FlatSlicedStrategy:
    """Variables of at least ``min_slice_bytes`` are sliced across every PS task (row-aligned,
    contiguous element ranges); smaller ones fall back to round-robin. This is the opt-in
    fast-path layout for the 8xB200 box."""

    def __init__(self, num_tasks: int, min_slice_bytes: int = 64 * 1024):
        self._num_tasks = num_tasks
        self._min = min_slice_bytes
        self._rr = RoundRobinStrategy(num_tasks)

    def __call__(self, op: OpInfo) -> int:
        if self._num_tasks > 1 and op.byte_size >= self._min and len(op.shape) >= 1:
            return SLICED
        return self._rr(op)


def slice_rows(shape: Sequence[int], num_slices: int, row_align: int = 64) -> List[Tuple[int, int]]:
    """Cut the leading dimension into ``num_slices`` contiguous [begin, end) row ranges whose
    boundaries are multiples of ``row_align`` (one TMA k-block), earlier slices first. Trailing
    slices may be empty."""
    rows = int(shape[0]) if len(shape) > 1 else 1
    if len(shape) == 1:
        # 1-D: treat elements as rows of width 1
        rows = int(shape[0])
    per = math.ceil(rows / num_slices)
    per = math.ceil(per / row_align) * row_align
    out = []
    for s in range(num_slices):
        b = min(rows, s * per)
        e = min(rows, (s + 1) * per)
        out.append((b, e))
    return out


# ------------------------------------------------------------------------------------------------
# replica_device_setter
# ------------------------------------------------------------------------------------------------
class _ReplicaDeviceChooser:
    def __init__(self, ps_tasks: int, ps_device: str, worker_device: str, merge_devices: bool,
                 ps_ops: Sequence[str], ps_strategy: Callable[[OpInfo], int]):
        self._ps_tasks = ps_tasks
        self._ps_device = ps_device
        self._worker_device = worker_device
        self._merge_devices = merge_devices
        self._ps_ops = tuple(ps_ops)
        self._ps_strategy = ps_strategy

    @property
    def ps_tasks(self) -> int:
        return self._ps_tasks

    def device_function(self, op: OpInfo) -> str:
        # If we don't return early here, either merge_devices is True, or op.device is empty
        # (in which case merging is a no-op).
        if not self._merge_devices and op.device:
            return op.device
        current = DeviceSpec.from_string(op.device or "")
        if self._ps_tasks and self._ps_device and op.type in self._ps_ops:
            ps = DeviceSpec.from_string(self._ps_device)
            cur_job, ps_job = current.job, ps.job
            if ps_job and (not cur_job or cur_job == ps_job):
                ps.task = self._ps_strategy(op)
            ps.merge_from(current)
            return ps.to_string()
        worker = DeviceSpec.from_string(self._worker_device or "")
        worker.merge_from(current)
        return worker.to_string()

    __call__ = device_function


def replica_device_setter(ps_tasks: int = 0, ps_device: str = "/job:ps",
                          worker_device: str = "/job:worker", merge_devices: bool = True,
                          cluster: Optional[Union[ClusterSpec, dict]] = None,
                          ps_ops: Optional[Sequence[str]] = None,
                          ps_strategy: Optional[Callable[[OpInfo], int]] = None):
    """Returns a device function to use in ``with device(...)`` (tf.train.replica_device_setter)."""
    if cluster is not None:
        cluster_spec = cluster if isinstance(cluster, ClusterSpec) else ClusterSpec(cluster)
        spec = cluster_spec.as_dict()
        ps_job_name = DeviceSpec.from_string(ps_device).job
        if ps_job_name not in spec or spec[ps_job_name] is None:
            return None
        ps_tasks = len(spec[ps_job_name])
    if ps_tasks == 0:
        return None
    if ps_ops is None:
        ps_ops = list(STANDARD_PS_OPS)
    if not merge_devices:
        import warnings
        warnings.warn("DEPRECATION: It is recommended to set merge_devices=true in replica_device_setter")
    if ps_strategy is None:
        ps_strategy = RoundRobinStrategy(ps_tasks)
    if not callable(ps_strategy):
        raise TypeError("ps_strategy must be callable")
    return _ReplicaDeviceChooser(ps_tasks, ps_device, worker_device, merge_devices, ps_ops, ps_strategy)


# ------------------------------------------------------------------------------------------------
# device scope stack (``with dtf.device(fn_or_string):``)
# ------------------------------------------------------------------------------------------------
_scope = threading.local()


def _stack() -> list:
    if not hasattr(_scope, "stack"):
        _scope.stack = []
    return _scope.stack


@contextlib.contextmanager
def device(device_name_or_function):
    _stack().append(device_name_or_function)
    try:
        yield
    finally:
        _stack().pop()


def apply_device_functions(op: OpInfo) -> str:
    """Runs the device stack innermost-first on ``op`` (TF's ``_apply_device_functions``)."""
    for spec in reversed(_stack()):
        if spec is None:
            continue
        if callable(spec):
            op.device = spec(op) or ""
        else:
            merged = DeviceSpec.from_string(spec)
            merged.merge_from(DeviceSpec.from_string(op.device))
            op.device = merged.to_string()
    return op.device
