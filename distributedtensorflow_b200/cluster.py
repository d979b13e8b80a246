"""Cluster description: ``config.json`` loader and ``ClusterSpec``.

Parity targets (reference file:line):

* ``Config`` — run_mnist_distributed.py:14-43. Same schema (``{"ps": {label: addr}, "workers":
  {label: addr}}``), same error text for a missing key, and the same rule that the *position* of
  an entry (dict insertion order) is its task index; the ``"ps:0"`` labels are never parsed.
* ``ClusterSpec`` — tf.train.ClusterSpec as used at run_mnist_distributed.py:93 and
  templates/00_mnist_replica.py:111.

B200 extension: an address may carry the GPU that hosts the task, ``host:port/gpu:N`` (or
``/cuda:N``, ``/cpu``). The ``host:port`` part is the TCP control-plane endpoint (rendezvous,
readiness, shutdown); the device part says where the PS shard's arena or the worker's compute
lives on the 8xB200 box.
"""
from __future__ import annotations

import json
import re
from dataclasses import dataclass
from typing import Dict, List, Mapping, Optional, Sequence, Tuple, Union


class Config:
    """Loads the cluster description from a JSON file (run_mnist_distributed.py:14-43)."""

    def __init__(self, file_path: str = "config.json"):
        self.file_path = file_path
        self.json = self._load_config()
        try:
            self.ps = self.json["ps"]
            self.workers = self.json["workers"]
        except KeyError as err:
            raise AttributeError(
                'Please provide a value for "{0}" configuration key in the config.json file!'.format(
                    err.args[0]))

    def _load_config(self) -> dict:
        with open(self.file_path) as f:
            return json.load(f)

    def get_workers_with_addresses(self) -> Tuple[Tuple[str, ...], Tuple[str, ...]]:
        workers, addresses = zip(*self.workers.items())
        return workers, addresses

    def get_ps_with_addresses(self) -> Tuple[Tuple[str, ...], Tuple[str, ...]]:
        ps, addresses = zip(*self.ps.items())
        return ps, addresses

    def get_ps_and_worker_hosts(self) -> Tuple[Tuple[str, ...], Tuple[str, ...]]:
        _, ps_hosts = self.get_ps_with_addresses()
        _, worker_hosts = self.get_workers_with_addresses()
        return ps_hosts, worker_hosts

    def cluster_spec(self) -> "ClusterSpec":
        ps_hosts, worker_hosts = self.get_ps_and_worker_hosts()
        return ClusterSpec({"ps": ps_hosts, "worker": worker_hosts})


_ADDR_RE = re.compile(r"^(?P<host>[^:/]+):(?P<port>\d+)(?:/(?P<dev>(?:gpu|cuda):\d+|cpu(?::\d+)?))?$")


@dataclass(frozen=True)
class TaskAddress:
    """Parsed ``host:port[/gpu:N]`` address."""
    host: str
    port: int
    device: Optional[str]  # "cuda:N", "cpu" or None (= decide from the task index)

    @property
    def endpoint(self) -> str:
        return f"{self.host}:{self.port}"

    def __str__(self) -> str:
        return self.endpoint + (f"/{self.device}" if self.device else "")


def parse_address(addr: str) -> TaskAddress:
    m = _ADDR_RE.match(addr.strip())
    if not m:
        raise ValueError(f"Malformed task address {addr!r}; expected host:port[/gpu:N]")
    dev = m.group("dev")
    if dev is not None:
        if dev.startswith("gpu:"):
            dev = "cuda:" + dev[4:]
        elif dev.startswith("cpu"):
            dev = "cpu"
    return TaskAddress(m.group("host"), int(m.group("port")), dev)


JobTasks = Union[Sequence[str], Mapping[int, str]]


class ClusterSpec:
    """Immutable ``job name -> {task index -> address}`` map (tf.train.ClusterSpec)."""

    def __init__(self, cluster: Union["ClusterSpec", Mapping[str, JobTasks]]):
        if isinstance(cluster, ClusterSpec):
            self._jobs: Dict[str, Dict[int, str]] = {j: dict(t) for j, t in cluster._jobs.items()}
            return
        if not isinstance(cluster, Mapping):
            raise TypeError("`cluster` must be a dictionary mapping one or more job names to lists "
                            "of network addresses, or a `ClusterSpec` object")
        self._jobs = {}
        for job, tasks in cluster.items():
            if isinstance(tasks, Mapping):
                job_tasks = {int(i): str(a) for i, a in tasks.items()}
            elif isinstance(tasks, (list, tuple)):
                job_tasks = {i: str(a) for i, a in enumerate(tasks)}
            else:
                raise TypeError(f"The tasks for job {job!r} must be a list or a dictionary from "
                                "integers to strings.")
            self._jobs[str(job)] = job_tasks

    # -- tf.train.ClusterSpec API -----------------------------------------------------------
    @property
    def jobs(self) -> List[str]:
        return list(self._jobs.keys())

    def num_tasks(self, job_name: str) -> int:
        return len(self._job(job_name))

    def task_indices(self, job_name: str) -> List[int]:
        return sorted(self._job(job_name).keys())

    def task_address(self, job_name: str, task_index: int) -> str:
        job = self._job(job_name)
        try:
            return job[task_index]
        except KeyError:
            raise ValueError(f"No task with index {task_index!r} in job {job_name!r}")

    def job_tasks(self, job_name: str) -> List[Optional[str]]:
        job = self._job(job_name)
        out: List[Optional[str]] = [None] * (max(job.keys()) + 1 if job else 0)
        for i, a in job.items():
            out[i] = a
        return out

    def as_dict(self) -> Dict[str, Union[List[str], Dict[int, str]]]:
        out: Dict[str, Union[List[str], Dict[int, str]]] = {}
        for job, tasks in self._jobs.items():
            idx = sorted(tasks.keys())
            if idx == list(range(len(idx))):
                out[job] = [tasks[i] for i in idx]
            else:
                out[job] = {i: tasks[i] for i in idx}
        return out

    def as_cluster_def(self) -> bytes:
        """Serialised ``ClusterDef`` proto (job = field 1 {name = 1, tasks = 2 map<int32,string>})."""
        from .utils import protowire as pw
        out = b""
        for job in sorted(self._jobs):
            body = pw.field_bytes(1, job.encode())
            for i in sorted(self._jobs[job]):
                entry = pw.field_varint(1, i) + pw.field_bytes(2, self._jobs[job][i].encode())
                body += pw.field_bytes(2, entry)
            out += pw.field_bytes(1, body)
        return out

    # -- extensions -------------------------------------------------------------------------
    def parsed_address(self, job_name: str, task_index: int) -> TaskAddress:
        return parse_address(self.task_address(job_name, task_index))

    def _job(self, job_name: str) -> Dict[int, str]:
        try:
            return self._jobs[job_name]
        except KeyError:
            raise ValueError(f"No such job in cluster: {job_name!r}")

    def __bool__(self) -> bool:
        return bool(self._jobs)

    def __eq__(self, other) -> bool:
        return isinstance(other, ClusterSpec) and self._jobs == other._jobs

    def __ne__(self, other) -> bool:
        return not self == other

    def __repr__(self) -> str:
        return "ClusterSpec({" + ", ".join(f"{j!r}: {t!r}" for j, t in sorted(self.as_dict().items())) + "})"


def cluster_from_flags(ps_hosts: str, worker_hosts: str) -> ClusterSpec:
    """``--ps_hosts a:1,b:2 --worker_hosts c:3`` form of templates/00_mnist_replica.py:105-111."""
    def split(s: str) -> List[str]:
        return [x for x in s.split(",") if x]
    return ClusterSpec({"ps": split(ps_hosts), "worker": split(worker_hosts)})
