"""``tf.train.Server``: one per task process, plus the in-process local cluster.

Parity targets: ``tf.train.Server(cluster, job_name, task_index)``, ``server.join()`` and
``server.target`` (run_mnist_distributed.py:96-103,128; templates/00_mnist_replica.py:113-118,233).
A ``ps`` server binds its own ``host:port`` from the cluster spec and serves the shard state
machine (parallel/ps_core.py) until told to shut down — and unlike the reference it CAN be shut
down cleanly: SIGINT/SIGTERM or a ``shutdown`` control message end ``join()`` (the reference's
README lists "catch Ctrl + C to terminate ParamServer" as a TODO, README.md:7). A ``worker`` server
binds its port too (liveness pings, ``existing_servers`` mode) but holds no state.

When the task's address names a GPU (``host:port/gpu:N``) the PS shard also owns a device arena
that workers map over NVLink (parallel/nvl.py); the TCP endpoint is then only the control plane.
"""
from __future__ import annotations

import signal
import threading
from typing import Dict, Optional, Union

from ..cluster import ClusterSpec, parse_address
from ..parallel.ps_core import PsShard
from ..parallel.tcp import PsRpcServer
from ..parallel.transport import LocalCluster


class Server:
    _registry: Dict[str, "Server"] = {}

    def __init__(self, server_or_cluster_def: Union[ClusterSpec, dict], job_name: Optional[str] = None,
                 task_index: Optional[int] = None, protocol: Optional[str] = None, config=None,
                 start: bool = True):
        self.cluster = server_or_cluster_def if isinstance(server_or_cluster_def, ClusterSpec) \
            else ClusterSpec(server_or_cluster_def)
        if job_name is None:
            jobs = self.cluster.jobs
            if len(jobs) != 1:
                raise ValueError("job_name must be given for a multi-job cluster")
            job_name = jobs[0]
        if task_index is None:
            if self.cluster.num_tasks(job_name) != 1:
                raise ValueError("task_index must be given for a multi-task job")
            task_index = 0
        self.job_name = job_name
        self.task_index = int(task_index)
        self.protocol = protocol or "grpc"
        self.config = config
        self.address = self.cluster.parsed_address(job_name, self.task_index)
        self.shard: Optional[PsShard] = None
        self._rpc: Optional[PsRpcServer] = None
        self._started = False
        self._installed_signals = False
        if start:
            self.start()

    # ----------------------------------------------------------------------------------------------
    def start(self) -> None:
        if self._started:
            return
        host = self.address.host
        bind_host = "0.0.0.0" if host not in ("localhost", "127.0.0.1") else host
        self.shard = PsShard(self.task_index)
        self.shard.job_name = self.job_name
        self.shard.device = self.address.device
        self._rpc = PsRpcServer(self.shard, bind_host, self.address.port)
        if self.address.port == 0:  # ephemeral port: publish the real one
            self.address = type(self.address)(self.address.host, self._rpc.port, self.address.device)
        Server._registry[self.target] = self
        self._started = True
        self._share_cores()

    def _share_cores(self) -> None:
        """Tasks of one cluster that live on the same host share its cores. The reference asks every
        task for ``op_parallelism_threads = os.cpu_count()`` (run_mnist_distributed.py:122-124) and
        TF's thread pools tolerate that; OpenMP-style pools (what the CPU backend's PyTorch ops use)
        collapse under 3 x 8 runnable threads on 8 cores (measured: 3.5 -> 8.5 global steps/s for
        1 PS + 2 workers). So each task takes its fair share unless DTF_B200_THREADS says otherwise."""
        import os
        local = {"localhost", "127.0.0.1", "::1", os.uname().nodename}
        try:
            n_local = sum(1 for job in self.cluster.jobs for i in self.cluster.task_indices(job)
                          if self.cluster.parsed_address(job, i).host in local)
        except Exception:
            n_local = 1
        want = os.environ.get("DTF_B200_THREADS")
        if want is None and n_local <= 1:
            return
        try:
            import torch
            n = int(want) if want else max(1, (os.cpu_count() or 1) // n_local)
            if torch.get_num_threads() > n or want:
                torch.set_num_threads(max(1, n))
        except Exception:
            pass

    @property
    def target(self) -> str:
        return "grpc://%s:%d" % (self.address.host, self.address.port)

    @property
    def server_def(self) -> dict:
        return {"cluster": self.cluster.as_dict(), "job_name": self.job_name,
                "task_index": self.task_index, "protocol": self.protocol}

    def join(self) -> None:
        """Blocks until shutdown is requested (Ctrl-C, SIGTERM or a control message)."""
        if threading.current_thread() is threading.main_thread() and not self._installed_signals:
            def handler(signum, frame):
                # no locks inside a signal handler (Event.set() could deadlock against the very
                # Event.wait() this thread was interrupted in): leave a flag for the join loop
                self.shard.signal_pending = True
            try:
                signal.signal(signal.SIGINT, handler)
                signal.signal(signal.SIGTERM, handler)
                self._installed_signals = True
            except ValueError:
                pass
        self._rpc.join()
        self.stop()

    def stop(self) -> None:
        if self._rpc is not None:
            self._rpc.stop()
            self._rpc = None
        Server._registry.pop(self.target, None)
        self._started = False

    shutdown = stop

    # ----------------------------------------------------------------------------------------------
    @staticmethod
    def create_local_server(config=None, start: bool = True, num_ps: int = 1) -> "LocalServer":
        return LocalServer(num_ps=num_ps)

    @classmethod
    def lookup(cls, target: str) -> Optional["Server"]:
        return cls._registry.get(target)


class LocalServer:
    """``tf.train.Server.create_local_server()``: a whole cluster inside this process."""

    def __init__(self, num_ps: int = 1):
        self.local_cluster = LocalCluster(num_ps)
        self.cluster = ClusterSpec({"ps": ["local:%d" % i for i in range(num_ps)], "worker": ["local:w0"]})
        self.job_name, self.task_index = "worker", 0

    @property
    def target(self) -> str:
        return self.local_cluster.name

    def join(self) -> None:
        threading.Event().wait()

    def stop(self) -> None:
        self.local_cluster.close()
