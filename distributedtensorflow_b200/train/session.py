"""Sessions: ``Session``, ``SessionManager``, ``Scaffold``, ``Coordinator`` and
``MonitoredTrainingSession`` with its hooked / coordinated / recoverable layers.

Parity targets (reference file:line): ``tf.train.MonitoredTrainingSession(master=server.target,
is_chief, config, hooks)`` + ``mon_sess.run`` + ``should_stop`` (run_mnist_distributed.py:128-161,
templates/00_between_graph_replication_async_mnist.py:43-52), ``tf.ConfigProto`` (:122-124,
templates/00_mnist_replica.py:213-217). Behaviour restated from TF1's
python/training/{monitored_session,session_manager,coordinator}.py:

* chief -> ``ChiefSessionCreator``: connect; restore the latest checkpoint of ``checkpoint_dir`` if
  there is one, else run ``init_op``; check ``ready_op``; run ``local_init_op``;
* non-chief -> ``WorkerSessionCreator``: poll ``ready_op`` until the chief has initialised the
  variables (every ``recovery_wait_secs``, up to ``max_wait_secs``);
* only ``is_chief and checkpoint_dir`` adds ``StepCounterHook``, ``SummarySaverHook`` and
  ``CheckpointSaverHook``;
* wrapper stack ``_RecoverableSession`` (an ``AbortedError`` = a PS went away -> recreate the
  session: the chief re-initialises or restores, the others wait, then the step is retried)
  around ``_CoordinatedSession`` (``should_stop``) around ``_HookedSession`` (hook lifecycle).
"""
from __future__ import annotations

import os
import threading
import time
from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple

import numpy as np

from ..cluster import ClusterSpec
from ..framework import graph as G
from ..parallel.layout import VarLayout, build_layout
from ..parallel.ps_core import AbortedError, OutOfRangeError  # re-exported
from ..parallel.transport import LocalCluster, PsClient
from . import hooks as H
from .optimizers import TrainOp

UnavailableError = AbortedError


class ConfigProto:
    """The ``tf.ConfigProto`` knobs the reference touches plus B200 extensions."""

    def __init__(self, allow_soft_placement: bool = False, log_device_placement: bool = False,
                 intra_op_parallelism_threads: int = 0, inter_op_parallelism_threads: int = 0,
                 device_filters: Optional[Sequence[str]] = None, cluster_def: Optional[ClusterSpec] = None,
                 backend: str = "auto", device: Optional[str] = None, **b200_options):
        self.allow_soft_placement = allow_soft_placement
        self.log_device_placement = log_device_placement
        self.intra_op_parallelism_threads = intra_op_parallelism_threads
        self.inter_op_parallelism_threads = inter_op_parallelism_threads
        self.device_filters = list(device_filters or [])
        self.cluster_def = cluster_def
        self.backend = backend          # "auto" | "cpu" | "b200"
        self.device = device            # e.g. "cuda:0"; default: from the worker's address
        self.b200_options = b200_options


# ------------------------------------------------------------------------------------------------
# special ops
# ------------------------------------------------------------------------------------------------
class InitOp:
    op = "InitVariables"
    name = "init"


class ReadyOp:
    op = "ReportUninitialized"
    name = "report_uninitialized_variables"


class NoOp:
    op = "NoOp"
    name = "NoOp"


def global_variables_initializer() -> InitOp:
    return InitOp()


def report_uninitialized_variables() -> ReadyOp:
    return ReadyOp()


def no_op(name=None) -> NoOp:
    return NoOp()


# ------------------------------------------------------------------------------------------------
# Session
# ------------------------------------------------------------------------------------------------
def _as_list(vals):
    return vals if type(vals) is list else list(vals)


def _flatten(fetches) -> Tuple[List[Any], Callable[[List[Any]], Any]]:
    """Flattens nested lists / tuples / dicts of fetches; returns (flat, rebuild)."""
    tf_ = type(fetches)
    if tf_ is list or tf_ is tuple:
        # the hot loop's shape, sess.run([train_op, loss, global_step]): no nesting, no recursion
        for f in fetches:
            if isinstance(f, (list, tuple, dict)):
                break
        else:
            return list(fetches), (_as_list if tf_ is list else tuple)
    if isinstance(fetches, dict):
        keys = list(fetches.keys())
        parts = [_flatten(fetches[k]) for k in keys]
    elif isinstance(fetches, (list, tuple)):
        keys = None
        parts = [_flatten(f) for f in fetches]
    else:
        return [fetches], lambda vals: vals[0]
    flat: List[Any] = []
    spans = []
    for fl, _ in parts:
        spans.append((len(flat), len(flat) + len(fl)))
        flat.extend(fl)
    kind = type(fetches)

    def rebuild(vals):
        outs = [rb(vals[a:b]) for (a, b), (_, rb) in zip(spans, parts)]
        if keys is not None:
            return dict(zip(keys, outs))
        return outs if kind is list else tuple(outs)
    return flat, rebuild


class Session:
    """Connects a worker to the parameter servers named by ``target`` and executes run requests."""

    def __init__(self, target: str = "", graph: Optional[G.Graph] = None,
                 config: Optional[ConfigProto] = None):
        self.graph = graph or G.get_default_graph()
        self.config = config or ConfigProto()
        self.target = target or ""
        self._closed = False
        # task index of this worker; an in-process (local://) cluster has no addresses to derive it from,
        # so several worker threads of one process name themselves with ConfigProto(task_index=i)
        self.worker_id = int((getattr(self.config, 'b200_options', None) or {}).get('task_index', 0))
        self.cluster: Optional[ClusterSpec] = None
        self._own_local: Optional[LocalCluster] = None
        self.ps_client: PsClient = self._connect()
        self.layout: Dict[str, VarLayout] = {}
        self._runner = None
        self._created = False
        self._bound_iterators: Dict[int, Tuple[np.ndarray, ...]] = {}

    # -- connection -----------------------------------------------------------------------------
    def _connect(self) -> PsClient:
        from .server import Server
        t = self.target
        if t == "":
            self._own_local = LocalCluster(1)
            return self._own_local.client()
        lc = LocalCluster.lookup(t)
        if lc is not None:
            return lc.client()
        if t.startswith("grpc://"):
            srv = Server.lookup(t)
            cluster = self.config.cluster_def
            if srv is not None:
                cluster = srv.cluster
                if srv.job_name == "worker":
                    self.worker_id = srv.task_index
            if cluster is None:
                raise ValueError(f"No in-process server for {t!r}: pass ConfigProto(cluster_def=ClusterSpec)")
            if srv is None and "worker" in cluster.jobs:
                for i in cluster.task_indices("worker"):
                    if "grpc://" + cluster.parsed_address("worker", i).endpoint == t:
                        self.worker_id = i
            self.cluster = cluster
            from ..parallel.tcp import TcpClient
            eps = [cluster.parsed_address("ps", i).endpoint for i in cluster.task_indices("ps")]
            return TcpClient(eps)
        raise ValueError(f"Unrecognised session target {t!r}")

    @property
    def num_ps(self) -> int:
        return self.ps_client.num_shards

    # -- variable lifecycle -----------------------------------------------------------------------
    def _train_op(self) -> Optional[TrainOp]:
        for n in self.graph.nodes:
            if isinstance(n, TrainOp):
                return n
        return None

    def _optimizer_spec(self):
        top = self._train_op()
        return top.optimizer.spec() if top is not None else None

    def _trainable_names(self) -> List[str]:
        top = self._train_op()
        return [v.name for v in top.var_list] if top is not None else []

    def ensure_created(self) -> None:
        """Builds the layout and (idempotently) creates the variables on their PS shards."""
        if self._created:
            return
        self.layout = build_layout(self.graph.variables(), self.num_ps)
        self._make_runner()
        self._runner_create()
        self._created = True

    def _backend(self) -> str:
        b = self.config.backend
        if b == "auto":
            try:
                import torch
                b = "b200" if torch.cuda.is_available() else "cpu"
            except Exception:
                b = "cpu"
        return b

    def _make_runner(self) -> None:
        if self._backend() == "b200":
            from ..framework.executor_b200 import B200StepRunner
            self._runner = B200StepRunner(self.graph, self.ps_client, self.layout, self.worker_id,
                                          self.config, self.cluster)
        else:
            from ..framework.executor_cpu import CpuStepRunner
            self._runner = CpuStepRunner(self.graph, self.ps_client, self.layout, self.worker_id)

    def _runner_create(self) -> None:
        if hasattr(self._runner, "create_variables"):
            self._runner.create_variables(self._optimizer_spec(), self._trainable_names())
        else:
            self.ps_client.create_variables(self.layout, self._optimizer_spec(), self._trainable_names())

    def initialize_variables(self) -> None:
        """init_op: the chief materialises every initial value and stores it on the owning PS."""
        self.ensure_created()
        values = {v.name: v.initial_numpy() for v in self.graph.variables()}
        if hasattr(self._runner, "init_variables"):
            self._runner.init_variables(values)
        else:
            self.ps_client.init_variables(self.layout, values)
            self.ps_client.reset_optimizer_state()
            self.ps_client.mark_initialized(True)

    def uninitialized_variables(self) -> List[str]:
        self.ensure_created()
        if hasattr(self._runner, "uninitialized_variables"):
            return self._runner.uninitialized_variables()
        return self.ps_client.uninitialized_variables(self.layout)

    # -- checkpoint plumbing (used by Saver) --------------------------------------------------------
    def save_state(self, prefix: str, var_names, sharded: bool = True) -> int:
        from . import saver as S
        self.ensure_created()
        if sharded and hasattr(self._runner, "save_parallel") and \
                (getattr(self.config, "b200_options", None) or {}).get("parallel_save", True):
            # device arenas: every PS task writes its own bytes, the chief only writes the index
            return self._runner.save_parallel(prefix, set(var_names) | {"global_step"})
        if hasattr(self._runner, "snapshot_shards"):
            shards = self._runner.snapshot_shards()
        else:
            shards = [self.ps_client.snapshot_shard(s) for s in range(self.num_ps)]
        shapes = {n: vl.shape for n, vl in self.layout.items()}
        spec = self._optimizer_spec()
        slot_names = spec.slot_names() if spec is not None else []
        # reassemble sliced variables (and their slots) on the chief; whole tensors stay with
        # the shard that owns them
        per_shard: List[Dict[str, np.ndarray]] = [dict() for _ in shards]
        sliced: Dict[str, np.ndarray] = {}
        for name, vl in self.layout.items():
            keys = [name] + [f"{name}/{s}" for s in slot_names]
            if name not in var_names and name != "global_step":
                continue
            live = [sl for sl in vl.slices if sl.n > 0]
            for key in keys:
                if len(live) == 1:
                    sh = live[0].shard
                    if key in shards[sh]:
                        per_shard[sh][key] = shards[sh][key].reshape(shapes[name])
                else:
                    pieces = [shards[sl.shard].get(key) for sl in live]
                    if all(p is not None for p in pieces):
                        sliced[key] = np.concatenate([p.reshape(-1) for p in pieces]).reshape(shapes[name])
        for sh, snap in enumerate(shards):
            for key in ("beta1_power", "beta2_power"):
                if key in snap:
                    per_shard[sh][key] = np.asarray(snap[key], np.float32).reshape(())
        per_shard[0].update(sliced)
        if not sharded:
            merged: Dict[str, np.ndarray] = {}
            for d in per_shard:
                merged.update(d)
            per_shard = [merged]
        parts = S.temp_part_prefixes(prefix, len(per_shard))
        total = 0
        for part, tensors in zip(parts, per_shard):
            total += S.write_part(part, tensors)
        S.finalize_parts(parts, prefix)
        return total

    def restore_state(self, prefix: str, var_names) -> None:
        from . import saver as S
        self.ensure_created()
        tensors = S.load_checkpoint_tensors(prefix)
        spec = self._optimizer_spec()
        slot_names = spec.slot_names() if spec is not None else []
        if hasattr(self._runner, "restore_tensors"):
            self._runner.restore_tensors(tensors, var_names, slot_names)
            return
        values = {n: tensors[n] for n in self.layout if n in tensors and (n in var_names or n == "global_step")}
        missing = [n for n in var_names if n not in tensors]
        if missing:
            raise KeyError("Tensors not found in checkpoint %s: %s" % (prefix, missing))
        self.ps_client.init_variables(self.layout, values)
        for name, vl in self.layout.items():
            for i, sn in enumerate(slot_names):
                key = f"{name}/{sn}"
                if key in tensors:
                    flat = tensors[key].reshape(-1)
                    for sl in vl.slices:
                        if sl.n > 0:
                            self.ps_client.call(sl.shard, "assign_slot", name, i, flat[sl.first:sl.first + sl.n])
        if spec is not None and spec.kind == "adam" and "beta1_power" in tensors:
            from ..parallel.layout import powers_shard
            ps = powers_shard(self.layout, self._trainable_names())
            self.ps_client.call(ps, "assign_power", "beta1_power", float(tensors["beta1_power"]))
            self.ps_client.call(ps, "assign_power", "beta2_power", float(tensors["beta2_power"]))
        self.ps_client.mark_initialized(True)

    # -- run ------------------------------------------------------------------------------------------
    def run(self, fetches, feed_dict: Optional[dict] = None, options=None, run_metadata=None):
        if self._closed:
            raise RuntimeError("Attempted to use a closed Session.")
        flat, rebuild = _flatten(fetches)
        results: List[Any] = [None] * len(flat)
        graph_fetch_idx: List[int] = []
        for i, f in enumerate(flat):
            op = getattr(f, "op", None)
            if isinstance(f, G.Tensor):
                graph_fetch_idx.append(i)
            elif op == "InitVariables":
                self.initialize_variables()
            elif op == "ReportUninitialized":
                results[i] = np.asarray(self.uninitialized_variables(), dtype=object)
            elif op == "MakeIterator":
                f.iterator._bind(f.dataset)
            elif op == "InitTokens":
                self.ensure_created()
                if not getattr(self._runner, "handles_sync", False):
                    f.opt.setup(self.ps_client, self.layout)
                    f.opt.init_tokens(self.ps_client, self.layout, f.num_tokens)
            elif op == "LocalStepInit":
                self.ensure_created()
                if not getattr(self._runner, "handles_sync", False):
                    f.opt.setup(self.ps_client, self.layout)
                    f.opt.init_local_step(self.ps_client, self.layout)
            elif op == "ReadyForLocalInit":
                results[i] = np.asarray(self.uninitialized_variables(), dtype=object)
            elif op == "NoOp" or f is None:
                pass
            else:
                raise TypeError(f"Fetch argument {f!r} has invalid type {type(f)}")
        if graph_fetch_idx:
            self.ensure_created()
            feeds = dict(feed_dict or {})
            gf = [flat[i] for i in graph_fetch_idx]
            self._feed_iterators(gf, feeds)
            vals = self._runner.run(gf, feeds)
            for i, v in zip(graph_fetch_idx, vals):
                results[i] = v
        return rebuild(results)

    def _feed_iterators(self, fetches: Sequence[G.Tensor], feeds: dict) -> None:
        """One ``get_next`` per run for every iterator the fetches depend on. The walk over the
        fetches' ancestors is cached per fetch set (a node's inputs never change), so the hot loop's
        ``sess.run([train_op, loss, global_step])`` does not re-traverse the graph every step."""
        key = tuple(id(f) for f in fetches)
        cache = self.__dict__.setdefault("_iter_cache", {})
        nodes = cache.get(key)
        if nodes is None:
            seen, stack, nodes = set(), list(fetches), []
            while stack:
                t = stack.pop()
                if id(t) in seen:
                    continue
                seen.add(id(t))
                if t.op == "IteratorGetNext":
                    nodes.append(t)
                stack.extend(t.inputs)
            cache[key] = nodes
        if not nodes:
            return
        iters = {}
        for t in nodes:
            if t not in feeds:
                iters.setdefault(id(t.attrs["iterator"]), (t.attrs["iterator"], []))[1].append(t)
        for it, _ in iters.values():
            elem = it.next_element()
            for out in (it._outputs or ()):
                feeds[out] = elem[out.attrs["index"]]

    def close(self) -> None:
        if self._closed:
            return
        self._closed = True
        if self._runner is not None and hasattr(self._runner, "close"):
            self._runner.close()
        try:
            self.ps_client.close()
        except Exception:
            pass
        if self._own_local is not None:
            self._own_local.close()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


# ------------------------------------------------------------------------------------------------
# Coordinator
# ------------------------------------------------------------------------------------------------
class Coordinator:
    """tf.train.Coordinator: cooperative stop flag + exception propagation for helper threads."""

    def __init__(self, clean_stop_exception_types=(OutOfRangeError,)):
        self._clean = tuple(clean_stop_exception_types or ())
        self._stop_event = threading.Event()
        self._lock = threading.Lock()
        self._exc_info: Optional[BaseException] = None
        self._threads: List[threading.Thread] = []

    def request_stop(self, ex: Optional[BaseException] = None) -> None:
        with self._lock:
            if ex is not None and not isinstance(ex, self._clean) and self._exc_info is None:
                self._exc_info = ex
            self._stop_event.set()

    def clear_stop(self) -> None:
        with self._lock:
            self._stop_event.clear()
            self._exc_info = None

    def should_stop(self) -> bool:
        return self._stop_event.is_set()

    def wait_for_stop(self, timeout: Optional[float] = None) -> bool:
        return self._stop_event.wait(timeout)

    def register_thread(self, thread: threading.Thread) -> None:
        self._threads.append(thread)

    def raise_requested_exception(self) -> None:
        with self._lock:
            if self._exc_info is not None:
                e, self._exc_info = self._exc_info, None
                raise e

    def join(self, threads: Optional[Sequence[threading.Thread]] = None,
             stop_grace_period_secs: float = 120.0) -> None:
        for t in list(threads or self._threads):
            t.join(stop_grace_period_secs)
        self.raise_requested_exception()

    def stop_on_exception(self):
        coord = self

        class _Ctx:
            def __enter__(self_inner):
                return self_inner

            def __exit__(self_inner, et, ev, tb):
                if ev is not None:
                    coord.request_stop(ev)
                    return True
                return False
        return _Ctx()


# ------------------------------------------------------------------------------------------------
# Scaffold + SessionManager
# ------------------------------------------------------------------------------------------------
class Scaffold:
    def __init__(self, init_op=None, init_feed_dict=None, init_fn: Optional[Callable] = None,
                 ready_op=None, ready_for_local_init_op=None, local_init_op=None, summary_op=None,
                 saver=None):
        self.init_op = init_op
        self.init_feed_dict = init_feed_dict
        self.init_fn = init_fn
        self.ready_op = ready_op
        self.ready_for_local_init_op = ready_for_local_init_op
        self.local_init_op = local_init_op
        self.summary_op = summary_op
        self.saver = saver
        self._finalized = False

    def finalize(self) -> "Scaffold":
        if self._finalized:
            return self
        if self.init_op is None:
            self.init_op = global_variables_initializer()
        if self.ready_op is None:
            self.ready_op = report_uninitialized_variables()
        if self.ready_for_local_init_op is None:
            self.ready_for_local_init_op = report_uninitialized_variables()
        if self.local_init_op is None:
            self.local_init_op = no_op()
        if self.saver is None:
            from .saver import Saver
            self.saver = Saver(sharded=True)
        self._finalized = True
        return self


class SessionManager:
    """Creates sessions the chief way (init or restore) or the worker way (wait)."""

    def __init__(self, local_init_op=None, ready_op=None, ready_for_local_init_op=None,
                 graph: Optional[G.Graph] = None, recovery_wait_secs: float = 30.0):
        self._local_init_op = local_init_op
        self._ready_op = ready_op
        self._ready_for_local_init_op = ready_for_local_init_op
        self._graph = graph
        self._recovery_wait_secs = recovery_wait_secs

    def _model_ready(self, sess: Session) -> Tuple[bool, Optional[str]]:
        if self._ready_op is None:
            return True, None
        try:
            missing = sess.run(self._ready_op)
        except AbortedError as e:
            return False, "PS unavailable: %s" % e
        if len(missing) == 0:
            return True, None
        return False, "Variables not initialized: " + ", ".join(map(str, missing))

    def _try_run_local_init_op(self, sess: Session) -> Tuple[bool, Optional[str]]:
        if self._local_init_op is not None:
            if self._ready_for_local_init_op is not None:
                missing = sess.run(self._ready_for_local_init_op)
                if len(missing):
                    return False, "Variables not initialized: " + ", ".join(map(str, missing))
            sess.run(self._local_init_op)
        return True, None

    def _restore_checkpoint(self, master, saver, checkpoint_dir, checkpoint_filename_with_path,
                            config) -> Tuple[Session, bool]:
        from .saver import latest_checkpoint
        sess = Session(master, graph=self._graph, config=config)
        if checkpoint_filename_with_path:
            saver.restore(sess, checkpoint_filename_with_path)
            return sess, True
        if not saver or not checkpoint_dir:
            return sess, False
        ckpt = latest_checkpoint(checkpoint_dir)
        if not ckpt:
            return sess, False
        saver.restore(sess, ckpt)
        st_paths = []
        from .saver import get_checkpoint_state
        st = get_checkpoint_state(checkpoint_dir)
        if st:
            st_paths = st.all_model_checkpoint_paths
        saver.recover_last_checkpoints(st_paths)
        return sess, True

    def prepare_session(self, master: str, init_op=None, saver=None, checkpoint_dir=None,
                        checkpoint_filename_with_path=None, wait_for_checkpoint=False,
                        max_wait_secs=7200, config=None, init_feed_dict=None, init_fn=None) -> Session:
        sess, is_loaded = self._restore_checkpoint(master, saver, checkpoint_dir,
                                                   checkpoint_filename_with_path, config)
        if not is_loaded:
            if init_op is None and not init_fn and self._local_init_op is None:
                raise RuntimeError("Model is not initialized and no init_op or init_fn or "
                                   "local_init_op was given")
            if init_op is not None:
                sess.run(init_op, feed_dict=init_feed_dict)
            if init_fn:
                init_fn(sess)
        ok, msg = self._try_run_local_init_op(sess)
        if not ok:
            raise RuntimeError("Init operations did not make model ready for local_init. %s" % msg)
        ready, msg = self._model_ready(sess)
        if not ready:
            raise RuntimeError("Init operations did not make model ready. Init op: %s, error: %s"
                               % (getattr(init_op, "name", init_op), msg))
        return sess

    def recover_session(self, master: str, saver=None, checkpoint_dir=None,
                        checkpoint_filename_with_path=None, wait_for_checkpoint=False,
                        max_wait_secs=7200, config=None) -> Tuple[Session, bool]:
        sess, is_loaded = self._restore_checkpoint(master, saver, checkpoint_dir,
                                                   checkpoint_filename_with_path, config)
        ok, _ = self._try_run_local_init_op(sess)
        if not ok:
            return sess, False
        ready, _ = self._model_ready(sess)
        return sess, (is_loaded and ready)

    def wait_for_session(self, master: str, config=None, max_wait_secs: float = float("inf")) -> Session:
        deadline = time.time() + max_wait_secs
        while True:
            sess = Session(master, graph=self._graph, config=config)
            not_ready_msg = None
            try:
                ok, not_ready_msg = self._try_run_local_init_op(sess)
                if ok:
                    ready, not_ready_msg = self._model_ready(sess)
                    if ready:
                        return sess
            except AbortedError as e:
                not_ready_msg = str(e)
            sess.close()
            remaining = deadline - time.time()
            if remaining <= 0:
                raise TimeoutError("Session was not ready after waiting %s secs: %s"
                                   % (max_wait_secs, not_ready_msg))
            time.sleep(min(self._recovery_wait_secs, max(remaining, 0.0)))


# ------------------------------------------------------------------------------------------------
# session creators
# ------------------------------------------------------------------------------------------------
class SessionCreator:
    def create_session(self) -> Session:
        raise NotImplementedError


class ChiefSessionCreator(SessionCreator):
    def __init__(self, scaffold: Optional[Scaffold] = None, master: str = "", config=None,
                 checkpoint_dir: Optional[str] = None, checkpoint_filename_with_path: Optional[str] = None,
                 recovery_wait_secs: float = 30.0):
        self._checkpoint_dir = checkpoint_dir
        self._checkpoint_filename_with_path = checkpoint_filename_with_path
        self._scaffold = scaffold or Scaffold()
        self._master = master
        self._config = config
        self._recovery_wait_secs = recovery_wait_secs

    def create_session(self) -> Session:
        sc = self._scaffold.finalize()
        sm = SessionManager(local_init_op=sc.local_init_op, ready_op=sc.ready_op,
                            ready_for_local_init_op=sc.ready_for_local_init_op,
                            recovery_wait_secs=self._recovery_wait_secs)
        return sm.prepare_session(self._master, saver=sc.saver, checkpoint_dir=self._checkpoint_dir,
                                  checkpoint_filename_with_path=self._checkpoint_filename_with_path,
                                  config=self._config, init_op=sc.init_op,
                                  init_feed_dict=sc.init_feed_dict, init_fn=sc.init_fn)


class WorkerSessionCreator(SessionCreator):
    def __init__(self, scaffold: Optional[Scaffold] = None, master: str = "", config=None,
                 max_wait_secs: float = 30 * 60, recovery_wait_secs: float = 30.0):
        self._scaffold = scaffold or Scaffold()
        self._master = master
        self._config = config
        self._max_wait_secs = max_wait_secs
        self._recovery_wait_secs = recovery_wait_secs

    def create_session(self) -> Session:
        sc = self._scaffold.finalize()
        sm = SessionManager(local_init_op=sc.local_init_op, ready_op=sc.ready_op,
                            ready_for_local_init_op=sc.ready_for_local_init_op,
                            recovery_wait_secs=self._recovery_wait_secs)
        return sm.wait_for_session(self._master, config=self._config, max_wait_secs=self._max_wait_secs)


# ------------------------------------------------------------------------------------------------
# wrapped sessions
# ------------------------------------------------------------------------------------------------
class _WrappedSession:
    def __init__(self, sess):
        self._sess = sess
        self._wrapped_is_stoppable = isinstance(sess, _WrappedSession)

    @property
    def graph(self):
        return self._sess.graph

    @property
    def ps_client(self):
        return self._sess.ps_client

    @property
    def layout(self):
        return self._sess.layout

    def __getattr__(self, name):
        return getattr(self._sess, name)

    def should_stop(self) -> bool:
        if self._check_stop():
            return True
        if self._sess is not None:
            return self._wrapped_is_stoppable and self._sess.should_stop()
        return True

    def _check_stop(self) -> bool:
        return False

    def close(self) -> None:
        if self._sess is not None:
            try:
                self._sess.close()
            finally:
                self._sess = None

    def run(self, *args, **kwargs):
        return self._sess.run(*args, **kwargs)


class _HookedSession(_WrappedSession):
    def __init__(self, sess, hooks: Sequence[H.SessionRunHook]):
        super().__init__(sess)
        self._hooks = list(hooks)
        self._should_stop = False

    def _check_stop(self) -> bool:
        return self._should_stop

    def run(self, fetches, feed_dict=None, options=None, run_metadata=None):
        if self.should_stop():
            raise RuntimeError("Run called even after should_stop requested.")
        actual = {"caller": fetches}
        run_context = H.SessionRunContext(H.SessionRunArgs(fetches, feed_dict), session=self._sess)
        feed = dict(feed_dict) if feed_dict else {}
        for hook in self._hooks:
            req = hook.before_run(run_context)
            if req is not None:
                if req.fetches is not None:
                    actual[id(hook)] = req.fetches
                if req.feed_dict:
                    for k in req.feed_dict:
                        if k in feed:
                            raise RuntimeError("Same tensor is fed by a SessionRunHook and user.")
                    feed.update(req.feed_dict)
        outputs = self._sess.run(actual, feed_dict=feed or None)
        for hook in self._hooks:
            hook.after_run(run_context, H.SessionRunValues(results=outputs.get(id(hook))))
        self._should_stop = self._should_stop or run_context.stop_requested
        return outputs["caller"]


class _CoordinatedSession(_WrappedSession):
    def __init__(self, sess, coord: Coordinator, stop_grace_period_secs: float = 120):
        super().__init__(sess)
        self._coord = coord
        self._stop_grace_period_secs = stop_grace_period_secs

    def _check_stop(self) -> bool:
        self._coord.raise_requested_exception()
        return self._coord.should_stop()

    def close(self) -> None:
        self._coord.request_stop()
        try:
            self._coord.join(stop_grace_period_secs=self._stop_grace_period_secs)
        finally:
            try:
                super().close()
            except Exception:
                pass

    def run(self, *args, **kwargs):
        try:
            return self._sess.run(*args, **kwargs)
        except (AbortedError,):
            raise
        except Exception:
            # prefer an exception raised by a helper thread (e.g. the sync queue runner)
            try:
                self._coord.raise_requested_exception()
            except Exception:
                raise
            raise


class _RecoverableSession(_WrappedSession):
    """Recreates the session when a PS/worker pre-emption surfaces as AbortedError
    (run_mnist_distributed.py:146 "mon_sess.run handles AbortedError in case of preempted PS")."""

    def __init__(self, sess_creator: SessionCreator):
        self._sess_creator = sess_creator
        self.recoveries = 0
        super().__init__(self._create_session())

    def _create_session(self):
        while True:
            try:
                return self._sess_creator.create_session()
            except AbortedError:
                from .. import logger
                logger.info("An error was raised while a session was being created. This may be "
                            "due to a preemption of a connected worker or parameter server. A new "
                            "session will be created.")
                time.sleep(0.2)

    def _check_stop(self) -> bool:
        try:
            if self._sess:
                return self._sess._check_stop()
            return True
        except AbortedError:
            self.close()
            self._sess = self._create_session()
            return False

    def run(self, fetches, feed_dict=None, options=None, run_metadata=None):
        while True:
            try:
                if not self._sess:
                    self._sess = self._create_session()
                return self._sess.run(fetches, feed_dict=feed_dict, options=options,
                                      run_metadata=run_metadata)
            except AbortedError:
                from .. import logger
                logger.info("An error was raised. This may be due to a preemption in a connected "
                            "worker or parameter server. The current session will be closed and a "
                            "new session will be created.")
                self.recoveries += 1
                self.close()
                self._sess = None


# ------------------------------------------------------------------------------------------------
# MonitoredSession
# ------------------------------------------------------------------------------------------------
class _MonitoredSession:
    def __init__(self, session_creator: Optional[SessionCreator], hooks: Optional[Sequence[H.SessionRunHook]],
                 should_recover: bool, stop_grace_period_secs: float = 120):
        self._graph_was_finalized = G.get_default_graph().finalized
        self._hooks = list(hooks or [])
        for h in self._hooks:
            h.begin()
        self._coordinated_creator = self._CoordinatedSessionCreator(
            session_creator=session_creator or ChiefSessionCreator(), hooks=self._hooks,
            stop_grace_period_secs=stop_grace_period_secs)
        if should_recover:
            self._sess = _RecoverableSession(self._coordinated_creator)
        else:
            self._sess = self._coordinated_creator.create_session()

    @property
    def graph(self):
        return self._coordinated_creator.tf_sess.graph if self._coordinated_creator.tf_sess else None

    class _CoordinatedSessionCreator(SessionCreator):
        def __init__(self, session_creator, hooks, stop_grace_period_secs):
            self._session_creator = session_creator
            self._hooks = hooks
            self.coord: Optional[Coordinator] = None
            self.tf_sess: Optional[Session] = None
            self._stop_grace_period_secs = stop_grace_period_secs

        def create_session(self):
            self.tf_sess = self._session_creator.create_session()
            self.coord = Coordinator(clean_stop_exception_types=[])
            for hook in self._hooks:
                hook.after_create_session(self.tf_sess, self.coord)
            return _CoordinatedSession(_HookedSession(self.tf_sess, self._hooks), self.coord,
                                       self._stop_grace_period_secs)

    def run(self, fetches, feed_dict=None, options=None, run_metadata=None):
        return self._sess.run(fetches, feed_dict=feed_dict, options=options, run_metadata=run_metadata)

    def run_step_fn(self, step_fn):
        class StepContext:
            def __init__(ctx, session, run_with_hooks_fn):
                ctx.session = session
                ctx.run_with_hooks = run_with_hooks_fn
                ctx._stop = False

            def request_stop(ctx):
                ctx._stop = True
        ctx = StepContext(self._coordinated_creator.tf_sess, self.run)
        out = step_fn(ctx)
        if ctx._stop:
            self._coordinated_creator.coord.request_stop()
        return out

    def should_stop(self) -> bool:
        return self._sess is None or self._sess.should_stop()

    def close(self) -> None:
        self._close_internal()

    def __enter__(self):
        return self

    def __exit__(self, exception_type, exception_value, traceback):
        if exception_type in (IndexError, OutOfRangeError, StopIteration):
            exception_type = None
        self._close_internal(exception_type)
        return exception_type is None

    def _close_internal(self, exception_type=None) -> None:
        try:
            if not exception_type:
                for h in self._hooks:
                    h.end(self._coordinated_creator.tf_sess)
        finally:
            try:
                if self._sess is not None:
                    self._sess.close()
            finally:
                self._sess = None
                self._coordinated_creator.tf_sess = None
                self._coordinated_creator.coord = None

    def _is_closed(self) -> bool:
        return self._coordinated_creator.tf_sess is None

    def _tf_sess(self) -> Optional[Session]:
        return self._coordinated_creator.tf_sess

    @property
    def recoveries(self) -> int:
        return getattr(self._sess, "recoveries", 0)


class MonitoredSession(_MonitoredSession):
    def __init__(self, session_creator=None, hooks=None, stop_grace_period_secs=120):
        super().__init__(session_creator, hooks, should_recover=True,
                         stop_grace_period_secs=stop_grace_period_secs)


class SingularMonitoredSession(_MonitoredSession):
    def __init__(self, hooks=None, scaffold=None, master="", config=None, checkpoint_dir=None,
                 stop_grace_period_secs=120, checkpoint_filename_with_path=None):
        creator = ChiefSessionCreator(scaffold=scaffold, master=master, config=config,
                                      checkpoint_dir=checkpoint_dir,
                                      checkpoint_filename_with_path=checkpoint_filename_with_path)
        super().__init__(creator, hooks, should_recover=False,
                         stop_grace_period_secs=stop_grace_period_secs)

    def raw_session(self):
        return self._tf_sess()


USE_DEFAULT = object()


def MonitoredTrainingSession(master: str = "", is_chief: bool = True, checkpoint_dir: Optional[str] = None,
                             scaffold: Optional[Scaffold] = None, hooks=None, chief_only_hooks=None,
                             save_checkpoint_secs=USE_DEFAULT, save_summaries_steps=USE_DEFAULT,
                             save_summaries_secs=USE_DEFAULT, config=None, stop_grace_period_secs=120,
                             log_step_count_steps=100, max_wait_secs=7200, save_checkpoint_steps=USE_DEFAULT,
                             summary_dir=None, recovery_wait_secs: float = 30.0) -> MonitoredSession:
    """Creates a ``MonitoredSession`` for training (run_mnist_distributed.py:128-132)."""
    if save_summaries_steps == USE_DEFAULT and save_summaries_secs == USE_DEFAULT:
        save_summaries_steps, save_summaries_secs = 100, None
    elif save_summaries_secs == USE_DEFAULT:
        save_summaries_secs = None
    elif save_summaries_steps == USE_DEFAULT:
        save_summaries_steps = None
    if save_checkpoint_steps == USE_DEFAULT and save_checkpoint_secs == USE_DEFAULT:
        save_checkpoint_steps, save_checkpoint_secs = None, 600
    elif save_checkpoint_secs == USE_DEFAULT:
        save_checkpoint_secs = None
    elif save_checkpoint_steps == USE_DEFAULT:
        save_checkpoint_steps = None

    scaffold = scaffold or Scaffold()
    if not is_chief:
        creator = WorkerSessionCreator(scaffold=scaffold, master=master, config=config,
                                       max_wait_secs=max_wait_secs, recovery_wait_secs=recovery_wait_secs)
        return MonitoredSession(session_creator=creator, hooks=hooks or [],
                                stop_grace_period_secs=stop_grace_period_secs)
    all_hooks: List[H.SessionRunHook] = []
    if chief_only_hooks:
        all_hooks.extend(chief_only_hooks)
    creator = ChiefSessionCreator(scaffold=scaffold, checkpoint_dir=checkpoint_dir, master=master,
                                  config=config, recovery_wait_secs=recovery_wait_secs)
    summary_dir = summary_dir or checkpoint_dir
    if summary_dir:
        if log_step_count_steps and log_step_count_steps > 0:
            all_hooks.append(H.StepCounterHook(output_dir=summary_dir, every_n_steps=log_step_count_steps))
        if (save_summaries_steps and save_summaries_steps > 0) or (save_summaries_secs and save_summaries_secs > 0):
            all_hooks.append(H.SummarySaverHook(scaffold=scaffold, save_steps=save_summaries_steps,
                                                save_secs=save_summaries_secs, output_dir=summary_dir))
    if checkpoint_dir:
        if (save_checkpoint_secs and save_checkpoint_secs > 0) or (save_checkpoint_steps and save_checkpoint_steps > 0):
            all_hooks.append(H.CheckpointSaverHook(checkpoint_dir, save_steps=save_checkpoint_steps,
                                                   save_secs=save_checkpoint_secs, scaffold=scaffold))
    if hooks:
        all_hooks.extend(hooks)
    return MonitoredSession(session_creator=creator, hooks=all_hooks,
                            stop_grace_period_secs=stop_grace_period_secs)
