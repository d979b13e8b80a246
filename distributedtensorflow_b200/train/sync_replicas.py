"""Synchronous data-parallel training: ``SyncReplicasOptimizer``.

Parity target: templates/00_mnist_replica.py:60-63,70-74,168-191,196-204,237-240 (and the
comment at run_mnist_distributed.py:143-145). Semantics of TF's
python/training/sync_replicas_optimizer.py:

* per variable a ``ConditionalAccumulator`` ON THE VARIABLE'S PS; workers ``apply_grad(g,
  local_step)`` — gradients with ``local_step`` older than the accumulator's global step are
  silently dropped (stale);
* workers then block on ``sync_token_q.dequeue()`` (queue on the ``global_step`` shard) and adopt
  the token as their new ``local_step`` — the barrier;
* the chief's queue-runner thread loops ``sync_op``: ``take_grad(replicas_to_aggregate)`` (mean)
  for every variable -> ``apply_gradients`` once -> ``global_step += 1`` ->
  ``accum.set_global_step`` -> ``enqueue_many(tokens_per_step x new global_step)`` with
  ``tokens_per_step = max(total_num_replicas, replicas_to_aggregate)``;
* ``get_init_tokens_op(num_tokens)`` primes the queue (default ``replicas_to_aggregate`` tokens;
  must be >= ``replicas_to_aggregate - total_num_replicas``);
* ``replicas_to_aggregate < total_num_replicas`` = backup workers: the slowest replicas'
  gradients of a round arrive stale and are dropped.

The mean+apply executes on the owning PS shard (``PsShard.accum_take_apply``), never through the
chief's memory. The B200 transport replaces this whole dance by ONE kernel per rank
(reduce-scatter + mean + apply + all-gather, csrc/kernels/sync_step.cu); this module is the
reference semantics, the CPU path and the fallback for backup-worker mode.
"""
from __future__ import annotations

import threading
from typing import Dict, List, Optional

import numpy as np

from ..framework import graph as G
from ..parallel.layout import powers_shard
from ..parallel.transport import SYNC_TOKEN_QUEUE, spec_to_dict
from .optimizers import Optimizer, TrainOp


class InitTokensOp:
    """Result of ``get_init_tokens_op``; ``session.run`` executes it on the chief."""

    def __init__(self, opt: "SyncReplicasOptimizer", num_tokens: int):
        self.opt = opt
        self.num_tokens = num_tokens
        self.op = "InitTokens"
        self.name = "sync_replicas/init_tokens"


class LocalStepInitOp:
    def __init__(self, opt: "SyncReplicasOptimizer", chief: bool):
        self.opt = opt
        self.chief = chief
        self.op = "LocalStepInit"
        self.name = "sync_replicas/" + ("chief_init_op" if chief else "local_step_init_op")


class ChiefQueueRunner:
    """The chief's background thread running ``sync_op`` until stopped."""

    def __init__(self, opt: "SyncReplicasOptimizer"):
        self.opt = opt
        self._thread: Optional[threading.Thread] = None
        self._stop = threading.Event()
        self._client = None
        self._layout = None
        self.rounds = 0
        self.error: Optional[BaseException] = None

    def start(self, client, layout) -> threading.Thread:
        self._client, self._layout = client, layout

        def loop():
            try:
                while not self._stop.is_set():
                    if self.opt.run_sync_op(client, layout, timeout=0.5):
                        self.rounds += 1
            except BaseException as e:  # surfaced through Coordinator / session.should_stop
                self.error = e
        self._thread = threading.Thread(target=loop, name="sync_replicas_chief_queue_runner", daemon=True)
        self._thread.start()
        return self._thread

    def stop(self, join: bool = True, close_queue: bool = True) -> None:
        """Stops the thread and (like TF's queue runner on coordinator stop) closes the token
        queue, so replicas blocked in ``dequeue`` end with OutOfRangeError instead of hanging."""
        self._stop.set()
        if join and self._thread is not None:
            self._thread.join(timeout=5.0)
        if close_queue and self._client is not None:
            try:
                gname = self.opt._train_op.global_step.name
                self._client.call(self._layout[gname].home_shard, "queue_close", SYNC_TOKEN_QUEUE)
            except Exception:
                pass


class SyncReplicasOptimizer(Optimizer):
    def __init__(self, opt: Optimizer, replicas_to_aggregate: int,
                 total_num_replicas: Optional[int] = None, variable_averages=None,
                 variables_to_average=None, use_locking: bool = False, name: str = "sync_replicas"):
        if total_num_replicas is None:
            total_num_replicas = replicas_to_aggregate
        super().__init__(opt._lr, use_locking, name)
        self._opt = opt
        self._replicas_to_aggregate = int(replicas_to_aggregate)
        self._total_num_replicas = int(total_num_replicas)
        self._tokens_per_step = max(self._total_num_replicas, self._replicas_to_aggregate)
        self._gradients_applied = False
        self._train_op: Optional[TrainOp] = None
        self._chief_queue_runner: Optional[ChiefQueueRunner] = None
        # sync_rep_local_step is a LOCAL variable: one per worker session (keyed by its client),
        # so threads sharing one graph object behave like separate worker processes
        self._local_steps: Dict[int, int] = {}
        self._setup_done = False

    kind = property(lambda self: self._opt.kind)  # type: ignore[assignment]

    def spec(self):
        return self._opt.spec()

    @property
    def replicas_to_aggregate(self) -> int:
        return self._replicas_to_aggregate

    @property
    def total_num_replicas(self) -> int:
        return self._total_num_replicas

    # -- graph construction -------------------------------------------------------------------------
    def compute_gradients(self, *args, **kwargs):
        return self._opt.compute_gradients(*args, **kwargs)

    def apply_gradients(self, grads_and_vars, global_step=None, name=None) -> TrainOp:
        if not grads_and_vars:
            raise ValueError("Must supply at least one variable")
        if global_step is None:
            raise ValueError("Global step is required to check staleness")
        top = self._opt.apply_gradients(grads_and_vars, global_step, name or self._name)
        top.sync = self
        self._train_op = top
        self._gradients_applied = True
        self._chief_queue_runner = ChiefQueueRunner(self)
        return top

    def get_chief_queue_runner(self) -> ChiefQueueRunner:
        if not self._gradients_applied:
            raise ValueError("Should be called after apply_gradients().")
        return self._chief_queue_runner

    def get_init_tokens_op(self, num_tokens: int = -1) -> InitTokensOp:
        if not self._gradients_applied:
            raise ValueError("get_init_tokens_op() should be called after apply_gradients().")
        tokens_needed = self._replicas_to_aggregate - self._total_num_replicas
        if num_tokens == -1:
            num_tokens = self._replicas_to_aggregate
        elif num_tokens < tokens_needed:
            raise ValueError("Too few tokens to finish the first step: %d (given) vs %d (needed)" %
                             (num_tokens, tokens_needed))
        return InitTokensOp(self, num_tokens)

    @property
    def chief_init_op(self) -> LocalStepInitOp:
        return LocalStepInitOp(self, True)

    @property
    def local_step_init_op(self) -> LocalStepInitOp:
        return LocalStepInitOp(self, False)

    @property
    def ready_for_local_init_op(self):
        return ReadyForLocalInitOp()

    def make_session_run_hook(self, is_chief: bool, num_tokens: int = -1):
        from .hooks import SyncReplicasOptimizerHook
        return SyncReplicasOptimizerHook(self, is_chief, num_tokens)

    # -- runtime ------------------------------------------------------------------------------------
    def _var_names(self) -> List[str]:
        return [v.name for v in self._train_op.var_list]

    def setup(self, client, layout) -> None:
        """Create the accumulators (on each variable slice's shard) and the token queue (on the
        global_step shard). Idempotent; every task may call it."""
        per_shard: Dict[int, list] = {}
        for n in self._var_names():
            for sl in layout[n].slices:
                if sl.n > 0:
                    per_shard.setdefault(sl.shard, []).append((n, sl.n))
        for s, items in per_shard.items():
            client.call(s, "accum_create", items)
        gshard = layout[self._train_op.global_step.name].home_shard
        client.call(gshard, "queue_create", SYNC_TOKEN_QUEUE)
        self._setup_done = True

    def init_local_step(self, client, layout) -> None:
        self._local_steps[id(client)] = client.read_global_step(layout, self._train_op.global_step.name)

    def get_local_step(self, client) -> int:
        return self._local_steps.get(id(client), 0)

    def init_tokens(self, client, layout, num_tokens: int) -> None:
        gname = self._train_op.global_step.name
        gshard = layout[gname].home_shard
        step = client.read_global_step(layout, gname)
        if num_tokens > 0:
            client.call(gshard, "queue_enqueue_many", SYNC_TOKEN_QUEUE, [step] * num_tokens)

    def worker_step(self, client, layout, grads: Dict[str, np.ndarray],
                    token_timeout: Optional[float] = None) -> int:
        """What a replica's train_op does: push grads into the accumulators, then wait for a token."""
        per_shard: Dict[int, Dict[str, np.ndarray]] = {}
        for n, g in grads.items():
            flat = np.ascontiguousarray(g).reshape(-1)
            for sl in layout[n].slices:
                if sl.n > 0:
                    per_shard.setdefault(sl.shard, {})[n] = flat[sl.first:sl.first + sl.n]
        for s, gs in per_shard.items():
            client.call(s, "accum_apply", gs, int(self.get_local_step(client)))
        gname = self._train_op.global_step.name
        gshard = layout[gname].home_shard
        token = client.call(gshard, "queue_dequeue", SYNC_TOKEN_QUEUE, token_timeout)
        self._local_steps[id(client)] = int(token)
        return client.read_global_step(layout, gname)

    def run_sync_op(self, client, layout, timeout: Optional[float] = None) -> bool:
        """One aggregation round, run by the chief's queue runner. Returns False on timeout
        (no full set of fresh gradients yet)."""
        spec = self._opt.spec()
        sd = spec_to_dict(spec)
        names = self._var_names()
        per_shard: Dict[int, List[str]] = {}
        for n in names:
            for sl in layout[n].slices:
                if sl.n > 0:
                    per_shard.setdefault(sl.shard, []).append(n)
        pshard = powers_shard(layout, names)
        powers = client.call(pshard, "read_powers") if spec.kind == "adam" else None
        shards = sorted(per_shard)
        done: List[int] = getattr(self, "_round_done", [])
        for s in shards:
            if s in done:
                continue
            try:
                client.call(s, "accum_take_apply", per_shard[s], self._replicas_to_aggregate, sd,
                            powers, timeout)
            except TimeoutError:
                self._round_done = done  # resume this round at the same shard next call
                return False
            done.append(s)
        self._round_done = []
        if spec.kind == "adam":
            client.call(pshard, "finish", sd)
        gname = self._train_op.global_step.name
        new_step = client.add_global_step(layout, gname, worker_id=0)
        for s in shards:
            client.call(s, "accum_set_global_step", new_step)
        gshard = layout[gname].home_shard
        client.call(gshard, "queue_enqueue_many", SYNC_TOKEN_QUEUE, [new_step] * self._tokens_per_step)
        return True


class ReadyForLocalInitOp:
    op = "ReadyForLocalInit"
    name = "sync_replicas/ready_for_local_init_op"


def run_sync_train_op(top: TrainOp, client, layout, grads: Dict[str, np.ndarray], worker_id: int) -> int:
    opt: SyncReplicasOptimizer = top.sync
    if not opt._setup_done:
        opt.setup(client, layout)
    return opt.worker_step(client, layout, grads)
