"""``SessionRunHook`` and the stock hooks of ``MonitoredTrainingSession``.

Parity targets: ``tf.train.StopAtStepHook(last_step=1000)`` (run_mnist_distributed.py:119-120,
templates/00_between_graph_replication_async_mnist.py:38) and the hooks
``MonitoredTrainingSession`` installs by itself when the chief has a ``checkpoint_dir``
(``StepCounterHook``, ``SummarySaverHook``, ``CheckpointSaverHook``). Lifecycle:
``begin -> after_create_session -> [before_run -> after_run]* -> end``. ``before_run`` may return
``SessionRunArgs`` with extra fetches; ``after_run`` sees their values and may
``run_context.request_stop()``.
"""
from __future__ import annotations

import os
import time
from typing import Any, Callable, Dict, List, Optional, Sequence

import numpy as np

from ..framework import graph as G


class SessionRunArgs:
    def __init__(self, fetches=None, feed_dict=None, options=None):
        self.fetches = fetches
        self.feed_dict = feed_dict
        self.options = options


class SessionRunContext:
    def __init__(self, original_args: SessionRunArgs, session):
        self._original_args = original_args
        self._session = session
        self._stop_requested = False

    @property
    def original_args(self) -> SessionRunArgs:
        return self._original_args

    @property
    def session(self):
        return self._session

    @property
    def stop_requested(self) -> bool:
        return self._stop_requested

    def request_stop(self) -> None:
        self._stop_requested = True


class SessionRunValues:
    def __init__(self, results, options=None, run_metadata=None):
        self.results = results
        self.options = options
        self.run_metadata = run_metadata


class SessionRunHook:
    def begin(self) -> None:
        pass

    def after_create_session(self, session, coord) -> None:
        pass

    def before_run(self, run_context: SessionRunContext) -> Optional[SessionRunArgs]:
        return None

    def after_run(self, run_context: SessionRunContext, run_values: SessionRunValues) -> None:
        pass

    def end(self, session) -> None:
        pass


def _global_step_or_raise() -> G.Variable:
    gs = G.get_global_step()
    if gs is None:
        raise RuntimeError("Global step should be created to use this hook.")
    return gs


class StopAtStepHook(SessionRunHook):
    """Requests stop at a given global step (``num_steps`` counts from the step seen at session
    creation; ``last_step`` is absolute). After the fetch says ``>= last_step`` the hook re-reads
    the step once to confirm, so with W async workers the overshoot is at most W-1 steps."""

    def __init__(self, num_steps: Optional[int] = None, last_step: Optional[int] = None):
        if num_steps is None and last_step is None:
            raise ValueError("One of num_steps or last_step must be specified.")
        if num_steps is not None and last_step is not None:
            raise ValueError("Only one of num_steps or last_step can be specified.")
        self._num_steps = num_steps
        self._last_step = last_step
        self._global_step_tensor = None

    def begin(self) -> None:
        self._global_step_tensor = _global_step_or_raise()

    def after_create_session(self, session, coord) -> None:
        if self._last_step is None:
            global_step = int(session.run(self._global_step_tensor))
            self._last_step = global_step + self._num_steps

    def before_run(self, run_context):
        return SessionRunArgs(self._global_step_tensor)

    def after_run(self, run_context, run_values):
        global_step = int(run_values.results)
        if global_step >= self._last_step:
            # the fetched value may be stale; check the latest one before stopping
            step = int(run_context.session.run(self._global_step_tensor))
            if step >= self._last_step:
                run_context.request_stop()


class SecondOrStepTimer:
    def __init__(self, every_secs: Optional[float] = None, every_steps: Optional[int] = None):
        if (every_secs is None) == (every_steps is None):
            raise ValueError("Exactly one of every_secs and every_steps should be provided.")
        self._every_secs, self._every_steps = every_secs, every_steps
        self.reset()

    def reset(self) -> None:
        self._last_triggered_step: Optional[int] = None
        self._last_triggered_time: Optional[float] = None

    def should_trigger_for_step(self, step: int) -> bool:
        if self._last_triggered_step is None:
            return True
        if self._last_triggered_step == step:
            return False
        if self._every_secs is not None and time.time() >= self._last_triggered_time + self._every_secs:
            return True
        if self._every_steps is not None and step >= self._last_triggered_step + self._every_steps:
            return True
        return False

    def update_last_triggered_step(self, step: int):
        now = time.time()
        if self._last_triggered_time is None:
            elapsed_secs, elapsed_steps = None, None
        else:
            elapsed_secs = now - self._last_triggered_time
            elapsed_steps = step - self._last_triggered_step
        self._last_triggered_time, self._last_triggered_step = now, step
        return elapsed_secs, elapsed_steps

    def last_triggered_step(self) -> Optional[int]:
        return self._last_triggered_step


class StepCounterHook(SessionRunHook):
    """Logs ``global_step/sec`` every N steps / seconds (to the python logger, a summary writer
    and ``self.last_rate``)."""

    def __init__(self, every_n_steps: Optional[int] = 100, every_n_secs: Optional[float] = None,
                 output_dir: Optional[str] = None, summary_writer=None):
        if (every_n_steps is None) == (every_n_secs is None):
            raise ValueError("exactly one of every_n_steps and every_n_secs should be provided.")
        self._timer = SecondOrStepTimer(every_secs=every_n_secs, every_steps=every_n_steps)
        self._summary_writer = summary_writer
        self._output_dir = output_dir
        self.last_rate: Optional[float] = None
        self._summary_tag = "global_step/sec"

    def begin(self) -> None:
        self._global_step_tensor = _global_step_or_raise()
        if self._summary_writer is None and self._output_dir:
            from ..utils.events import EventFileWriter
            self._summary_writer = EventFileWriter(self._output_dir)

    def before_run(self, run_context):
        return SessionRunArgs(self._global_step_tensor)

    def after_run(self, run_context, run_values):
        global_step = int(run_values.results)
        if self._timer.should_trigger_for_step(global_step):
            elapsed_time, elapsed_steps = self._timer.update_last_triggered_step(global_step)
            if elapsed_time is not None and elapsed_time > 0:
                self.last_rate = elapsed_steps / elapsed_time
                from .. import logger
                logger.info("%s: %g" % (self._summary_tag, self.last_rate))
                if self._summary_writer is not None:
                    self._summary_writer.add_scalars({self._summary_tag: self.last_rate}, global_step)

    def end(self, session) -> None:
        if self._summary_writer is not None:
            self._summary_writer.flush()


class DataPlaneStatsHook(SessionRunHook):
    """Observability for the B200 data plane ("global_step/sec, per-path achieved GB/s
    and %-of-roofline as logged keys"). Every ``every_n_steps`` LOCAL steps it puts into the
    reference-style logger (``logger.logkv``; emitted by the next ``dumpkvs``):

        local_step/sec, pull_GB/s, push_GB/s, pull_pct_nvlink, push_pct_nvlink, push_mode

    from the byte counts the executor derives from its layout (remote bytes only: what actually
    crosses NVLink). A CPU-backend session logs the step rate only. No device synchronisation and no
    extra fetch: the hook only counts ``after_run`` calls and reads a clock."""

    NVLINK_GBS = 770.0   # measured peer-copy rate per direction (profiles/pushpull_sweep_n8.json)

    def __init__(self, every_n_steps: int = 100, dump: bool = False):
        self._every = max(1, int(every_n_steps))
        self._dump = dump
        self._n = 0
        self._t0: Optional[float] = None
        self.last: Dict[str, Any] = {}

    def begin(self) -> None:
        self._n = 0
        self._t0 = None

    def after_run(self, run_context, run_values) -> None:
        now = time.time()
        if self._t0 is None:
            self._t0, self._n = now, 0
            return
        self._n += 1
        if self._n < self._every:
            return
        dt = max(now - self._t0, 1e-9)
        rate = self._n / dt
        from .. import logger
        stats: Dict[str, Any] = {"local_step/sec": rate}
        sess = run_context.session
        raw = sess._tf_sess() if hasattr(sess, "_tf_sess") else sess
        runner = getattr(raw, "_runner", None)
        traffic = getattr(runner, "traffic_per_step", None)
        if callable(traffic):
            t = traffic()
            stats["pull_GB/s"] = t["pull_remote_bytes"] * rate / 1e9
            stats["push_GB/s"] = t["push_remote_bytes"] * rate / 1e9
            stats["pull_pct_nvlink"] = 100.0 * stats["pull_GB/s"] / self.NVLINK_GBS
            stats["push_pct_nvlink"] = 100.0 * stats["push_GB/s"] / self.NVLINK_GBS
            stats["push_mode"] = t["push_mode"]
        for k, v in stats.items():
            logger.logkv(k, v)
        if self._dump:
            logger.dumpkvs()
        self.last = stats
        self._t0, self._n = now, 0


class NanLossDuringTrainingError(RuntimeError):
    def __str__(self):
        return "NaN loss during training."


class NanTensorHook(SessionRunHook):
    def __init__(self, loss_tensor, fail_on_nan_loss: bool = True):
        self._loss_tensor = loss_tensor
        self._fail_on_nan_loss = fail_on_nan_loss

    def before_run(self, run_context):
        return SessionRunArgs(self._loss_tensor)

    def after_run(self, run_context, run_values):
        if np.isnan(run_values.results):
            if self._fail_on_nan_loss:
                raise NanLossDuringTrainingError
            run_context.request_stop()


class LoggingTensorHook(SessionRunHook):
    def __init__(self, tensors, every_n_iter: Optional[int] = None, every_n_secs: Optional[float] = None,
                 at_end: bool = False, formatter: Optional[Callable[[dict], str]] = None):
        if not isinstance(tensors, dict):
            tensors = {t.name: t for t in tensors}
        if not at_end and (every_n_iter is None) == (every_n_secs is None):
            raise ValueError("either at_end and/or exactly one of every_n_iter and every_n_secs must be provided.")
        self._tensors = tensors
        self._tag_order = sorted(tensors.keys())
        self._formatter = formatter
        self._timer = (SecondOrStepTimer(every_secs=every_n_secs, every_steps=every_n_iter)
                       if (every_n_iter is not None or every_n_secs is not None) else None)
        self._log_at_end = at_end
        self.logged: List[dict] = []

    def begin(self):
        self._iter_count = 0

    def before_run(self, run_context):
        self._should_trigger = self._timer is not None and self._timer.should_trigger_for_step(self._iter_count)
        if self._should_trigger:
            return SessionRunArgs(dict(self._tensors))
        return None

    def _log(self, values: dict) -> None:
        from .. import logger
        self.logged.append(values)
        if self._formatter:
            logger.info(self._formatter(values))
        else:
            logger.info(", ".join("%s = %s" % (t, values[t]) for t in self._tag_order))

    def after_run(self, run_context, run_values):
        if self._should_trigger:
            self._log(run_values.results)
            self._timer.update_last_triggered_step(self._iter_count)
        self._iter_count += 1

    def end(self, session):
        if self._log_at_end:
            self._log(session.run(dict(self._tensors)))


class CheckpointSaverListener:
    def begin(self): pass
    def before_save(self, session, global_step_value): pass
    def after_save(self, session, global_step_value): pass
    def end(self, session, global_step_value): pass


class CheckpointSaverHook(SessionRunHook):
    """Saves TF1-format checkpoints every N seconds / steps and at ``end`` (chief only)."""

    def __init__(self, checkpoint_dir: str, save_secs: Optional[float] = None,
                 save_steps: Optional[int] = None, saver=None, checkpoint_basename: str = "model.ckpt",
                 scaffold=None, listeners: Optional[Sequence[CheckpointSaverListener]] = None):
        if saver is not None and scaffold is not None:
            raise ValueError("You cannot provide both saver and scaffold.")
        self._saver = saver
        self._scaffold = scaffold
        self._checkpoint_dir = checkpoint_dir
        self._save_path = os.path.join(checkpoint_dir, checkpoint_basename)
        self._timer = SecondOrStepTimer(every_secs=save_secs, every_steps=save_steps)
        self._listeners = list(listeners or [])
        self.saved_steps: List[int] = []

    def begin(self):
        self._global_step_tensor = _global_step_or_raise()
        for l in self._listeners:
            l.begin()

    def _get_saver(self):
        if self._saver is not None:
            return self._saver
        if self._scaffold is not None and self._scaffold.saver is not None:
            return self._scaffold.saver
        from .saver import Saver
        self._saver = Saver()
        return self._saver

    def after_create_session(self, session, coord):
        global_step = int(session.run(self._global_step_tensor))
        os.makedirs(self._checkpoint_dir, exist_ok=True)
        self._save(session, global_step)
        self._timer.update_last_triggered_step(global_step)

    def before_run(self, run_context):
        return SessionRunArgs(self._global_step_tensor)

    def after_run(self, run_context, run_values):
        stale = int(run_values.results)
        if self._timer.should_trigger_for_step(stale):
            global_step = int(run_context.session.run(self._global_step_tensor))
            if self._timer.should_trigger_for_step(global_step):
                self._timer.update_last_triggered_step(global_step)
                self._save(run_context.session, global_step)

    def end(self, session):
        last_step = int(session.run(self._global_step_tensor))
        if last_step != self._timer.last_triggered_step():
            self._save(session, last_step)
        for l in self._listeners:
            l.end(session, last_step)

    def _save(self, session, step: int) -> None:
        for l in self._listeners:
            l.before_save(session, step)
        self._get_saver().save(session, self._save_path, global_step=step)
        self.saved_steps.append(step)
        for l in self._listeners:
            l.after_save(session, step)


class SummarySaverHook(SessionRunHook):
    """Writes scalar summaries (a dict tag -> tensor, or the graph's SUMMARIES collection) every
    N steps / seconds."""

    def __init__(self, save_steps: Optional[int] = None, save_secs: Optional[float] = None,
                 output_dir: Optional[str] = None, summary_writer=None, scaffold=None,
                 summary_op: Optional[Dict[str, G.Tensor]] = None):
        self._summary_op = summary_op
        self._summary_writer = summary_writer
        self._output_dir = output_dir
        self._scaffold = scaffold
        self._timer = SecondOrStepTimer(every_secs=save_secs, every_steps=save_steps)

    def begin(self):
        self._next_step: Optional[int] = None
        self._global_step_tensor = _global_step_or_raise()
        if self._summary_writer is None and self._output_dir:
            from ..utils.events import EventFileWriter
            self._summary_writer = EventFileWriter(self._output_dir)

    def _get_summary_op(self) -> Dict[str, G.Tensor]:
        if self._summary_op is not None:
            return self._summary_op
        col = G.get_default_graph().get_collection(G.GraphKeys.SUMMARIES)
        return {tag: t for tag, t in col}

    def before_run(self, run_context):
        self._request_summary = self._next_step is None or self._timer.should_trigger_for_step(self._next_step)
        requests = {"global_step": self._global_step_tensor}
        if self._request_summary:
            ops = self._get_summary_op()
            if ops:
                requests["summary"] = dict(ops)
        return SessionRunArgs(requests)

    def after_run(self, run_context, run_values):
        if self._summary_writer is None:
            return
        stale = int(run_values.results["global_step"])
        global_step = stale
        if self._request_summary:
            self._timer.update_last_triggered_step(global_step)
            if "summary" in run_values.results:
                vals = {k: float(v) for k, v in run_values.results["summary"].items()}
                self._summary_writer.add_scalars(vals, global_step)
        self._next_step = global_step + 1

    def end(self, session=None):
        if self._summary_writer is not None:
            self._summary_writer.flush()


class FinalOpsHook(SessionRunHook):
    def __init__(self, final_ops, final_ops_feed_dict=None):
        self._final_ops = final_ops
        self._final_ops_feed_dict = final_ops_feed_dict
        self.final_ops_values = None

    def end(self, session):
        if self._final_ops is not None:
            self.final_ops_values = session.run(self._final_ops, feed_dict=self._final_ops_feed_dict)


class FeedFnHook(SessionRunHook):
    def __init__(self, feed_fn):
        self.feed_fn = feed_fn

    def before_run(self, run_context):
        return SessionRunArgs(fetches=None, feed_dict=self.feed_fn())


class GlobalStepWaiterHook(SessionRunHook):
    """Delays a worker until the global step reaches ``wait_until_step``."""

    def __init__(self, wait_until_step: int):
        self._wait_until_step = wait_until_step

    def begin(self):
        self._worker_is_started = False
        self._global_step_tensor = _global_step_or_raise()

    def before_run(self, run_context):
        if self._worker_is_started or self._wait_until_step <= 0:
            self._worker_is_started = True
            return None
        while True:
            if int(run_context.session.run(self._global_step_tensor)) >= self._wait_until_step:
                self._worker_is_started = True
                return None
            time.sleep(0.05)


class SyncReplicasOptimizerHook(SessionRunHook):
    """``SyncReplicasOptimizer.make_session_run_hook``: local-step init, token priming and the
    chief queue runner, for use with ``MonitoredTrainingSession``."""

    def __init__(self, sync_optimizer, is_chief: bool, num_tokens: int):
        self._sync_optimizer = sync_optimizer
        self._is_chief = is_chief
        self._num_tokens = num_tokens
        self._runner = None

    def begin(self):
        if not self._sync_optimizer._gradients_applied:
            raise ValueError("SyncReplicasOptimizer.apply_gradient should be called before using the hook.")

    def after_create_session(self, session, coord):
        opt = self._sync_optimizer
        session.ensure_created()
        if getattr(session._runner, "handles_sync", False):
            return   # B200 backend: the whole round is one device kernel, no host-side queues
        opt.setup(session.ps_client, session.layout)
        opt.init_local_step(session.ps_client, session.layout)
        if self._is_chief:
            session.run(opt.get_init_tokens_op(self._num_tokens))
            self._runner = opt.get_chief_queue_runner()
            self._runner.start(session.ps_client, session.layout)

    def end(self, session):
        if self._runner is not None:
            self._runner.stop()


class FaultInjectionHook(SessionRunHook):
    """``--fault_inject`` support: kill or stall this worker at a given local step
    (the reference has no fault injection)."""

    def __init__(self, at_local_step: int, action: str = "kill", stall_secs: float = 5.0):
        self._at = at_local_step
        self._action = action
        self._stall = stall_secs
        self._n = 0

    def after_run(self, run_context, run_values):
        self._n += 1
        if self._n == self._at:
            if self._action == "kill":
                os._exit(17)
            elif self._action == "stall":
                time.sleep(self._stall)
            elif self._action == "raise":
                from ..parallel.ps_core import AbortedError
                raise AbortedError("injected fault")
