"""``tf.train.Supervisor`` — the (deprecated in TF, but used by the reference) training helper of
templates/00_mnist_replica.py:193-240: ``Supervisor(is_chief, logdir, init_op, local_init_op,
ready_for_local_init_op, recovery_wait_secs=1, global_step)``,
``sv.prepare_or_wait_for_session(master, config)``, ``sv.start_queue_runners(sess,
[chief_queue_runner])``, ``sv.should_stop()``, ``sv.stop()``.

Chief: prepares the session (restore from ``logdir`` or run ``init_op``) and starts the background
services — checkpoint saver every ``save_model_secs`` (600) to ``logdir/model.ckpt``, a step
counter every ``save_summaries_secs`` (120) writing ``global_step/sec`` to the event file.
Non-chief: waits (polling every ``recovery_wait_secs``) until the chief initialised the variables.
"""
from __future__ import annotations

import os
import threading
import time
from typing import List, Optional, Sequence

from ..framework import graph as G
from .saver import Saver
from .session import (Coordinator, Scaffold, Session, SessionManager, global_variables_initializer,
                      report_uninitialized_variables)

USE_DEFAULT = 0


class Supervisor:
    def __init__(self, graph: Optional[G.Graph] = None, ready_op=USE_DEFAULT,
                 ready_for_local_init_op=USE_DEFAULT, is_chief: bool = True, init_op=USE_DEFAULT,
                 init_feed_dict=None, local_init_op=USE_DEFAULT, logdir: Optional[str] = None,
                 summary_op=USE_DEFAULT, saver=USE_DEFAULT, global_step=USE_DEFAULT,
                 save_summaries_secs: float = 120, save_model_secs: float = 600,
                 recovery_wait_secs: float = 30, stop_grace_secs: float = 120,
                 checkpoint_basename: str = "model.ckpt", session_manager=None, summary_writer=USE_DEFAULT,
                 init_fn=None, local_init_run_options=None):
        self._graph = graph or G.get_default_graph()
        self._is_chief = is_chief
        self._coord = Coordinator()
        self._recovery_wait_secs = recovery_wait_secs
        self._stop_grace_secs = stop_grace_secs
        self._init_fn = init_fn
        self._logdir = logdir if is_chief else None
        self._save_summaries_secs = save_summaries_secs if is_chief else None
        self._save_model_secs = save_model_secs if is_chief else None
        self._save_path = os.path.join(self._logdir, checkpoint_basename) if self._logdir else None
        self._ready_op = report_uninitialized_variables() if ready_op is USE_DEFAULT else ready_op
        self._ready_for_local_init_op = None if ready_for_local_init_op is USE_DEFAULT else ready_for_local_init_op
        self._init_op = global_variables_initializer() if init_op is USE_DEFAULT else init_op
        self._init_feed_dict = init_feed_dict
        self._local_init_op = None if local_init_op is USE_DEFAULT else local_init_op
        self._saver = Saver() if saver is USE_DEFAULT else saver
        self._global_step = G.get_global_step(self._graph) if global_step is USE_DEFAULT else global_step
        self._summary_writer = None
        if self._logdir and summary_writer is USE_DEFAULT:
            from ..utils.events import EventFileWriter
            self._summary_writer = EventFileWriter(self._logdir)
        self._session_manager = session_manager or SessionManager(
            local_init_op=self._local_init_op, ready_op=self._ready_op,
            ready_for_local_init_op=self._ready_for_local_init_op, graph=self._graph,
            recovery_wait_secs=self._recovery_wait_secs)
        self._threads: List[threading.Thread] = []
        self._runners = []
        self.last_global_step_per_sec: Optional[float] = None

    # -- properties ---------------------------------------------------------------------------------
    is_chief = property(lambda self: self._is_chief)
    coord = property(lambda self: self._coord)
    saver = property(lambda self: self._saver)
    global_step = property(lambda self: self._global_step)
    save_path = property(lambda self: self._save_path)
    session_manager = property(lambda self: self._session_manager)
    summary_writer = property(lambda self: self._summary_writer)
    init_op = property(lambda self: self._init_op)
    ready_op = property(lambda self: self._ready_op)

    # -- session ------------------------------------------------------------------------------------
    def prepare_or_wait_for_session(self, master: str = "", config=None, wait_for_checkpoint=False,
                                    max_wait_secs: float = 7200, start_standard_services: bool = True) -> Session:
        self._coord.clear_stop()
        if self._is_chief:
            sess = self._session_manager.prepare_session(
                master, init_op=self._init_op, saver=self._saver, checkpoint_dir=self._logdir,
                wait_for_checkpoint=wait_for_checkpoint, max_wait_secs=max_wait_secs, config=config,
                init_feed_dict=self._init_feed_dict, init_fn=self._init_fn)
            if start_standard_services:
                self.start_standard_services(sess)
        else:
            sess = self._session_manager.wait_for_session(master, config=config, max_wait_secs=max_wait_secs)
        return sess

    def start_standard_services(self, sess: Session) -> List[threading.Thread]:
        if not self._is_chief:
            raise RuntimeError("Only chief supervisor can start standard services. "
                               "Because only chief supervisors can write events.")
        if not self._logdir:
            return []
        threads = []
        if self._save_model_secs and self._saver is not None:
            threads.append(self._looper(self._save_model_secs, lambda: self._save(sess), "sv_saver"))
        if self._save_summaries_secs and self._global_step is not None:
            state = {"t": time.time(), "s": None}

            def count():
                step = int(sess.run(self._global_step))
                now = time.time()
                if state["s"] is not None and now > state["t"]:
                    self.last_global_step_per_sec = (step - state["s"]) / (now - state["t"])
                    if self._summary_writer is not None:
                        self._summary_writer.add_scalars({"global_step/sec": self.last_global_step_per_sec}, step)
                state["t"], state["s"] = now, step
            threads.append(self._looper(self._save_summaries_secs, count, "sv_step_counter"))
        self._threads.extend(threads)
        return threads

    def _save(self, sess: Session) -> None:
        self._saver.save(sess, self._save_path, global_step=self._global_step)

    def _looper(self, period: float, fn, name: str) -> threading.Thread:
        coord = self._coord

        def loop():
            while not coord.wait_for_stop(period):
                try:
                    fn()
                except Exception as e:  # propagate through the coordinator like TF's LooperThread
                    coord.request_stop(e)
                    return
        t = threading.Thread(target=loop, name=name, daemon=True)
        coord.register_thread(t)
        t.start()
        return t

    def start_queue_runners(self, sess: Session, queue_runners: Optional[Sequence] = None) -> List[threading.Thread]:
        threads = []
        for qr in (queue_runners or []):
            sess.ensure_created()
            if getattr(sess._runner, "handles_sync", False):
                continue   # B200 backend: no chief-side aggregation thread is needed
            if hasattr(qr.opt, "setup"):
                qr.opt.setup(sess.ps_client, sess.layout)
            t = qr.start(sess.ps_client, sess.layout)
            self._runners.append(qr)
            self._coord.register_thread(t)
            threads.append(t)
        self._threads.extend(threads)
        return threads

    # -- stop ---------------------------------------------------------------------------------------
    def should_stop(self) -> bool:
        return self._coord.should_stop()

    def request_stop(self, ex=None) -> None:
        self._coord.request_stop(ex)

    def stop(self, threads=None, close_summary_writer: bool = True, ignore_live_threads: bool = False) -> None:
        self._coord.request_stop()
        for qr in self._runners:
            qr.stop(join=True)
        try:
            self._coord.join(threads or self._threads, stop_grace_period_secs=self._stop_grace_secs)
        finally:
            if close_summary_writer and self._summary_writer is not None:
                self._summary_writer.flush()
                self._summary_writer.close()
                self._summary_writer = None

    def wait_for_stop(self) -> None:
        self._coord.wait_for_stop()

    def managed_session(self, master: str = "", config=None, start_standard_services: bool = True,
                        close_summary_writer: bool = True):
        sv = self

        class _Managed:
            def __enter__(self_inner):
                self_inner.sess = sv.prepare_or_wait_for_session(
                    master, config=config, start_standard_services=start_standard_services)
                return self_inner.sess

            def __exit__(self_inner, et, ev, tb):
                if ev is not None:
                    sv.request_stop(ev)
                try:
                    sv.stop(close_summary_writer=close_summary_writer)
                finally:
                    self_inner.sess.close()
                return False
        return _Managed()
