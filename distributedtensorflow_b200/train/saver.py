"""``tf.train.Saver`` (V2 / tensor-bundle format) and checkpoint-state management.

Parity targets: the checkpointing ``MonitoredTrainingSession(checkpoint_dir="/tmp/train_logs")``
performs in templates/00_between_graph_replication_async_mnist.py:43-45 (disabled by a comment in
run_mnist_distributed.py:130) and ``tf.train.Supervisor(logdir=...)`` in
templates/00_mnist_replica.py:196-211. Files written are TF1-readable:

* ``model.ckpt-N.index`` / ``model.ckpt-N.data-0000i-of-0000M`` — sharded like TF's default
  ``Saver(sharded=True)``: each PS shard's variables (+ their slots, + beta powers) go to that
  shard's data file, written to ``model.ckpt-N_temp_<uuid>/part-0000i-of-0000M`` first and then
  merged (``MergeV2Checkpoints``) by the chief;
* ``checkpoint`` — text proto ``CheckpointState`` with ``model_checkpoint_path`` and
  ``all_model_checkpoint_paths``; ``max_to_keep`` (default 5) old checkpoints are deleted.

Tensor names follow TF: variable names from the graph (``conv2d/kernel`` ...), ``global_step``,
``<var>/Adam``, ``<var>/Adam_1`` (``/Adagrad``, ``/Momentum``), ``beta1_power``, ``beta2_power``.
Restore re-shards: an M-shard checkpoint can be loaded into any number of PS tasks.
"""
from __future__ import annotations

import glob
import os
import re
import shutil
import time
import uuid
from typing import Dict, List, Optional, Sequence

import numpy as np

from ..ckpt.bundle import BundleReader, BundleWriter, merge_bundles
from ..framework import graph as G

_STATE_FILE = "checkpoint"


# ------------------------------------------------------------------------------------------------
# CheckpointState text proto
# ------------------------------------------------------------------------------------------------
class CheckpointState:
    def __init__(self, model_checkpoint_path: str = "", all_model_checkpoint_paths: Optional[List[str]] = None):
        self.model_checkpoint_path = model_checkpoint_path
        self.all_model_checkpoint_paths = list(all_model_checkpoint_paths or [])


def _escape(s: str) -> str:
    return s.replace("\\", "\\\\").replace('"', '\\"')


def _unescape(s: str) -> str:
    return s.replace('\\"', '"').replace("\\\\", "\\")


def update_checkpoint_state(save_dir: str, model_checkpoint_path: str,
                            all_model_checkpoint_paths: Optional[Sequence[str]] = None,
                            latest_filename: Optional[str] = None) -> None:
    paths = list(all_model_checkpoint_paths or [])
    if not paths or paths[-1] != model_checkpoint_path:
        paths.append(model_checkpoint_path)

    def rel(p: str) -> str:  # TF stores paths relative to the state file's directory
        if os.path.isabs(p) and os.path.dirname(p) == os.path.abspath(save_dir):
            return os.path.basename(p)
        return p

    lines = ['model_checkpoint_path: "%s"' % _escape(rel(model_checkpoint_path))]
    lines += ['all_model_checkpoint_paths: "%s"' % _escape(rel(p)) for p in paths]
    fname = os.path.join(save_dir, latest_filename or _STATE_FILE)
    tmp = fname + ".tmp%s" % uuid.uuid4().hex
    with open(tmp, "w") as f:
        f.write("\n".join(lines) + "\n")
    os.replace(tmp, fname)


def get_checkpoint_state(checkpoint_dir: str, latest_filename: Optional[str] = None) -> Optional[CheckpointState]:
    fname = os.path.join(checkpoint_dir, latest_filename or _STATE_FILE)
    if not os.path.exists(fname):
        return None
    st = CheckpointState()
    with open(fname) as f:
        for line in f:
            m = re.match(r'\s*(\w+)\s*:\s*"(.*)"\s*$', line)
            if not m:
                continue
            val = _unescape(m.group(2))
            if not os.path.isabs(val):
                val = os.path.join(checkpoint_dir, val)
            if m.group(1) == "model_checkpoint_path":
                st.model_checkpoint_path = val
            elif m.group(1) == "all_model_checkpoint_paths":
                st.all_model_checkpoint_paths.append(val)
    return st


def checkpoint_exists(prefix: str) -> bool:
    return os.path.exists(prefix + ".index")


def latest_checkpoint(checkpoint_dir: str, latest_filename: Optional[str] = None) -> Optional[str]:
    st = get_checkpoint_state(checkpoint_dir, latest_filename)
    if st and st.model_checkpoint_path and checkpoint_exists(st.model_checkpoint_path):
        return st.model_checkpoint_path
    return None


def remove_checkpoint(prefix: str) -> None:
    for p in glob.glob(glob.escape(prefix) + ".index") + glob.glob(glob.escape(prefix) + ".data-*") + \
            glob.glob(glob.escape(prefix) + ".meta"):
        try:
            os.remove(p)
        except OSError:
            pass


# ------------------------------------------------------------------------------------------------
# Saver
# ------------------------------------------------------------------------------------------------
class Saver:
    def __init__(self, var_list: Optional[Sequence[G.Variable]] = None, sharded: bool = True,
                 max_to_keep: int = 5, keep_checkpoint_every_n_hours: float = 10000.0,
                 save_relative_paths: bool = True):
        self._var_list = list(var_list) if var_list is not None else None
        self._sharded = sharded
        self._max_to_keep = max_to_keep
        self._last_checkpoints: List[str] = []
        self.last_save_seconds = 0.0
        self.last_save_bytes = 0

    @property
    def last_checkpoints(self) -> List[str]:
        return list(self._last_checkpoints)

    def _variables(self, sess) -> List[G.Variable]:
        return self._var_list if self._var_list is not None else sess.graph.variables()

    # -- save ---------------------------------------------------------------------------------------
    def save(self, sess, save_path: str, global_step=None, latest_filename: Optional[str] = None,
             write_state: bool = True, write_meta_graph: bool = True) -> str:
        t0 = time.time()
        if global_step is not None:
            if not isinstance(global_step, (int, np.integer)):
                global_step = int(sess.run(global_step))
            prefix = "%s-%d" % (save_path, int(global_step))
        else:
            prefix = save_path
        prefix = os.path.abspath(prefix)
        save_dir = os.path.dirname(prefix)
        os.makedirs(save_dir, exist_ok=True)
        names = {v.name for v in self._variables(sess)}
        self.last_save_bytes = sess.save_state(prefix, names, sharded=self._sharded)
        if write_meta_graph:      # TF's Saver.save exports <prefix>.meta with every checkpoint
            from .meta_graph import write_meta_graph as _write_meta
            _write_meta(prefix, self._variables(sess), sharded=self._sharded, max_to_keep=self._max_to_keep or 0)
        if write_state:
            self._last_checkpoints.append(prefix)
            while self._max_to_keep and len(self._last_checkpoints) > self._max_to_keep:
                remove_checkpoint(self._last_checkpoints.pop(0))
            update_checkpoint_state(save_dir, prefix, self._last_checkpoints, latest_filename)
        self.last_save_seconds = time.time() - t0
        return prefix

    # -- restore ------------------------------------------------------------------------------------
    def restore(self, sess, save_path: str) -> None:
        if save_path is None:
            raise ValueError("Can't load save_path when it is None.")
        if not checkpoint_exists(save_path):
            raise ValueError("The passed save_path is not a valid checkpoint: %s" % save_path)
        print("INFO:tensorflow:Restoring parameters from %s" % save_path, flush=True)   # TF1's log line
        names = {v.name for v in self._variables(sess)}
        sess.restore_state(save_path, names)

    def recover_last_checkpoints(self, checkpoint_paths: Sequence[str]) -> None:
        self._last_checkpoints = [p for p in checkpoint_paths if checkpoint_exists(p)]


# ------------------------------------------------------------------------------------------------
# helpers shared by the session implementations
# ------------------------------------------------------------------------------------------------
def temp_part_prefixes(prefix: str, num_parts: int) -> List[str]:
    tmp_dir = "%s_temp_%s" % (prefix, uuid.uuid4().hex)
    os.makedirs(tmp_dir, exist_ok=True)
    return [os.path.join(tmp_dir, "part-%05d-of-%05d" % (i, num_parts)) for i in range(num_parts)]


def write_part(part_prefix: str, tensors: Dict[str, np.ndarray]) -> int:
    n = 0
    with BundleWriter(part_prefix) as w:
        for name in sorted(tensors):
            w.add(name, tensors[name])
            n += tensors[name].nbytes
    return n


def finalize_parts(parts: Sequence[str], prefix: str) -> None:
    merge_bundles(list(parts), prefix, delete_old_dirs=True)
    tmp_dir = os.path.dirname(parts[0])
    if os.path.isdir(tmp_dir):
        shutil.rmtree(tmp_dir, ignore_errors=True)


def load_checkpoint_tensors(prefix: str, wanted: Optional[Sequence[str]] = None) -> Dict[str, np.ndarray]:
    with BundleReader(prefix) as r:
        keys = r.keys() if wanted is None else [k for k in wanted if r.has_tensor(k)]
        return {k: r.get_tensor(k) for k in keys}
