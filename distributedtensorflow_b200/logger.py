"""Key/value + text logger with stdout / log.txt / JSON-lines / CSV / TensorBoard sinks.

API-compatible with the reference's ``logger.py`` (the OpenAI-baselines logger; citations are
logger.py:line): output formats (:26-187), module-level ``logkv / logkv_mean / logkvs / dumpkvs /
getkvs / log / debug / info / warn / error / set_level / get_dir`` (:193-259), ``ProfileKV`` and
``profile`` wall-clock helpers (:261-285), the ``Logger`` backend with ``DEFAULT`` / ``CURRENT``
(:292-346), ``configure / reset / scoped_configure`` (:348-402) and the ``read_json / read_csv /
read_tb`` readers (:439-483). Written from scratch for this framework; differences on purpose:

* ``TensorBoardOutputFormat`` needs no TensorFlow: events go through the native C++ writer
  (``csrc/io/events_writer.cc``), and ``writekvs(kvs, global_step=None)`` falls back to an
  internal step counter, so ``dumpkvs()`` can drive it (the reference's signature mismatch,
  logger.py:157 vs :321, made that combination crash);
* ``flush_every`` lets the hot path avoid a flush per step (logger.py:165 flushes every call;
  that remains the default);
* the rank suffix also honours ``RANK`` (torchrun), next to ``PMI_RANK`` /
  ``OMPI_COMM_WORLD_RANK`` (logger.py:361).
"""
from __future__ import annotations

import datetime
import functools
import json
import os
import os.path as osp
import shutil
import sys
import tempfile
import time
from collections import defaultdict
from typing import List

DEBUG = 10
INFO = 20
WARN = 30
ERROR = 40
DISABLED = 50


class KVWriter(object):
    def writekvs(self, kvs):
        raise NotImplementedError


class SeqWriter(object):
    def writeseq(self, seq):
        raise NotImplementedError


def _clip(text: str) -> str:
    """Cells longer than 23 characters are cut to 20 + '...'."""
    return text[:20] + "..." if len(text) > 23 else text


class HumanOutputFormat(KVWriter, SeqWriter):
    """ASCII table of the current key/values, and plain text lines."""

    def __init__(self, filename_or_file):
        if isinstance(filename_or_file, str):
            self.file = open(filename_or_file, "wt")
            self.own_file = True
        else:
            assert hasattr(filename_or_file, "read") or hasattr(filename_or_file, "write"), \
                "expected file or str, got %s" % filename_or_file
            self.file = filename_or_file
            self.own_file = False

    def writekvs(self, kvs):
        cells = {}
        for key in sorted(kvs.keys()):
            val = kvs[key]
            text = "%-8.3g" % (val,) if isinstance(val, float) else str(val)
            cells[_clip(key)] = _clip(text)
        if not cells:
            print("WARNING: tried to write empty key-value dict")
            return
        kw = max(len(k) for k in cells)
        vw = max(len(v) for v in cells.values())
        rule = "-" * (kw + vw + 7)
        rows = [rule]
        for key in sorted(cells, key=str.lower):
            rows.append("| %s | %s |" % (key.ljust(kw), cells[key].ljust(vw)))
        rows.append(rule)
        self.file.write("\n".join(rows) + "\n")
        self.file.flush()

    def writeseq(self, seq):
        self.file.write(" ".join(list(seq)) + "\n")
        self.file.flush()

    def close(self):
        if self.own_file:
            self.file.close()


class JSONOutputFormat(KVWriter):
    """One JSON object per dump (``progress.json``)."""

    def __init__(self, filename):
        self.file = open(filename, "wt")

    def writekvs(self, kvs):
        row = {}
        for k in sorted(kvs.keys()):
            v = kvs[k]
            if hasattr(v, "dtype"):
                v = float(v.tolist())
            row[k] = v
        self.file.write(json.dumps(row) + "\n")
        self.file.flush()

    def close(self):
        self.file.close()


class CSVOutputFormat(KVWriter):
    """``progress.csv``; when a dump brings new keys the header is rewritten and earlier rows are
    padded with empty cells."""

    def __init__(self, filename):
        self.file = open(filename, "w+t")
        self.keys: List[str] = []
        self.sep = ","

    def writekvs(self, kvs):
        new_keys = sorted(k for k in kvs.keys() if k not in self.keys)
        if new_keys:
            self.keys.extend(new_keys)
            self.file.seek(0)
            old = self.file.readlines()
            self.file.seek(0)
            self.file.write(self.sep.join(self.keys) + "\n")
            for line in old[1:]:
                self.file.write(line.rstrip("\n") + self.sep * len(new_keys) + "\n")
        out = []
        for k in self.keys:
            v = kvs.get(k)
            out.append("" if v is None else str(v))
        self.file.write(self.sep.join(out) + "\n")
        self.file.flush()

    def close(self):
        self.file.close()


class TensorBoardOutputFormat(KVWriter):
    """Scalars into ``<dir>/events.out.tfevents.<unix>.<host>`` (the only logger piece the
    reference's training script uses: run_mnist_distributed.py:139,159)."""

    def __init__(self, dir, flush_every: int = 1):
        os.makedirs(dir, exist_ok=True)
        self.dir = dir
        self.step = 1
        self.flush_every = max(1, int(flush_every))
        self._since_flush = 0
        from .utils.events import EventsWriter
        self.writer = EventsWriter(osp.join(osp.abspath(dir), "events"))

    def writekvs(self, kvs, global_step=None):
        if global_step is None:
            global_step = self.step
            self.step += 1
        scalars = {k: float(v) for k, v in kvs.items()}
        self.writer.write_scalars(scalars, int(global_step), time.time())
        self._since_flush += 1
        if self._since_flush >= self.flush_every:
            self.writer.Flush()
            self._since_flush = 0

    def close(self):
        if self.writer:
            self.writer.Close()
            self.writer = None


def make_output_format(format, ev_dir, log_suffix=""):
    os.makedirs(ev_dir, exist_ok=True)
    if format == "stdout":
        return HumanOutputFormat(sys.stdout)
    if format == "log":
        return HumanOutputFormat(osp.join(ev_dir, "log%s.txt" % log_suffix))
    if format == "json":
        return JSONOutputFormat(osp.join(ev_dir, "progress%s.json" % log_suffix))
    if format == "csv":
        return CSVOutputFormat(osp.join(ev_dir, "progress%s.csv" % log_suffix))
    if format == "tensorboard":
        return TensorBoardOutputFormat(osp.join(ev_dir, "tb%s" % log_suffix))
    raise ValueError("Unknown format specified: %s" % (format,))


# ================================================================================================
# module-level API (forwards to Logger.CURRENT)
# ================================================================================================
def logkv(key, val):
    """Record ``val`` under ``key`` for this iteration; the last value wins."""
    Logger.CURRENT.logkv(key, val)


def logkv_mean(key, val):
    """Like ``logkv`` but repeated calls within one iteration are averaged."""
    Logger.CURRENT.logkv_mean(key, val)


def logkvs(d):
    for k, v in d.items():
        logkv(k, v)


def dumpkvs():
    """Write this iteration's diagnostics to every KV sink and clear them."""
    Logger.CURRENT.dumpkvs()


def getkvs():
    return Logger.CURRENT.name2val


def log(*args, level=INFO):
    """Write ``args`` (space separated) to the console and text sinks if ``level`` passes."""
    Logger.CURRENT.log(*args, level=level)


def debug(*args):
    log(*args, level=DEBUG)


def info(*args):
    log(*args, level=INFO)


def warn(*args):
    log(*args, level=WARN)


def error(*args):
    log(*args, level=ERROR)


def set_level(level):
    Logger.CURRENT.set_level(level)


def get_dir():
    """Directory the log files go to (None when no directory was configured)."""
    return Logger.CURRENT.get_dir()


record_tabular = logkv
dump_tabular = dumpkvs


class ProfileKV:
    """``with ProfileKV("scope"): ...`` adds the elapsed wall-clock seconds to ``wait_scope``."""

    def __init__(self, n):
        self.n = "wait_" + n

    def __enter__(self):
        self.t1 = time.time()

    def __exit__(self, type, value, traceback):
        Logger.CURRENT.name2val[self.n] += time.time() - self.t1


def profile(n):
    """Decorator form of :class:`ProfileKV`."""
    def decorate(func):
        @functools.wraps(func)
        def wrapped(*args, **kwargs):
            with ProfileKV(n):
                return func(*args, **kwargs)
        return wrapped
    return decorate


# ================================================================================================
# backend
# ================================================================================================
class Logger(object):
    DEFAULT = None  # stdout-only logger created at import
    CURRENT = None  # the logger the module-level functions talk to

    def __init__(self, dir, output_formats):
        self.name2val = defaultdict(float)
        self.name2cnt = defaultdict(int)
        self.level = INFO
        self.dir = dir
        self.output_formats = output_formats

    def logkv(self, key, val):
        self.name2val[key] = val

    def logkv_mean(self, key, val):
        if val is None:
            self.name2val[key] = None
            return
        cnt = self.name2cnt[key]
        self.name2val[key] = self.name2val[key] * cnt / (cnt + 1) + val / (cnt + 1)
        self.name2cnt[key] = cnt + 1

    def dumpkvs(self):
        if self.level == DISABLED:
            return
        for fmt in self.output_formats:
            if isinstance(fmt, KVWriter):
                fmt.writekvs(self.name2val)
        self.name2val.clear()
        self.name2cnt.clear()

    def log(self, *args, level=INFO):
        if self.level <= level:
            self._do_log(args)

    def set_level(self, level):
        self.level = level

    def get_dir(self):
        return self.dir

    def close(self):
        for fmt in self.output_formats:
            fmt.close()

    def _do_log(self, args):
        for fmt in self.output_formats:
            if isinstance(fmt, SeqWriter):
                fmt.writeseq(map(str, args))


def _detect_rank() -> int:
    rank = 0
    for var in ("RANK", "PMI_RANK", "OMPI_COMM_WORLD_RANK"):
        if var in os.environ:
            try:
                rank = int(os.environ[var])
            except ValueError:
                pass
    return rank


def configure(dir=None, format_strs=None):
    if dir is None:
        dir = os.getenv("OPENAI_LOGDIR")
    if dir is None:
        dir = osp.join(tempfile.gettempdir(),
                       datetime.datetime.now().strftime("openai-%Y-%m-%d-%H-%M-%S-%f"))
    assert isinstance(dir, str)
    os.makedirs(dir, exist_ok=True)

    rank = _detect_rank()
    log_suffix = "-rank%03i" % rank if rank > 0 else ""
    if format_strs is None:
        if rank == 0:
            format_strs = os.getenv("OPENAI_LOG_FORMAT", "stdout,log,csv").split(",")
        else:
            format_strs = os.getenv("OPENAI_LOG_FORMAT_MPI", "log").split(",")
    output_formats = [make_output_format(f, dir, log_suffix) for f in format_strs if f]
    Logger.CURRENT = Logger(dir=dir, output_formats=output_formats)
    log("Logging to %s" % dir)


def _configure_default_logger():
    format_strs = None
    if "OPENAI_LOG_FORMAT" not in os.environ:
        format_strs = ["stdout"]  # historic default: console only
    configure(format_strs=format_strs)
    Logger.DEFAULT = Logger.CURRENT


def reset():
    if Logger.CURRENT is not Logger.DEFAULT:
        Logger.CURRENT.close()
        Logger.CURRENT = Logger.DEFAULT
        log("Reset logger")


class scoped_configure(object):
    def __init__(self, dir=None, format_strs=None):
        self.dir = dir
        self.format_strs = format_strs
        self.prevlogger = None

    def __enter__(self):
        self.prevlogger = Logger.CURRENT
        configure(dir=self.dir, format_strs=self.format_strs)

    def __exit__(self, *args):
        Logger.CURRENT.close()
        Logger.CURRENT = self.prevlogger


def _demo():
    info("hi")
    debug("shouldn't appear")
    set_level(DEBUG)
    debug("should appear")
    dir = osp.join(tempfile.gettempdir(), "testlogging")
    if os.path.exists(dir):
        shutil.rmtree(dir)
    configure(dir=dir)
    logkv("a", 3)
    logkv("b", 2.5)
    dumpkvs()
    logkv("b", -2.5)
    logkv("a", 5.5)
    dumpkvs()
    info("^^^ should see a = 5.5")
    logkv_mean("b", -22.5)
    logkv_mean("b", -44.4)
    logkv("a", 5.5)
    dumpkvs()
    info("^^^ should see b = -33.45 (the mean of the two values)")
    logkv("b", -2.5)
    dumpkvs()
    logkv("a", "longasslongasslongasslongasslongasslongassvalue")
    dumpkvs()


# ================================================================================================
# readers
# ================================================================================================
def read_json(fname):
    import pandas
    rows = []
    with open(fname, "rt") as fh:
        for line in fh:
            if line.strip():
                rows.append(json.loads(line))
    return pandas.DataFrame(rows)


def read_csv(fname):
    import pandas
    return pandas.read_csv(fname, index_col=None, comment="#")


def read_tb(path):
    """``path``: one TensorBoard event file or a directory holding ``events.*`` files. Returns a
    DataFrame with one row per step (index = step - 1), one column per tag, NaN where a tag was
    not logged at that step."""
    import numpy as np
    import pandas
    from glob import glob
    from .utils.events import summary_iterator
    if osp.isdir(path):
        fnames = glob(osp.join(path, "events.*"))
    elif osp.basename(path).startswith("events."):
        fnames = [path]
    else:
        raise NotImplementedError(
            "Expected tensorboard file or directory containing them. Got %s" % path)
    tag2pairs = defaultdict(list)
    maxstep = 0
    for fname in fnames:
        for ev in summary_iterator(fname):
            if ev.step > 0:
                for v in ev.summary.value:
                    tag2pairs[v.tag].append((ev.step, v.simple_value))
                maxstep = max(ev.step, maxstep)
    tags = sorted(tag2pairs.keys())
    data = np.full((maxstep, len(tags)), np.nan)
    for col, tag in enumerate(tags):
        for step, value in tag2pairs[tag]:
            data[step - 1, col] = value
    return pandas.DataFrame(data, columns=tags)


_configure_default_logger()

if __name__ == "__main__":
    _demo()
