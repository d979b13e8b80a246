"""TensorBoard event files: writer (native C++ with a pure-Python fallback) and reader.

Parity target: ``pywrap_tensorflow.EventsWriter`` as driven by the reference's
``TensorBoardOutputFormat`` (logger.py:138-171) and ``tf.train.summary_iterator`` used by
``read_tb`` (logger.py:451-483). File ``<prefix>.out.tfevents.<unix>.<host>``; records are
TFRecord frames ``u64 len | u32 masked_crc(len) | bytes | u32 masked_crc(bytes)``; the first
record is ``Event(file_version="brain.Event:2")``.
"""
from __future__ import annotations

import ctypes as C
import os
import socket
import struct
import time
from dataclasses import dataclass, field
from typing import Dict, Iterator, List, Optional

from . import protowire as pw

# ------------------------------------------------------------------------------------------------
# crc32c
# ------------------------------------------------------------------------------------------------
_TABLE: Optional[List[int]] = None


def _table() -> List[int]:
    global _TABLE
    if _TABLE is None:
        t = []
        for i in range(256):
            c = i
            for _ in range(8):
                c = (c >> 1) ^ 0x82F63B78 if c & 1 else c >> 1
            t.append(c)
        _TABLE = t
    return _TABLE


def crc32c_py(data: bytes, crc: int = 0) -> int:
    t = _table()
    c = crc ^ 0xFFFFFFFF
    for b in data:
        c = t[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def crc32c(data: bytes) -> int:
    try:
        from . import hostlib
        return int(hostlib.lib().dtf_crc32c(data, len(data)))
    except Exception:
        return crc32c_py(data)


def mask_crc(crc: int) -> int:
    return (((crc >> 15) | (crc << 17)) + 0xA282EAD8) & 0xFFFFFFFF


def unmask_crc(masked: int) -> int:
    rot = (masked - 0xA282EAD8) & 0xFFFFFFFF
    return ((rot >> 17) | (rot << 15)) & 0xFFFFFFFF


def masked_crc32c(data: bytes) -> int:
    return mask_crc(crc32c(data))


# ------------------------------------------------------------------------------------------------
# Event / Summary protos
# ------------------------------------------------------------------------------------------------
def encode_scalar_event(kvs: Dict[str, float], step: int, wall_time: Optional[float] = None) -> bytes:
    summary = b""
    for tag, v in kvs.items():
        val = pw.field_string(1, tag) + pw.field_float(2, float(v))
        summary += pw.field_bytes(1, val)
    ev = pw.field_double(1, time.time() if wall_time is None else wall_time)
    ev += pw.field_varint(2, int(step))
    ev += pw.field_bytes(5, summary)
    return ev


def encode_file_version_event(wall_time: float) -> bytes:
    return pw.field_double(1, wall_time) + pw.field_string(3, "brain.Event:2")


def frame_record(data: bytes) -> bytes:
    header = struct.pack("<Q", len(data))
    return header + struct.pack("<I", masked_crc32c(header)) + data + struct.pack("<I", masked_crc32c(data))


@dataclass
class SummaryValue:
    tag: str = ""
    simple_value: float = 0.0


@dataclass
class Summary:
    value: List[SummaryValue] = field(default_factory=list)


@dataclass
class Event:
    wall_time: float = 0.0
    step: int = 0
    file_version: str = ""
    summary: Summary = field(default_factory=Summary)


def decode_event(buf: bytes) -> Event:
    ev = Event()
    for f, wt, v in pw.iter_fields(buf):
        if f == 1 and wt == pw.WT_FIXED64:
            ev.wall_time = pw.as_double(v)
        elif f == 2 and wt == pw.WT_VARINT:
            ev.step = pw.as_signed(v)
        elif f == 3 and wt == pw.WT_LEN:
            ev.file_version = v.decode()
        elif f == 5 and wt == pw.WT_LEN:
            for f2, wt2, v2 in pw.iter_fields(v):
                if f2 == 1 and wt2 == pw.WT_LEN:
                    sv = SummaryValue()
                    for f3, wt3, v3 in pw.iter_fields(v2):
                        if f3 == 1 and wt3 == pw.WT_LEN:
                            sv.tag = v3.decode()
                        elif f3 == 2 and wt3 == pw.WT_FIXED32:
                            sv.simple_value = pw.as_float(v3)
                    ev.summary.value.append(sv)
    return ev


# ------------------------------------------------------------------------------------------------
# writer
# ------------------------------------------------------------------------------------------------
class EventsWriter:
    """``EventsWriter(prefix)``: native when libdtf_b200_host is available, else pure Python."""

    def __init__(self, file_prefix: str, use_native: Optional[bool] = None):
        self._native = None
        self._file = None
        self._lib = None
        if use_native is None or use_native:
            try:
                from . import hostlib
                self._lib = hostlib.lib()
                h = self._lib.dtf_events_open(os.fsencode(file_prefix), None)
                if not h:
                    raise OSError(hostlib.last_error())
                self._native = h
                self.filename = self._lib.dtf_events_filename(h).decode()
            except Exception:
                if use_native:
                    raise
                self._native = None
        if self._native is None:
            now = time.time()
            self.filename = "%s.out.tfevents.%010d.%s" % (file_prefix, int(now), socket.gethostname())
            self._file = open(self.filename, "wb")
            self._file.write(frame_record(encode_file_version_event(float(int(now)))))
            self._file.flush()

    @property
    def is_native(self) -> bool:
        return self._native is not None

    def write_scalars(self, kvs: Dict[str, float], step: int, wall_time: Optional[float] = None) -> None:
        wall_time = time.time() if wall_time is None else wall_time
        if self._native is not None:
            n = len(kvs)
            tags = (C.c_char_p * n)(*[k.encode() for k in kvs.keys()])
            vals = (C.c_float * n)(*[float(v) for v in kvs.values()])
            rc = self._lib.dtf_events_write_scalars(self._native, wall_time, int(step), n, tags, vals)
            if rc != 0:
                raise OSError("event write failed")
        else:
            self._file.write(frame_record(encode_scalar_event(kvs, step, wall_time)))

    def WriteEvent(self, event_bytes: bytes) -> None:
        if self._native is not None:
            self._lib.dtf_events_write_raw(self._native, event_bytes, len(event_bytes))
        else:
            self._file.write(frame_record(event_bytes))

    def Flush(self) -> None:
        if self._native is not None:
            self._lib.dtf_events_flush(self._native)
        elif self._file is not None:
            self._file.flush()

    def Close(self) -> None:
        if self._native is not None:
            self._lib.dtf_events_close(self._native)
            self._native = None
        if self._file is not None:
            self._file.close()
            self._file = None

    flush, close = Flush, Close

    def __del__(self):
        try:
            self.Close()
        except Exception:
            pass


class EventFileWriter:
    """``tf.summary.FileWriter``-like convenience used by the summary / step-counter hooks.
    Flushes on a timer (``flush_secs``) instead of after every event (the reference flushes every
    step on the chief's hot path, logger.py:165)."""

    def __init__(self, logdir: str, flush_secs: float = 120.0, filename_prefix: str = "events"):
        os.makedirs(logdir, exist_ok=True)
        self._writer = EventsWriter(os.path.join(os.path.abspath(logdir), filename_prefix))
        self._flush_secs = flush_secs
        self._last_flush = time.time()

    def add_scalars(self, kvs: Dict[str, float], global_step: int) -> None:
        self._writer.write_scalars(kvs, global_step)
        if time.time() - self._last_flush >= self._flush_secs:
            self.flush()

    def flush(self) -> None:
        self._writer.Flush()
        self._last_flush = time.time()

    def close(self) -> None:
        self._writer.Close()

    @property
    def filename(self) -> str:
        return self._writer.filename


# ------------------------------------------------------------------------------------------------
# reader
# ------------------------------------------------------------------------------------------------
def tf_record_iterator(path: str, verify: bool = True) -> Iterator[bytes]:
    with open(path, "rb") as f:
        while True:
            header = f.read(12)
            if not header:
                return
            if len(header) != 12:
                raise IOError("truncated record header in %s" % path)
            (length,) = struct.unpack("<Q", header[:8])
            (lcrc,) = struct.unpack("<I", header[8:])
            if verify and masked_crc32c(header[:8]) != lcrc:
                raise IOError("corrupted record length in %s" % path)
            data = f.read(length)
            tail = f.read(4)
            if len(data) != length or len(tail) != 4:
                raise IOError("truncated record in %s" % path)
            if verify and masked_crc32c(data) != struct.unpack("<I", tail)[0]:
                raise IOError("corrupted record data in %s" % path)
            yield data


def summary_iterator(path: str) -> Iterator[Event]:
    """``tf.train.summary_iterator``: yields the Events of one event file."""
    for rec in tf_record_iterator(path):
        yield decode_event(rec)
