"""Tracing helpers.

The reference only accumulates wall-clock seconds into its logger (``ProfileKV`` / ``@profile``,
logger.py:261-285 — kept in :mod:`distributedtensorflow_b200.logger`). On the B200 the useful
tools are device timing and NVTX ranges that ``ncu --nvtx`` (or any CUPTI consumer) can filter on:

* ``nvtx_range("name")``     host-side NVTX range; a no-op unless ``DTF_B200_NVTX=1`` (or
                             ``enable_nvtx(True)``), so the 160 us step pays nothing for it.
* ``DeviceTimer``            CUDA-event timer on the launching stream (the timing rule of
                             B200_PROFILING.md: warm up, synchronize both sides, never wall clock).
"""
from __future__ import annotations

import contextlib
import os
from typing import Iterator, List

_NVTX = os.environ.get("DTF_B200_NVTX", "0") == "1"


def enable_nvtx(on: bool = True) -> None:
    global _NVTX
    _NVTX = bool(on)


def nvtx_enabled() -> bool:
    return _NVTX


class _Null:
    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False


_NULL = _Null()


class _Range:
    __slots__ = ("name",)

    def __init__(self, name: str):
        self.name = name

    def __enter__(self):
        import torch
        torch.cuda.nvtx.range_push(self.name)
        return self

    def __exit__(self, *exc):
        import torch
        torch.cuda.nvtx.range_pop()
        return False


def nvtx_range(name: str):
    """``with nvtx_range("dtf/fwd"): ...`` — free when NVTX is off."""
    return _Range(name) if _NVTX else _NULL


class DeviceTimer:
    """Accumulates CUDA-event timings of regions on one stream.

    >>> t = DeviceTimer(stream)
    >>> with t.region(): launch()
    >>> t.milliseconds()     # synchronises, returns the list of region times
    """

    def __init__(self, stream=None):
        import torch
        self._torch = torch
        self.stream = stream
        self._pairs: List[tuple] = []

    @contextlib.contextmanager
    def region(self) -> Iterator[None]:
        torch = self._torch
        s = self.stream if self.stream is not None else torch.cuda.current_stream()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(s)
        try:
            yield
        finally:
            b.record(s)
            self._pairs.append((a, b))

    def milliseconds(self, reset: bool = True) -> List[float]:
        out = []
        for a, b in self._pairs:
            b.synchronize()
            out.append(a.elapsed_time(b))
        if reset:
            self._pairs = []
        return out

    def total_ms(self, reset: bool = True) -> float:
        return float(sum(self.milliseconds(reset)))
