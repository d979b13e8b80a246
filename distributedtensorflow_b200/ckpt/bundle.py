"""Python façade over the native TF1 tensor-bundle reader/writer (``csrc/io/tensor_bundle.cc``).

Format notes live in the C++ file and in SURVEY.md §5.4. ``DataType`` values are TensorFlow's.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, List, Optional, Sequence

import numpy as np

from ..utils import hostlib

DT_FLOAT, DT_DOUBLE, DT_INT32, DT_UINT8, DT_INT64, DT_BOOL, DT_BFLOAT16 = 1, 2, 3, 4, 9, 10, 14
_NP_TO_DT = {np.dtype(np.float32): DT_FLOAT, np.dtype(np.float64): DT_DOUBLE,
             np.dtype(np.int32): DT_INT32, np.dtype(np.uint8): DT_UINT8,
             np.dtype(np.int64): DT_INT64, np.dtype(np.bool_): DT_BOOL}
_DT_TO_NP = {v: k for k, v in _NP_TO_DT.items()}
_DT_TO_NP[DT_BFLOAT16] = np.dtype(np.uint16)  # raw bf16 bits


class BundleWriter:
    """Writes ``<prefix>.index`` + ``<prefix>.data-00000-of-00001``."""

    def __init__(self, prefix: str):
        os.makedirs(os.path.dirname(os.path.abspath(prefix)), exist_ok=True)
        self._lib = hostlib.lib()
        self._h = self._lib.dtf_bundle_writer_open(os.fsencode(prefix))
        if not self._h:
            raise IOError(hostlib.last_error())
        self.prefix = prefix
        self.bytes_written = 0

    def add(self, name: str, array: np.ndarray, dtype_enum: Optional[int] = None) -> None:
        src = np.asarray(array)
        a = np.ascontiguousarray(src).reshape(src.shape)  # ascontiguousarray promotes 0-d to 1-d
        dt = dtype_enum if dtype_enum is not None else _NP_TO_DT[a.dtype]
        dims = (C.c_int64 * max(1, a.ndim))(*a.shape)
        rc = self._lib.dtf_bundle_writer_add(self._h, name.encode(), dt, a.ndim, dims,
                                             a.ctypes.data_as(C.c_void_p), a.nbytes)
        hostlib.check(rc, f"bundle add {name}")
        self.bytes_written += a.nbytes

    def finish(self) -> None:
        h, self._h = self._h, None
        hostlib.check(self._lib.dtf_bundle_writer_finish(h), "bundle finish")

    def __enter__(self):
        return self

    def __exit__(self, exc_type, exc, tb):
        if self._h:
            self.finish()


class BundleIndexWriter:
    """Writes only ``<prefix>.index``: the data shards were written by their owners (every PS task
    stores the bytes it holds straight into ``<prefix>.data-0000i-of-0000N`` at planned offsets)."""

    def __init__(self, prefix: str, num_shards: int):
        self._lib = hostlib.lib()
        self._h = self._lib.dtf_bundle_index_open(os.fsencode(prefix), int(num_shards))
        if not self._h:
            raise IOError(hostlib.last_error())
        self.prefix, self.num_shards = prefix, int(num_shards)

    def data_path(self, shard_id: int) -> str:
        buf = C.create_string_buffer(4096)
        hostlib.check(self._lib.dtf_bundle_index_data_path(self._h, int(shard_id), buf, 4096), "data path")
        return os.fsdecode(buf.value)

    def add(self, name: str, dtype_enum: int, shape: Sequence[int], shard_id: int, offset: int, size: int,
            crc32c: int) -> None:
        dims = (C.c_int64 * max(1, len(shape)))(*[int(d) for d in shape])
        rc = self._lib.dtf_bundle_index_add(self._h, name.encode(), int(dtype_enum), len(shape), dims,
                                            int(shard_id), int(offset), int(size), int(crc32c) & 0xffffffff)
        hostlib.check(rc, f"bundle index add {name}")

    def finish(self) -> None:
        h, self._h = self._h, None
        hostlib.check(self._lib.dtf_bundle_index_finish(h), "bundle index finish")


def crc32c(data) -> int:
    a = np.ascontiguousarray(data)
    return int(hostlib.lib().dtf_crc32c(a.ctypes.data_as(C.c_void_p), a.nbytes))


def crc32c_combine(crc_a: int, crc_b: int, len_b: int) -> int:
    return int(hostlib.lib().dtf_crc32c_combine(int(crc_a), int(crc_b), int(len_b)))


class BundleReader:
    def __init__(self, prefix: str):
        self._lib = hostlib.lib()
        self._h = self._lib.dtf_bundle_reader_open(os.fsencode(prefix))
        if not self._h:
            raise IOError(hostlib.last_error())
        self.prefix = prefix
        self.num_shards = self._lib.dtf_bundle_reader_num_shards(self._h)
        self.entries: Dict[str, dict] = {}
        name_buf = C.create_string_buffer(4096)
        dims = (C.c_int64 * 8)()
        dtype, ndim, shard = C.c_int(), C.c_int(), C.c_int()
        nbytes = C.c_int64()
        for i in range(self._lib.dtf_bundle_reader_num_entries(self._h)):
            rc = self._lib.dtf_bundle_reader_entry(self._h, i, name_buf, 4096, C.byref(dtype),
                                                   C.byref(ndim), dims, C.byref(nbytes), C.byref(shard))
            hostlib.check(rc, "bundle entry")
            self.entries[name_buf.value.decode()] = dict(
                dtype=dtype.value, shape=tuple(dims[k] for k in range(ndim.value)),
                nbytes=nbytes.value, shard_id=shard.value)

    def keys(self) -> List[str]:
        return list(self.entries.keys())

    def has_tensor(self, name: str) -> bool:
        return name in self.entries

    def get_tensor(self, name: str, verify_crc: bool = True) -> np.ndarray:
        e = self.entries[name]
        out = np.empty(e["shape"], _DT_TO_NP[e["dtype"]])
        assert out.nbytes == e["nbytes"], (name, out.nbytes, e)
        rc = self._lib.dtf_bundle_reader_read(self._h, name.encode(), out.ctypes.data_as(C.c_void_p),
                                              out.nbytes, int(verify_crc))
        hostlib.check(rc, f"bundle read {name}")
        return out

    def get_variable_to_shape_map(self) -> Dict[str, tuple]:
        return {k: v["shape"] for k, v in self.entries.items()}

    def close(self) -> None:
        if self._h:
            self._lib.dtf_bundle_reader_close(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def merge_bundles(part_prefixes: Sequence[str], merged_prefix: str, delete_old_dirs: bool = True) -> None:
    """MergeV2Checkpoints: per-PS single-shard bundles -> one N-shard checkpoint."""
    lib = hostlib.lib()
    arr = (C.c_char_p * len(part_prefixes))(*[os.fsencode(p) for p in part_prefixes])
    rc = lib.dtf_bundle_merge(arr, len(part_prefixes), os.fsencode(merged_prefix), int(delete_old_dirs))
    hostlib.check(rc, "bundle merge")
