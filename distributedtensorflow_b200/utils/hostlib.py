"""ctypes binding of ``libdtf_b200_host.so`` (C ABI in ``include/dtf_b200_io.h``): crc32c,
TensorBoard event files, LevelDB tables, TF1 tensor-bundle checkpoints. Built on demand with g++
when missing (seconds; no CUDA needed)."""
from __future__ import annotations

import ctypes as C
from pathlib import Path

LIB_PATH = Path(__file__).resolve().parent.parent / "lib" / "libdtf_b200_host.so"

_lib = None


class HostLibError(RuntimeError):
    pass


def _declare(lib):
    vp, i, sz, cp = C.c_void_p, C.c_int, C.c_size_t, C.c_char_p
    u32, i64 = C.c_uint32, C.c_int64
    lib.dtf_io_last_error.restype = cp
    for name, res, args in [
        ("dtf_crc32c", u32, [vp, sz]),
        ("dtf_crc32c_extend", u32, [u32, vp, sz]),
        ("dtf_crc32c_mask", u32, [u32]),
        ("dtf_crc32c_unmask", u32, [u32]),
        ("dtf_crc32c_is_accelerated", i, []),
        ("dtf_crc32c_combine", u32, [u32, u32, C.c_uint64]),
        ("dtf_bundle_index_open", vp, [cp, i]),
        ("dtf_bundle_index_add", i, [vp, cp, i, i, C.POINTER(i64), i, i64, i64, u32]),
        ("dtf_bundle_index_data_path", i, [vp, i, cp, sz]),
        ("dtf_bundle_index_finish", i, [vp]),
        ("dtf_events_open", vp, [cp, cp]),
        ("dtf_events_filename", cp, [vp]),
        ("dtf_events_write_scalars", i, [vp, C.c_double, i64, i, C.POINTER(cp), C.POINTER(C.c_float)]),
        ("dtf_events_write_raw", i, [vp, vp, sz]),
        ("dtf_events_flush", i, [vp]),
        ("dtf_events_close", i, [vp]),
        ("dtf_tfrecord_open", vp, [cp]),
        ("dtf_tfrecord_next", i, [vp, C.POINTER(vp), C.POINTER(sz)]),
        ("dtf_tfrecord_close", i, [vp]),
        ("dtf_bundle_writer_open", vp, [cp]),
        ("dtf_bundle_writer_add", i, [vp, cp, i, i, C.POINTER(i64), vp, sz]),
        ("dtf_bundle_writer_finish", i, [vp]),
        ("dtf_bundle_merge", i, [C.POINTER(cp), i, cp, i]),
        ("dtf_bundle_reader_open", vp, [cp]),
        ("dtf_bundle_reader_num_shards", i, [vp]),
        ("dtf_bundle_reader_num_entries", i, [vp]),
        ("dtf_bundle_reader_entry", i, [vp, i, cp, sz, C.POINTER(i), C.POINTER(i), C.POINTER(i64),
                                       C.POINTER(i64), C.POINTER(i)]),
        ("dtf_bundle_reader_read", i, [vp, cp, vp, sz, i]),
        ("dtf_bundle_reader_close", i, [vp]),
        ("dtf_table_builder_open", vp, [cp, i]),
        ("dtf_table_builder_add", i, [vp, vp, sz, vp, sz]),
        ("dtf_table_builder_finish", i, [vp]),
    ]:
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    return lib


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        from .. import _build
        _build.build_host()
    try:
        _lib = _declare(C.CDLL(str(LIB_PATH)))
    except OSError as e:
        raise HostLibError(f"cannot load {LIB_PATH}: {e}") from e
    return _lib


def available() -> bool:
    try:
        lib()
        return True
    except Exception:
        return False


def last_error() -> str:
    return lib().dtf_io_last_error().decode(errors="replace")


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise HostLibError(f"{what} failed ({rc}): {last_error()}")
