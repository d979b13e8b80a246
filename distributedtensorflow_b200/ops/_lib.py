"""ctypes binding of ``libdtf_b200_cuda.so`` (C ABI in ``include/dtf_b200.h``).

The library is plain CUDA (no PyTorch headers). Tensors cross the boundary as raw device
pointers (``tensor.data_ptr()``) and kernels are enqueued on the caller's current CUDA stream, so
they compose with ``torch.cuda.CUDAGraph`` capture. If a GPU is present and the library is
missing or fails to load, importing raises: there is no silent eager fallback on a GPU box.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

LIB_PATH = Path(__file__).resolve().parent.parent / "lib" / "libdtf_b200_cuda.so"

MAX_WORKERS = 16
MAX_SHARDS = 8
MAX_APPLY_ITEMS = 24
MAX_COPY_ITEMS = 32
IPC_HANDLE_BYTES = 64

OPT_SGD, OPT_MOMENTUM, OPT_ADAGRAD, OPT_ADAM = 0, 1, 2, 3


class PsCtrl(C.Structure):
    """Mirror of ``dtf_ps_ctrl`` (256 bytes)."""
    _fields_ = [
        ("global_step", C.c_ulonglong),
        ("beta1_power", C.c_float),
        ("beta2_power", C.c_float),
        ("initialized", C.c_uint),
        ("epoch", C.c_uint),
        ("shutdown", C.c_uint),
        ("apply_count", C.c_uint),
        ("heartbeat", C.c_ulonglong * MAX_WORKERS),
        ("arrive", C.c_uint * MAX_WORKERS),
        ("done", C.c_uint * MAX_SHARDS),
    ]


assert C.sizeof(PsCtrl) == 256
CTRL_BYTES = 256
CTRL_OFF_GLOBAL_STEP = 0
CTRL_OFF_BETA1 = 8
CTRL_OFF_BETA2 = 12
CTRL_OFF_INITIALIZED = 16
CTRL_OFF_EPOCH = 20
CTRL_OFF_SHUTDOWN = 24
CTRL_OFF_APPLY_COUNT = 28
CTRL_OFF_HEARTBEAT = 32


class GemmEpilogue(C.Structure):
    _fields_ = [
        ("out", C.c_void_p),
        ("ldo", C.c_int),
        ("out_f32", C.c_int),
        ("bias", C.c_void_p),
        ("relu", C.c_int),
        ("mask", C.c_void_p),
        ("ldmask", C.c_int),
        ("reduce", C.c_int),
        ("alpha", C.c_float),
        ("splitk_ws", C.c_void_p),
        ("ldws", C.c_int),
        ("tile_counters", C.c_void_p),
        ("apply_enable", C.c_int),
        ("apply_opt", C.c_int),
        ("apply_rows_per_slice", C.c_int),
        ("apply_shadow_ld", C.c_int),
        ("apply_lr", C.c_float), ("apply_beta1", C.c_float), ("apply_beta2", C.c_float),
        ("apply_epsilon", C.c_float), ("apply_momentum", C.c_float), ("apply_grad_scale", C.c_float),
        ("apply_use_nesterov", C.c_int),
        ("apply_reserved2", C.c_int),
        ("apply_ctrl_powers", C.c_void_p),
        ("apply_var", C.c_void_p * 8),
        ("apply_slot0", C.c_void_p * 8),
        ("apply_slot1", C.c_void_p * 8),
        ("apply_shadow", C.c_void_p * 8),
        ("scatter_enable", C.c_int),
        ("scatter_rows_per_slice", C.c_int),
        ("scatter_depth", C.c_int),
        ("scatter_reserved", C.c_int),
        ("scatter_seq", C.c_void_p),
        ("scatter_base", C.c_void_p * 8),
        ("scatter_buf_stride", C.c_longlong * 8),
        ("pool_out", C.c_void_p),
        ("pool_amax", C.c_void_p),
        ("pool_dbias", C.c_void_p),
    ]


class ApplyItem(C.Structure):
    _fields_ = [
        ("var", C.c_void_p),
        ("slot0", C.c_void_p),
        ("slot1", C.c_void_p),
        ("shadow", C.c_void_p),
        ("grad", C.c_void_p),
        ("n", C.c_longlong),
        ("first", C.c_longlong),
        ("cols", C.c_int),
        ("shadow_ld", C.c_int),
        ("zero_grad", C.c_int),
        ("reserved", C.c_int),
    ]


class ApplyHyper(C.Structure):
    _fields_ = [
        ("lr", C.c_float),
        ("beta1", C.c_float),
        ("beta2", C.c_float),
        ("epsilon", C.c_float),
        ("momentum", C.c_float),
        ("grad_scale", C.c_float),
        ("use_nesterov", C.c_int),
        ("atomic_sgd", C.c_int),
    ]


class SyncSegment(C.Structure):
    """Mirror of ``dtf_sync_segment``."""
    _fields_ = [
        ("owner", C.c_int), ("cols", C.c_int), ("shadow_ld", C.c_int), ("reserved", C.c_int),
        ("n", C.c_longlong), ("first", C.c_longlong), ("first_row", C.c_longlong),
        ("grad_local", C.c_ulonglong), ("mailbox_off", C.c_ulonglong), ("mailbox_stride", C.c_ulonglong),
        ("var_off", C.c_ulonglong), ("slot0_off", C.c_ulonglong), ("slot1_off", C.c_ulonglong),
        ("shadow_off", C.c_ulonglong), ("mirror_f32_off", C.c_ulonglong), ("mirror_bf16_off", C.c_ulonglong),
    ]


class CopyItem(C.Structure):
    _fields_ = [("src", C.c_void_p), ("dst", C.c_void_p), ("bytes", C.c_longlong)]


class MboxPushItem(C.Structure):
    """Mirror of ``dtf_mbox_push_item``."""
    _fields_ = [("src", C.c_void_p), ("dst0", C.c_void_p), ("bytes", C.c_longlong),
                ("buf_stride", C.c_longlong), ("zero_src", C.c_int), ("owner", C.c_int)]


MAX_MBOX_ITEMS = 64
MBOX_FLAGS_BYTES = 512
MBOX_OFF_PRODUCED, MBOX_OFF_CONSUMED, MBOX_OFF_DONE = 0, 128, 256
MBOX_OFF_TAG0, MBOX_OFF_TAG1 = 64, 192
MBOX_OFF_LOCK, MBOX_OFF_PAUSED, MBOX_OFF_DRAINS, MBOX_OFF_APPLIED, MBOX_OFF_MIRROR_BASE = 320, 324, 328, 332, 384

# worker mirror header (dtf_mirror_hdr, 512 bytes at offset 0 of every worker mirror)
MIRROR_HDR_BYTES = 512
MIRROR_MAX_SHARDS = 32
MIRROR_OFF_GLOBAL_STEP, MIRROR_OFF_EPOCH, MIRROR_OFF_ABORT, MIRROR_OFF_VERSION, MIRROR_OFF_CONSUMED = 0, 8, 12, 128, 256


class DrainMirrorItem(C.Structure):
    """Mirror of ``dtf_drain_mirror_item``."""
    _fields_ = [("bf16_off", C.c_longlong), ("f32_off", C.c_longlong), ("first", C.c_longlong)]


class ResultSink(C.Structure):
    """Mirror of ``dtf_result_sink``."""
    _fields_ = [("loss_src", C.c_void_p), ("host_dst", C.c_void_p), ("seq_dev", C.c_void_p),
                ("clear_loss", C.c_int), ("reserved", C.c_int)]


class NativeError(RuntimeError):
    pass


_lib = None


def available() -> bool:
    return LIB_PATH.exists()


def _declare(lib):
    vp, i, f, ll = C.c_void_p, C.c_int, C.c_float, C.c_longlong
    lib.dtf_last_error.restype = C.c_char_p
    lib.dtf_abi_version.restype = i
    sigs = {
        "dtf_gemm_bf16": [vp, i, i, vp, i, i, i, i, i, C.POINTER(GemmEpilogue), i, i, vp],
        "dtf_softmax_xent_sparse": [vp, i, vp, i, i, vp, vp, i, vp, vp],
        "dtf_softmax_xent_dense": [vp, i, vp, i, i, i, vp, vp, i, vp, vp],
        "dtf_push_apply": [i, C.POINTER(ApplyItem), i, C.POINTER(ApplyHyper), vp, vp, i, i, vp, vp, vp],
        "dtf_push_apply_ex": [i, C.POINTER(ApplyItem), i, C.POINTER(ApplyHyper), vp, vp, i, i, vp, vp,
                              C.POINTER(ResultSink), i, vp],
        "dtf_debug_div_sqrt": [vp, vp, vp, vp, vp, vp, ll, vp],
        "dtf_pull_bulk": [C.POINTER(CopyItem), i, vp, vp, vp],
        "dtf_refresh_shadow": [vp, vp, ll, ll, i, i, vp],
        "dtf_u8_to_bf16": [vp, vp, i, i, i, f, vp],
        "dtf_f32_to_bf16": [vp, i, vp, i, i, i, vp],
        "dtf_colsum_bf16": [vp, i, i, i, vp, vp],
        "dtf_fill_u32": [vp, C.c_uint, ll, vp],
        "dtf_conv1_fwd": [vp, vp, f, vp, vp, vp, vp, i, vp],
        "dtf_conv1_fwd_ingest": [vp, vp, f, vp, vp, vp, vp, vp, vp, vp, vp, i, i, vp],
        "dtf_conv1_bwd": [vp, vp, f, vp, vp, vp, vp, vp, i, vp],
        "dtf_conv2_fwd": [vp, vp, vp, vp, vp, i, vp],
        "dtf_conv2_fwd_halo": [vp, vp, vp, vp, vp, i, vp],
        "dtf_conv2_dgrad_halo": [vp, vp, vp, i, vp],
        "dtf_pool2_bwd": [vp, i, vp, vp, vp, vp, i, vp],
        "dtf_conv2_wgrad": [vp, vp, vp, i, vp],
        "dtf_conv2_dgrad": [vp, vp, vp, i, vp],
        "dtf_arena_alloc": [C.POINTER(vp), C.c_size_t],
        "dtf_arena_free": [vp],
        "dtf_ipc_export": [vp, C.c_char_p],
        "dtf_ipc_import": [C.c_char_p, C.POINTER(vp)],
        "dtf_ipc_close": [vp],
        "dtf_enable_peer_access": [i],
        "dtf_can_access_peer": [i, i],
        "dtf_device_count": [],
        "dtf_set_device": [i],
        "dtf_memcpy_d2h": [vp, vp, C.c_size_t, vp],
        "dtf_memcpy_h2d": [vp, vp, C.c_size_t, vp],
        "dtf_stream_sync": [vp],
        "dtf_set_pdl": [i],
        "dtf_sync_round": [i, vp, i, i, i, C.c_uint, C.POINTER(ApplyHyper), C.POINTER(C.c_ulonglong), vp, vp, i, vp],
        "dtf_mailbox_push": [C.POINTER(MboxPushItem), i, i, i, C.POINTER(vp), i, i, vp, vp, vp, vp, vp,
                             C.POINTER(ResultSink), vp, vp, vp],
        "dtf_mailbox_drain_ex": [i, C.POINTER(ApplyItem), i, C.c_longlong, i, i, C.POINTER(ApplyHyper), vp, vp,
                                 vp, vp, vp, vp, C.POINTER(DrainMirrorItem), i, i, vp],
        "dtf_mailbox_drain": [i, C.POINTER(ApplyItem), i, C.c_longlong, i, i, C.POINTER(ApplyHyper), vp, vp,
                              vp, vp, vp, vp, vp],
        "dtf_mbox_daemon_start": [C.POINTER(vp), i, i, C.POINTER(ApplyItem), i, C.c_longlong, i, i,
                                  C.POINTER(ApplyHyper), vp, vp, vp, i, C.POINTER(DrainMirrorItem), i, i, i],
        "dtf_mailbox_drain_sync": [i, C.POINTER(ApplyItem), i, C.c_longlong, i, i, C.POINTER(ApplyHyper), vp, vp,
                                   vp, vp, vp, C.POINTER(DrainMirrorItem), i, i, vp],
        "dtf_sync_token_wait": [vp, i, vp, vp, vp, vp],
        "dtf_mbox_daemon_set_backstop": [vp, i],
        "dtf_mbox_daemon_set_max_ctas": [vp, i],
        "dtf_mbox_daemon_pause": [vp, i],
        "dtf_mbox_daemon_stats": [vp, C.POINTER(C.c_ulonglong), C.POINTER(C.c_ulonglong), C.POINTER(C.c_ulonglong)],
        "dtf_mbox_daemon_stop": [vp],
        "dtf_graph_launch": [vp, vp],
        "dtf_step_report": [vp, i, i, vp, C.POINTER(ResultSink), vp],
        "dtf_host_wait_u32": [vp, C.c_uint, C.c_longlong],
        "dtf_head_fused": [vp, i, i, vp, i, vp, vp, vp, i, i, i, vp, i, vp, vp, i, vp, vp, i, vp, vp, vp],
        "dtf_host_alloc_pinned": [C.POINTER(vp), C.c_size_t],
        "dtf_host_free_pinned": [vp],
    }
    for name, args in sigs.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = i
    return lib


def lib():
    """Load (once) and return the CUDA library; raises NativeError when it cannot be loaded."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise NativeError(
            f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or python -m distributedtensorflow_b200._build) first")
    try:
        _lib = _declare(C.CDLL(str(LIB_PATH), mode=os.RTLD_GLOBAL if hasattr(os, "RTLD_GLOBAL") else 0))
    except OSError as e:  # pragma: no cover - depends on the box
        raise NativeError(f"cannot load {LIB_PATH}: {e}") from e
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().dtf_last_error().decode(errors="replace")
        raise NativeError(f"{what or 'native call'} failed ({rc}): {msg}")


# Selective programmatic dependent launch: while set to a stream handle, kernels enqueued on THAT
# stream are launched with the PDL attribute and kernels on any other stream without it (every
# wrapper fetches current_stream() immediately before its launch).
pdl_main_stream = None


def current_stream() -> int:
    import torch
    h = torch.cuda.current_stream().cuda_stream
    if pdl_main_stream is not None:
        lib().dtf_set_pdl(1 if h == pdl_main_stream else 0)
    return h
