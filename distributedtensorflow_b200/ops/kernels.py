"""Torch-tensor front ends of the hand-written sm_100a kernels.

Each function validates shapes/dtypes, then enqueues exactly one kernel of
``libdtf_b200_cuda.so`` on the current CUDA stream. ``LAUNCHES`` counts those launches (bench.py
reports it as ``gpu_launches``).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence

import torch

from . import _lib as L

LAUNCHES = 0


def _count(n: int = 1) -> None:
    global LAUNCHES
    LAUNCHES += n


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def pad8(n: int) -> int:
    return (n + 7) // 8 * 8


# ------------------------------------------------------------------------------------------------
# GEMM
# ------------------------------------------------------------------------------------------------
def gemm(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor, M: int, N: int, K: int, *,
         a_mn: bool = False, b_mn: bool = False, bias: Optional[torch.Tensor] = None,
         relu: bool = False, mask: Optional[torch.Tensor] = None, reduce: bool = False,
         alpha: float = 1.0, splits: int = 1, block_n: int = 0,
         a_ptr: Optional[int] = None, b_ptr: Optional[int] = None, lda: Optional[int] = None,
         ldb: Optional[int] = None, ws: Optional[torch.Tensor] = None,
         counters: Optional[torch.Tensor] = None, apply: Optional[dict] = None,
         scatter: Optional[dict] = None, pool: Optional[dict] = None) -> None:
    """out[M,N] = alpha * A @ B^T with the fused epilogue of ``dtf_gemm_bf16``.

    ``a``/``b`` are 2-D bf16 tensors whose row pitch (stride(0)) is a multiple of 8; with
    ``a_mn``/``b_mn`` they are stored transposed ([K,M] / [K,N]). ``a_ptr``/``b_ptr`` override
    the base address (peer-mapped PS memory) while ``lda``/``ldb`` give its pitch.
    """
    if apply is None and scatter is None and pool is None:
        assert out.dim() == 2 and out.stride(1) == 1
    if a_ptr is None:
        assert a.dtype == torch.bfloat16 and a.stride(1) == 1
        a_ptr, lda = a.data_ptr(), a.stride(0)
    if b_ptr is None:
        assert b.dtype == torch.bfloat16 and b.stride(1) == 1
        b_ptr, ldb = b.data_ptr(), b.stride(0)
    epi = L.GemmEpilogue()
    if pool is not None:
        # dX GEMM whose epilogue is conv2's max-pool + ReLU gradient (see dtf_gemm_epilogue.pool_*)
        epi.out, epi.ldo, epi.out_f32 = None, N, 0
        epi.pool_out = pool["out"].data_ptr()
        epi.pool_amax = pool["amax"].data_ptr()
        epi.pool_dbias = pool["dbias"].data_ptr()
    elif scatter is not None:
        # weight-gradient GEMM whose epilogue stores straight into the owners' mailboxes
        epi.out, epi.ldo, epi.out_f32 = None, N, 1
        epi.scatter_enable = 1
        epi.scatter_rows_per_slice = scatter["rows_per_slice"]
        epi.scatter_depth = scatter["depth"]
        epi.scatter_seq = scatter["seq"]
        for i, (base, stride) in enumerate(scatter["slices"]):
            epi.scatter_base[i] = base
            epi.scatter_buf_stride[i] = stride
    elif apply is None:
        epi.out = out.data_ptr()
        epi.ldo = out.stride(0)
        epi.out_f32 = 1 if out.dtype == torch.float32 else 0
        assert out.dtype in (torch.float32, torch.bfloat16)
    else:
        # fused push+apply epilogue: see dtf_gemm_epilogue.apply_* in include/dtf_b200.h
        epi.out, epi.ldo, epi.out_f32 = None, N, 1
        h = apply["hyper"]
        epi.apply_enable, epi.apply_opt = 1, apply["opt"]
        epi.apply_rows_per_slice = apply["rows_per_slice"]
        epi.apply_shadow_ld = apply.get("shadow_ld", N)
        epi.apply_lr, epi.apply_beta1, epi.apply_beta2 = h.lr, h.beta1, h.beta2
        epi.apply_epsilon, epi.apply_momentum, epi.apply_grad_scale = h.epsilon, h.momentum, h.grad_scale
        epi.apply_use_nesterov = h.use_nesterov
        epi.apply_ctrl_powers = apply.get("ctrl_powers")
        for i, sl in enumerate(apply["slices"]):
            epi.apply_var[i] = sl["var"]
            epi.apply_slot0[i] = sl.get("slot0")
            epi.apply_slot1[i] = sl.get("slot1")
            epi.apply_shadow[i] = sl.get("shadow")
    epi.bias = _ptr(bias)
    epi.relu = int(relu)
    epi.mask = _ptr(mask)
    epi.ldmask = mask.stride(0) if mask is not None else 0
    epi.reduce = int(reduce)
    epi.alpha = float(alpha)
    if ws is not None:
        # split-K with in-kernel fix-up: ws is a zeroed fp32 [M, ldws] view, counters zeroed u32
        assert ws.dtype == torch.float32 and ws.dim() == 2 and ws.stride(1) == 1
        epi.splitk_ws, epi.ldws = ws.data_ptr(), ws.stride(0)
        epi.tile_counters = counters.data_ptr()
    rc = L.lib().dtf_gemm_bf16(a_ptr, lda, int(a_mn), b_ptr, ldb, int(b_mn), M, N, K,
                               C.byref(epi), splits, block_n, L.current_stream())
    L.check(rc, "dtf_gemm_bf16")
    _count()


# ------------------------------------------------------------------------------------------------
# loss
# ------------------------------------------------------------------------------------------------
def softmax_xent_sparse(logits: torch.Tensor, labels: torch.Tensor, loss_accum: torch.Tensor,
                        dlogits: torch.Tensor, dbias_accum: Optional[torch.Tensor] = None,
                        num_classes: Optional[int] = None) -> None:
    B = logits.shape[0]
    Cn = num_classes or logits.shape[1]
    assert logits.dtype == torch.float32 and labels.dtype == torch.int32
    assert dlogits.dtype == torch.bfloat16
    rc = L.lib().dtf_softmax_xent_sparse(logits.data_ptr(), logits.stride(0), labels.data_ptr(),
                                         B, Cn, loss_accum.data_ptr(), dlogits.data_ptr(),
                                         dlogits.stride(0), _ptr(dbias_accum), L.current_stream())
    L.check(rc, "dtf_softmax_xent_sparse")
    _count()


def softmax_xent_dense(logits: torch.Tensor, labels: torch.Tensor, loss_accum: torch.Tensor,
                       dlogits: torch.Tensor, dbias_accum: Optional[torch.Tensor] = None,
                       num_classes: Optional[int] = None) -> None:
    B = logits.shape[0]
    Cn = num_classes or logits.shape[1]
    assert logits.dtype == torch.float32 and labels.dtype == torch.float32
    rc = L.lib().dtf_softmax_xent_dense(logits.data_ptr(), logits.stride(0), labels.data_ptr(),
                                        labels.stride(0), B, Cn, loss_accum.data_ptr(),
                                        dlogits.data_ptr(), dlogits.stride(0),
                                        _ptr(dbias_accum), L.current_stream())
    L.check(rc, "dtf_softmax_xent_dense")
    _count()


# ------------------------------------------------------------------------------------------------
# push + apply, pull
# ------------------------------------------------------------------------------------------------
def make_apply_items(entries: Sequence[dict]) -> "C.Array":
    arr = (L.ApplyItem * max(1, len(entries)))()
    for i, e in enumerate(entries):
        it = arr[i]
        it.var = e["var"]
        it.slot0 = e.get("slot0")
        it.slot1 = e.get("slot1")
        it.shadow = e.get("shadow")
        it.grad = e["grad"]
        it.n = e["n"]
        it.first = e.get("first", 0)
        it.cols = e.get("cols", 1)
        it.shadow_ld = e.get("shadow_ld", e.get("cols", 1))
        it.zero_grad = int(e.get("zero_grad", 0))
    return arr


def make_hyper(lr: float, beta1: float = 0.9, beta2: float = 0.999, epsilon: float = 1e-8,
               momentum: float = 0.0, grad_scale: float = 1.0, use_nesterov: bool = False,
               atomic_sgd: bool = False) -> L.ApplyHyper:
    h = L.ApplyHyper()
    h.lr, h.beta1, h.beta2, h.epsilon = lr, beta1, beta2, epsilon
    h.momentum, h.grad_scale = momentum, grad_scale
    h.use_nesterov, h.atomic_sgd = int(use_nesterov), int(atomic_sgd)
    return h


def step_report(ctrl_step: int, worker_id: int, bump: bool, step_out: Optional[int],
                sink: Optional[L.ResultSink]) -> None:
    rc = L.lib().dtf_step_report(ctrl_step, worker_id, int(bump), step_out,
                                 C.byref(sink) if sink is not None else None, L.current_stream())
    L.check(rc, "dtf_step_report")
    _count()


def make_sink(loss_ptr: int, host_ptr: int, seq_ptr: int, clear_loss: bool = False) -> L.ResultSink:
    sk = L.ResultSink()
    sk.loss_src, sk.host_dst, sk.seq_dev = loss_ptr, host_ptr, seq_ptr
    sk.clear_loss = int(clear_loss)
    return sk


def push_apply(opt: int, items, n_items: int, hyper: L.ApplyHyper, ctrl_powers: Optional[int],
               ctrl_step: Optional[int], finish: bool, worker_id: int,
               step_out: Optional[int], block_counter: Optional[int],
               sink: Optional[L.ResultSink] = None, max_ctas: int = 0) -> None:
    rc = L.lib().dtf_push_apply_ex(opt, items, n_items, C.byref(hyper), ctrl_powers, ctrl_step,
                                   int(finish), worker_id, step_out, block_counter,
                                   C.byref(sink) if (sink is not None and int(finish) == 1) else None,
                                   int(max_ctas), L.current_stream())
    L.check(rc, "dtf_push_apply")
    _count()


def mailbox_push(items: Sequence[tuple], depth: int, worker_id: int, owner_flags: Sequence[int],
                 publish: bool, seq: int, block_counter: Optional[int],
                 min_consumed_out: Optional[int], ctrl_step: Optional[int] = None,
                 step_out: Optional[int] = None, sink: Optional[L.ResultSink] = None,
                 my_hdr: Optional[int] = None, sync_tag: Optional[int] = None) -> None:
    """Worker side of the "owner applies" push (csrc/kernels/mailbox.cu).

    items: (src_ptr, dst0_ptr, nbytes, buf_stride[, zero_src]) — gradient slice -> its place in MY
    buffer 0 on the owning shard (zero_src: clear the source after reading it; owner: the owning
    shard, which with ``my_hdr`` turns on device-side flow control). ``publish`` makes the last CTA release the sequence number to every owner in
    ``owner_flags`` (pointers to their ``dtf_mbox_flags``)."""
    n = len(items)
    chunks = [items[i:i + L.MAX_MBOX_ITEMS] for i in range(0, n, L.MAX_MBOX_ITEMS)] or [[]]
    flags = (C.c_void_p * max(1, len(owner_flags)))(*owner_flags)
    for ci, chunk in enumerate(chunks):
        arr = (L.MboxPushItem * max(1, len(chunk)))()
        for i, item in enumerate(chunk):
            s, d, b, st = item[:4]
            arr[i].src, arr[i].dst0, arr[i].bytes, arr[i].buf_stride = s, d, b, st
            arr[i].zero_src = int(item[4]) if len(item) > 4 else 0
            arr[i].owner = int(item[5]) if len(item) > 5 else -1
        last = publish and ci == len(chunks) - 1
        if not chunk and not last:
            continue
        rc = L.lib().dtf_mailbox_push(arr, len(chunk), depth, worker_id, flags, len(owner_flags),
                                      int(last), seq, block_counter, min_consumed_out if last else None,
                                      ctrl_step if last else None, step_out if last else None,
                                      C.byref(sink) if (sink is not None and last) else None,
                                      my_hdr, sync_tag if last else None, L.current_stream())
        L.check(rc, "dtf_mailbox_push")
        _count()


def mailbox_drain_sync(opt: int, items, n_items: int, slot_stride: int, n_workers: int, replicas_to_aggregate: int,
                       hyper: L.ApplyHyper, ctrl: int, ctrl_step: Optional[int], flags: int, todo: int,
                       block_counter: int, mirror_items=None, shard_index: int = 0, max_ctas: int = 0) -> None:
    """One synchronous-round attempt on a shard (snapshot + drain): retires stale gradients, and when
    ``replicas_to_aggregate`` fresh ones are pending applies their mean once and advances the epoch."""
    rc = L.lib().dtf_mailbox_drain_sync(opt, items, n_items, slot_stride, n_workers, replicas_to_aggregate,
                                        C.byref(hyper), ctrl, ctrl_step, flags, todo, block_counter,
                                        mirror_items, shard_index, int(max_ctas), L.current_stream())
    L.check(rc, "dtf_mailbox_drain_sync")
    _count(2)


def sync_token_wait(my_hdr: int, n_shards: int, last_round: int, local_step_out: int,
                    step_out: Optional[int]) -> None:
    rc = L.lib().dtf_sync_token_wait(my_hdr, n_shards, last_round, local_step_out, step_out, L.current_stream())
    L.check(rc, "dtf_sync_token_wait")
    _count()


def make_mirror_items(entries: Sequence[tuple]) -> "C.Array":
    """entries: (bf16_off or -1, f32_off or -1, first) per drained slice."""
    arr = (L.DrainMirrorItem * max(1, len(entries)))()
    for i, (b, f, first) in enumerate(entries):
        arr[i].bf16_off, arr[i].f32_off, arr[i].first = int(b), int(f), int(first)
    return arr


def mailbox_drain(opt: int, items, n_items: int, slot_stride: int, depth: int, n_workers: int,
                  hyper: L.ApplyHyper, ctrl_powers: Optional[int], ctrl_step: Optional[int],
                  flags: int, todo: int, block_counter: int, step_out: Optional[int],
                  mirror_items=None, shard_index: int = 0, max_ctas: int = 0) -> None:
    """Owner side: claim the lock + snapshot pending mailboxes, then apply them from local HBM and
    (``mirror_items``) broadcast the new values into the registered worker mirrors. Two launches.
    ``todo`` needs MAX_WORKERS + 1 words."""
    rc = L.lib().dtf_mailbox_drain_ex(opt, items, n_items, slot_stride, depth, n_workers, C.byref(hyper),
                                      ctrl_powers, ctrl_step, flags, todo, block_counter, step_out,
                                      mirror_items, shard_index, int(max_ctas), L.current_stream())
    L.check(rc, "dtf_mailbox_drain")
    _count(2)


def pull_bulk(copies: Sequence[tuple], ctrl_step: Optional[int] = None,
              pulled_step_out: Optional[int] = None) -> None:
    """copies: (src_ptr, dst_ptr, nbytes) triples, nbytes % 16 == 0."""
    n = len(copies)
    for start in range(0, max(n, 1), L.MAX_COPY_ITEMS):
        chunk = copies[start:start + L.MAX_COPY_ITEMS]
        arr = (L.CopyItem * max(1, len(chunk)))()
        for i, (s, d, b) in enumerate(chunk):
            arr[i].src, arr[i].dst, arr[i].bytes = s, d, b
        first = start == 0
        if not chunk and not (first and pulled_step_out):
            break
        rc = L.lib().dtf_pull_bulk(arr, len(chunk), ctrl_step if first else None,
                                   pulled_step_out if first else None, L.current_stream())
        L.check(rc, "dtf_pull_bulk")
        _count()


def refresh_shadow(var_ptr: int, shadow_ptr: int, n: int, first: int, cols: int,
                   shadow_ld: int) -> None:
    rc = L.lib().dtf_refresh_shadow(var_ptr, shadow_ptr, n, first, cols, shadow_ld,
                                    L.current_stream())
    L.check(rc, "dtf_refresh_shadow")
    _count()


# ------------------------------------------------------------------------------------------------
# elementwise helpers
# ------------------------------------------------------------------------------------------------
def u8_to_bf16(src: torch.Tensor, dst: torch.Tensor, scale: float = 1.0 / 255.0) -> None:
    rows, cols = src.shape
    rc = L.lib().dtf_u8_to_bf16(src.data_ptr(), dst.data_ptr(), rows, cols, dst.stride(0), scale,
                                L.current_stream())
    L.check(rc, "dtf_u8_to_bf16")
    _count()


def f32_to_bf16(src: torch.Tensor, dst: torch.Tensor) -> None:
    rows, cols = src.shape
    rc = L.lib().dtf_f32_to_bf16(src.data_ptr(), src.stride(0), dst.data_ptr(), dst.stride(0),
                                 rows, cols, L.current_stream())
    L.check(rc, "dtf_f32_to_bf16")
    _count()


def head_fused(act: torch.Tensor, H: int, w_ptr: int, ldw: int, bias_ptr: Optional[int],
               labels: torch.Tensor, B: int, num_classes: int, logits: torch.Tensor, loss_accum: torch.Tensor,
               dlogits: Optional[torch.Tensor] = None, dbias_accum: Optional[torch.Tensor] = None,
               dact: Optional[torch.Tensor] = None, dprev_bias_accum: Optional[torch.Tensor] = None,
               dw_accum: Optional[torch.Tensor] = None) -> None:
    """Classifier head in one launch (csrc/kernels/head_fused.cu): logits + loss (+ dlogits, bias
    gradients and the previous layer's masked activation gradient when training)."""
    sparse = labels.dtype == torch.int32
    rc = L.lib().dtf_head_fused(act.data_ptr(), act.stride(0), H, w_ptr, ldw, bias_ptr,
                                labels.data_ptr() if sparse else None,
                                None if sparse else labels.data_ptr(), 0 if sparse else labels.stride(0),
                                B, num_classes, logits.data_ptr(), logits.stride(0), loss_accum.data_ptr(),
                                _ptr(dlogits), dlogits.stride(0) if dlogits is not None else 0,
                                _ptr(dbias_accum), _ptr(dact), dact.stride(0) if dact is not None else 0,
                                _ptr(dprev_bias_accum), _ptr(dw_accum), L.current_stream())
    L.check(rc, "dtf_head_fused")
    _count()


def colsum_bf16(x: torch.Tensor, rows: int, cols: int, out_accum: torch.Tensor) -> None:
    rc = L.lib().dtf_colsum_bf16(x.data_ptr(), x.stride(0), rows, cols, out_accum.data_ptr(),
                                 L.current_stream())
    L.check(rc, "dtf_colsum_bf16")
    _count()


def fill_u32(t: torch.Tensor, value: int = 0) -> None:
    nbytes = t.numel() * t.element_size()
    assert nbytes % 4 == 0
    rc = L.lib().dtf_fill_u32(t.data_ptr(), value, nbytes // 4, L.current_stream())
    L.check(rc, "dtf_fill_u32")
    _count()


# ------------------------------------------------------------------------------------------------
# CNN layers
# ------------------------------------------------------------------------------------------------
def conv1_fwd(x: torch.Tensor, w1: torch.Tensor, b1: torch.Tensor, pool1p: torch.Tensor,
              amax1: torch.Tensor, scale: float, x_copy: Optional[torch.Tensor] = None,
              ingest: bool = False, y_src: Optional[int] = None, y_dst: Optional[int] = None,
              y_bytes: int = 0) -> None:
    """``ingest``: x is pinned host memory read in place by the kernel (and mirrored to x_copy)."""
    B = x.shape[0]
    u8 = x.dtype == torch.uint8
    if ingest:
        cp = _ptr(x_copy)
        rc = L.lib().dtf_conv1_fwd_ingest(x.data_ptr() if u8 else None, None if u8 else x.data_ptr(), scale,
                                          w1.data_ptr(), b1.data_ptr(), pool1p.data_ptr(), amax1.data_ptr(),
                                          cp if u8 else None, None if u8 else cp, y_src, y_dst, y_bytes, B,
                                          L.current_stream())
        L.check(rc, "dtf_conv1_fwd_ingest")
        _count()
        return
    rc = L.lib().dtf_conv1_fwd(x.data_ptr() if u8 else None, None if u8 else x.data_ptr(), scale,
                               w1.data_ptr(), b1.data_ptr(), pool1p.data_ptr(), amax1.data_ptr(),
                               B, L.current_stream())
    L.check(rc, "dtf_conv1_fwd")
    _count()


def conv1_bwd(x: torch.Tensor, dpool1: torch.Tensor, pool1p: torch.Tensor, amax1: torch.Tensor,
              dw1_accum: torch.Tensor, db1_accum: torch.Tensor, scale: float) -> None:
    B = x.shape[0]
    u8 = x.dtype == torch.uint8
    rc = L.lib().dtf_conv1_bwd(x.data_ptr() if u8 else None, None if u8 else x.data_ptr(), scale,
                               dpool1.data_ptr(), pool1p.data_ptr(), amax1.data_ptr(),
                               dw1_accum.data_ptr(), db1_accum.data_ptr(), B, L.current_stream())
    L.check(rc, "dtf_conv1_bwd")
    _count()


CONV2_HALO = os.environ.get("DTF_B200_CONV2_HALO", "1") != "0"


def conv2_fwd(pool1p: torch.Tensor, w2_ptr: int, b2: torch.Tensor, pool2: torch.Tensor,
              amax2: torch.Tensor, halo: Optional[bool] = None) -> None:
    B = pool1p.shape[0]
    fn = L.lib().dtf_conv2_fwd_halo if (CONV2_HALO if halo is None else halo) else L.lib().dtf_conv2_fwd
    rc = fn(pool1p.data_ptr(), w2_ptr, b2.data_ptr(), pool2.data_ptr(),
                               amax2.data_ptr(), B, L.current_stream())
    L.check(rc, "dtf_conv2_fwd")
    _count()


def pool2_bwd(dpool2: torch.Tensor, pool2: torch.Tensor, amax2: torch.Tensor,
              dconv2p: torch.Tensor, db2_accum: torch.Tensor) -> None:
    B = pool2.shape[0]
    rc = L.lib().dtf_pool2_bwd(dpool2.data_ptr(), dpool2.stride(0), pool2.data_ptr(),
                               amax2.data_ptr(), dconv2p.data_ptr(), db2_accum.data_ptr(), B,
                               L.current_stream())
    L.check(rc, "dtf_pool2_bwd")
    _count()


def conv2_wgrad(pool1p: torch.Tensor, dconv2p: torch.Tensor, dw2_accum: torch.Tensor) -> None:
    B = pool1p.shape[0]
    rc = L.lib().dtf_conv2_wgrad(pool1p.data_ptr(), dconv2p.data_ptr(), dw2_accum.data_ptr(), B,
                                 L.current_stream())
    L.check(rc, "dtf_conv2_wgrad")
    _count()


def conv2_dgrad(dconv2p: torch.Tensor, w2_ptr: int, dpool1: torch.Tensor, halo: Optional[bool] = None) -> None:
    B = dconv2p.shape[0]
    fn = L.lib().dtf_conv2_dgrad_halo if (CONV2_HALO if halo is None else halo) else L.lib().dtf_conv2_dgrad
    rc = fn(dconv2p.data_ptr(), w2_ptr, dpool1.data_ptr(), B,
                                 L.current_stream())
    L.check(rc, "dtf_conv2_dgrad")
    _count()
