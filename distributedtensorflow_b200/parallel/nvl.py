"""NVLink peer-memory parameter store — the B200 product path's data plane.

Every PS shard owns ONE device arena (``cudaMalloc`` on the shard's GPU):

    [ dtf_ps_ctrl (256 B) | var slice fp32 | slot0 | slot1 | bf16 serving copy | next variable ... ]

whose layout every task derives independently from the graph's variables + placement + optimizer
(between-graph replication: nobody ships a graph). The owning process exports a 64-byte CUDA IPC
handle through the TCP control plane (``PsShard.nvl_alloc``) and every worker maps the arena into
its own address space; from then on parameters and gradients move ONLY by loads / stores /
atomics issued from inside kernels over NVLink 5 / NVSwitch:

* pull   — ``dtf_pull_bulk`` (one launch, 16-byte peer loads) or, for the first layer, TMA
  straight from the peer-mapped bf16 serving copy into the tcgen05 GEMM's smem ring;
* push   — ``dtf_push_apply``: the worker's kernel reads its local gradient, reads the slice's
  optimizer state from the owner, applies the TF1 update and writes var / slots / serving copy
  back; the last CTA bumps ``global_step`` with a system-scope atomic.

This replaces the reference's per-tensor gRPC ``RecvTensor`` traffic: 8
parameter tensors down and 8 gradient tensors up per step per worker, through host memory,
protobuf and TCP.
"""
from __future__ import annotations

import ctypes as C
import os
import struct
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from ..ops import _lib as L
from ..train.optimizers import OptimizerSpec
from .layout import VarLayout, VarSlice, powers_shard

ALIGN = 256


def _align(n: int, a: int = ALIGN) -> int:
    return (n + a - 1) // a * a


def pad8(n: int) -> int:
    return (n + 7) // 8 * 8


@dataclass
class SliceRegion:
    """Byte offsets of one variable slice inside its shard's arena."""
    name: str
    shard: int
    sl: VarSlice
    cols: int
    rows: int            # rows of the serving copy held by this slice
    shadow_ld: int       # elements
    var_off: int
    slot_off: List[int]
    shadow_off: Optional[int]

    @property
    def var_bytes(self) -> int:
        return self.sl.n * 4

    @property
    def shadow_bytes(self) -> int:
        return self.rows * self.shadow_ld * 2


@dataclass
class ArenaLayout:
    num_shards: int
    shard_bytes: List[int]
    regions: Dict[str, List[SliceRegion]]   # variable name -> regions of its non-empty slices
    step_var: Optional[str]                 # the global_step variable (lives in ctrl.global_step)
    step_shard: int
    powers_shard: int
    num_slots: int
    # synchronous mode (csrc/kernels/sync_step.cu): per-shard gradient mailboxes and, at the SAME
    # offset in every arena, a mirror of all parameters that the owners all-gather into
    sync_workers: int = 0
    mailbox_off: Dict[Tuple[str, int], int] = field(default_factory=dict)   # (name, shard) -> slot-0 offset
    mailbox_stride: List[int] = field(default_factory=list)                 # bytes per worker slot, per shard
    mirror_f32: Dict[str, int] = field(default_factory=dict)                # name -> offset (all arenas)
    mirror_bf16: Dict[str, int] = field(default_factory=dict)
    # asynchronous "owner applies" push (csrc/kernels/mailbox.cu): the same per-worker mailboxes,
    # `mailbox_depth` buffers deep, plus one dtf_mbox_flags block per shard; no mirrors
    mailbox_depth: int = 1
    mbox_flags_off: List[int] = field(default_factory=list)
    mbox_slot_base: List[int] = field(default_factory=list)


def build_arena_layout(layout: Dict[str, VarLayout], num_shards: int, spec: Optional[OptimizerSpec],
                       trainable_names: Sequence[str], sync_workers: int = 0, mailbox_depth: int = 1,
                       mirrors: bool = True) -> ArenaLayout:
    num_slots = spec.num_slots if spec is not None else 0
    offs = [L.CTRL_BYTES] * num_shards
    regions: Dict[str, List[SliceRegion]] = {}
    step_var, step_shard = None, 0
    tset = set(trainable_names)
    for vl in sorted(layout.values(), key=lambda v: v.index):
        if vl.dtype in ("int64", "int32") and vl.num_elements == 1:
            if vl.name == "global_step" or step_var is None:
                step_var, step_shard = vl.name, vl.home_shard
            continue
        if vl.dtype != "float32":
            raise NotImplementedError(f"variable {vl.name} has dtype {vl.dtype}; the device arena "
                                      "holds float32 tensors and the integer global_step")
        want_shadow = len(vl.shape) >= 2
        cols = vl.cols
        out = []
        for sl in vl.slices:
            if sl.n == 0:
                continue
            s = sl.shard
            var_off = _align(offs[s])
            o = var_off + _align(sl.n * 4)
            slot_off = []
            if vl.name in tset:
                for _ in range(num_slots):
                    slot_off.append(o)
                    o += _align(sl.n * 4)
            rows = sl.n // cols
            ld = pad8(cols)
            shadow_off = None
            if want_shadow:
                shadow_off = o
                o += _align(rows * ld * 2)
            offs[s] = o
            out.append(SliceRegion(vl.name, s, sl, cols, rows, ld, var_off, slot_off, shadow_off))
        regions[vl.name] = out
    pshard = powers_shard(layout, list(trainable_names)) if trainable_names else 0
    al = ArenaLayout(num_shards, [], regions, step_var, step_shard, pshard, num_slots)
    if sync_workers > 0:
        al.sync_workers = sync_workers
        al.mailbox_depth = mailbox_depth
        al.mailbox_stride = [0] * num_shards
        al.mbox_flags_off = [_align(o) for o in offs]
        slot_base = [o + L.MBOX_FLAGS_BYTES for o in al.mbox_flags_off]
        al.mbox_slot_base = slot_base
        for name in trainable_names:                       # one packed slot per worker per shard
            for r in regions.get(name, []):
                al.mailbox_off[(name, r.shard)] = slot_base[r.shard] + al.mailbox_stride[r.shard]
                al.mailbox_stride[r.shard] += _align(r.sl.n * 4)
        ends = [slot_base[s] + sync_workers * mailbox_depth * al.mailbox_stride[s] for s in range(num_shards)]
        o = _align(max(ends))                              # mirrors: same offset in EVERY arena
        for name, regs in (regions.items() if mirrors else ()):
            if not regs:
                continue
            vl = layout[name]
            al.mirror_f32[name] = o
            o += _align(vl.num_elements * 4)
            if regs[0].shadow_off is not None:
                al.mirror_bf16[name] = o
                o += _align(vl.rows * regs[0].shadow_ld * 2)
        offs = [o] * num_shards
    al.shard_bytes = [_align(o, 4096) for o in offs]
    return al


@dataclass
class MirrorLayout:
    """A worker's private mirror of the parameters (``dtf_mirror_hdr`` + copies): byte offsets
    from the mirror base. Matrices get the bf16 row-padded form the tensor-core kernels read,
    vectors and small tensors an fp32 copy (a small matrix gets both)."""
    nbytes: int
    f32: Dict[str, int]
    bf16: Dict[str, int]


MIRROR_F32_MAX_ELEMS = 4096


def build_mirror_layout(layout: Dict[str, VarLayout], regions: Dict[str, List[SliceRegion]]) -> MirrorLayout:
    o = L.MIRROR_HDR_BYTES
    f32: Dict[str, int] = {}
    bf16: Dict[str, int] = {}
    for vl in sorted(layout.values(), key=lambda v: v.index):
        regs = regions.get(vl.name)
        if not regs:
            continue
        if len(vl.shape) < 2 or vl.num_elements <= MIRROR_F32_MAX_ELEMS:
            f32[vl.name] = o
            o += _align(vl.num_elements * 4)
        if regs[0].shadow_off is not None:
            bf16[vl.name] = o
            o += _align(vl.rows * regs[0].shadow_ld * 2)
    return MirrorLayout(_align(o, 4096), f32, bf16)


class NvlStore:
    """Maps every shard's arena into this process and answers "where is X" with raw pointers."""

    def __init__(self, client, layout: Dict[str, VarLayout], spec: Optional[OptimizerSpec],
                 trainable_names: Sequence[str], local_device: int,
                 shard_devices: Optional[Sequence[Optional[int]]] = None, sync_workers: int = 0,
                 mailbox_depth: int = 1, mirrors: bool = True):
        self.client = client
        self.layout = layout
        self.spec = spec
        self.num_shards = client.num_shards
        self.local_device = local_device
        self.arena = build_arena_layout(layout, self.num_shards, spec, trainable_names, sync_workers,
                                        mailbox_depth, mirrors)
        self.base: List[int] = [0] * self.num_shards
        self.same_process: List[bool] = [False] * self.num_shards
        self.mirror: Optional[MirrorLayout] = None
        self.mirror_base = 0
        self._mirror_worker: Optional[int] = None
        self.is_local: List[bool] = [False] * self.num_shards
        self.shard_device: List[int] = [local_device] * self.num_shards
        self._imported: List[int] = []
        lib = L.lib()
        for s in range(self.num_shards):
            dev = local_device
            if shard_devices is not None and shard_devices[s] is not None:
                dev = int(shard_devices[s])
            info = client.call(s, "nvl_alloc", int(self.arena.shard_bytes[s]), dev)
            self.shard_device[s] = int(info["device"])
            if int(info["pid"]) == os.getpid():
                self.base[s] = int(info["ptr"])
                self.same_process[s] = True
                if self.shard_device[s] != local_device:
                    L.check(lib.dtf_set_device(local_device), "set_device")
                    L.check(lib.dtf_enable_peer_access(self.shard_device[s]), "enable_peer_access")
            else:
                L.check(lib.dtf_set_device(local_device), "set_device")
                ptr = C.c_void_p()
                L.check(lib.dtf_ipc_import(bytes(info["handle"]), C.byref(ptr)), "ipc_import")
                self.base[s] = int(ptr.value)
                self._imported.append(self.base[s])
            self.is_local[s] = self.shard_device[s] == local_device
        L.check(lib.dtf_set_device(local_device), "set_device")
        names_per_shard: Dict[int, List[str]] = {}
        for name, regs in self.arena.regions.items():
            for r in regs:
                names_per_shard.setdefault(r.shard, []).append(name)
        if self.arena.step_var:
            names_per_shard.setdefault(self.arena.step_shard, []).append(self.arena.step_var)
        for s, names in names_per_shard.items():
            client.call(s, "declare", names)

    # -- pointers -----------------------------------------------------------------------------------
    def ctrl_ptr(self, shard: int) -> int:
        return self.base[shard]

    def var_ptr(self, r: SliceRegion) -> int:
        return self.base[r.shard] + r.var_off

    def slot_ptr(self, r: SliceRegion, i: int) -> Optional[int]:
        return self.base[r.shard] + r.slot_off[i] if i < len(r.slot_off) else None

    def shadow_ptr(self, r: SliceRegion) -> Optional[int]:
        return None if r.shadow_off is None else self.base[r.shard] + r.shadow_off

    def mbox_flags_ptr(self, shard: int) -> int:
        return self.base[shard] + self.arena.mbox_flags_off[shard]

    def mbox_slot_ptr(self, shard: int, worker: int, buf: int = 0) -> int:
        """Start of mailbox slot (worker, buffer) inside ``shard``'s arena."""
        al = self.arena
        return self.base[shard] + al.mbox_slot_base[shard] + \
            (worker * al.mailbox_depth + buf) * al.mailbox_stride[shard]

    def mbox_slice_off(self, name: str, shard: int) -> int:
        """Byte offset of a variable's slice inside every mailbox slot of ``shard``."""
        al = self.arena
        return al.mailbox_off[(name, shard)] - al.mbox_slot_base[shard]

    @property
    def step_ctrl(self) -> int:
        return self.ctrl_ptr(self.arena.step_shard)

    @property
    def powers_ctrl(self) -> int:
        return self.ctrl_ptr(self.arena.powers_shard)

    # -- host <-> arena (init, restore, checkpoint): plain synchronous copies ------------------------
    # set by the executor: host-side reads / writes of shard memory must wait for the step stream
    # (a training step reports its results to the host before its applies have finished)
    sync_hook = None

    def _h2d(self, dst: int, arr: np.ndarray) -> None:
        if self.sync_hook is not None:
            self.sync_hook()
        a = np.ascontiguousarray(arr)
        lib = L.lib()
        L.check(lib.dtf_memcpy_h2d(dst, a.ctypes.data_as(C.c_void_p), a.nbytes, None), "h2d")
        L.check(lib.dtf_stream_sync(None), "sync")

    def _d2h(self, src: int, nbytes: int, dtype) -> np.ndarray:
        if self.sync_hook is not None:
            self.sync_hook()
        out = np.empty(nbytes // np.dtype(dtype).itemsize, dtype)
        lib = L.lib()
        L.check(lib.dtf_memcpy_d2h(out.ctypes.data_as(C.c_void_p), src, nbytes, None), "d2h")
        L.check(lib.dtf_stream_sync(None), "sync")
        return out

    def write_ctrl(self, shard: int, global_step: int = 0, beta1_power: float = 0.0,
                   beta2_power: float = 0.0, initialized: int = 0) -> None:
        buf = bytearray(L.CTRL_BYTES)
        struct.pack_into("<QffI", buf, 0, int(global_step), float(beta1_power), float(beta2_power),
                         int(initialized))
        self._h2d(self.ctrl_ptr(shard), np.frombuffer(bytes(buf), np.uint8))

    def read_ctrl(self, shard: int) -> dict:
        raw = self._d2h(self.ctrl_ptr(shard), L.CTRL_BYTES, np.uint8).tobytes()
        gs, b1, b2, init, epoch, shutdown, applies = struct.unpack_from("<QffIIII", raw, 0)
        hb = struct.unpack_from("<%dQ" % L.MAX_WORKERS, raw, L.CTRL_OFF_HEARTBEAT)
        return dict(global_step=gs, beta1_power=b1, beta2_power=b2, initialized=init, epoch=epoch,
                    shutdown=shutdown, apply_count=applies, heartbeat=list(hb))

    def set_field(self, shard: int, offset: int, fmt: str, value) -> None:
        self._h2d(self.ctrl_ptr(shard) + offset, np.frombuffer(struct.pack(fmt, value), np.uint8))

    def read_global_step(self) -> int:
        raw = self._d2h(self.step_ctrl, 8, np.uint8).tobytes()
        return struct.unpack("<Q", raw)[0]

    def write_variable(self, name: str, value: np.ndarray) -> None:
        """fp32 master + refreshed bf16 serving copy, slice by slice."""
        from ..ops import kernels as K
        flat = np.ascontiguousarray(value, np.float32).reshape(-1)
        for r in self.arena.regions[name]:
            self._h2d(self.var_ptr(r), flat[r.sl.first:r.sl.first + r.sl.n])
            if r.shadow_off is not None:
                K.refresh_shadow(self.var_ptr(r), self.shadow_ptr(r), r.sl.n, 0, r.cols, r.shadow_ld)
        import torch
        torch.cuda.current_stream().synchronize()

    def write_slot(self, name: str, slot: int, value) -> None:
        for r in self.arena.regions[name]:
            if slot >= len(r.slot_off):
                continue
            if np.isscalar(value):
                arr = np.full(r.sl.n, value, np.float32)
            else:
                arr = np.ascontiguousarray(value, np.float32).reshape(-1)[r.sl.first:r.sl.first + r.sl.n]
            self._h2d(self.slot_ptr(r, slot), arr)

    def read_variable(self, name: str) -> np.ndarray:
        vl = self.layout[name]
        out = np.empty(vl.num_elements, np.float32)
        for r in self.arena.regions[name]:
            out[r.sl.first:r.sl.first + r.sl.n] = self._d2h(self.var_ptr(r), r.sl.n * 4, np.float32)
        return out.reshape(vl.shape)

    def read_slot(self, name: str, slot: int) -> Optional[np.ndarray]:
        vl = self.layout[name]
        regs = self.arena.regions[name]
        if not regs or slot >= len(regs[0].slot_off):
            return None
        out = np.empty(vl.num_elements, np.float32)
        for r in regs:
            out[r.sl.first:r.sl.first + r.sl.n] = self._d2h(self.slot_ptr(r, slot), r.sl.n * 4, np.float32)
        return out.reshape(vl.shape)

    def read_shadow(self, name: str) -> np.ndarray:
        """bf16 serving copy as float32 (tests)."""
        import torch
        vl = self.layout[name]
        rows = []
        for r in self.arena.regions[name]:
            raw = self._d2h(self.shadow_ptr(r), r.shadow_bytes, np.uint16)
            t = torch.from_numpy(raw.view(np.int16).copy()).view(torch.bfloat16)
            rows.append(t.float().numpy().reshape(r.rows, r.shadow_ld)[:, :r.cols])
        return np.concatenate(rows, 0).reshape(vl.shape)

    # -- worker mirror ("owner pushes, worker never pulls") -------------------------------------------
    def alloc_mirror(self) -> MirrorLayout:
        """Allocate this worker's mirror on its own GPU (zeroed; the header starts at 0)."""
        if self.mirror is None:
            self.mirror = build_mirror_layout(self.layout, self.arena.regions)
            lib = L.lib()
            L.check(lib.dtf_set_device(self.local_device), "set_device")
            ptr = C.c_void_p()
            L.check(lib.dtf_arena_alloc(C.byref(ptr), self.mirror.nbytes), "mirror alloc")
            self.mirror_base = int(ptr.value)
        return self.mirror

    def register_mirror(self, worker_id: int) -> None:
        """Hand the mirror to every shard's owner process (CUDA IPC handle through the control
        plane, or the plain pointer inside one process): from now on each drain of a shard also
        stores the fresh parameters and this worker's flow-control counters into it."""
        lib = L.lib()
        handle = C.create_string_buffer(L.IPC_HANDLE_BYTES)
        L.check(lib.dtf_ipc_export(self.mirror_base, handle), "ipc_export(mirror)")
        info = dict(handle=bytes(handle.raw), pid=os.getpid(), ptr=self.mirror_base,
                    device=int(self.local_device), nbytes=int(self.mirror.nbytes))
        for s in range(self.num_shards):
            self.client.call(s, "mirror_register", int(worker_id), info)
        self._mirror_worker = int(worker_id)

    def unregister_mirror(self) -> None:
        if self._mirror_worker is None:
            return
        w, self._mirror_worker = self._mirror_worker, None
        for s in range(self.num_shards):
            try:
                self.client.call(s, "mirror_unregister", w)
            except Exception:
                pass

    def mirror_hdr_u32(self, byte_off: int, count: int) -> np.ndarray:
        return self._d2h(self.mirror_base + byte_off, 4 * count, np.uint32)

    def read_mirror(self, name: str) -> Dict[str, np.ndarray]:
        """This worker's mirrored copies of a variable as float32 arrays (tests / diagnostics):
        ``{"bf16": ..., "f32": ...}`` with the forms the mirror holds."""
        import torch
        vl = self.layout[name]
        out: Dict[str, np.ndarray] = {}
        if name in self.mirror.f32:
            out["f32"] = self._d2h(self.mirror_base + self.mirror.f32[name], vl.num_elements * 4,
                                   np.float32).reshape(vl.shape)
        if name in self.mirror.bf16:
            ld = self.arena.regions[name][0].shadow_ld
            raw = self._d2h(self.mirror_base + self.mirror.bf16[name], vl.rows * ld * 2, np.uint16)
            t = torch.from_numpy(raw.view(np.int16).copy()).view(torch.bfloat16)
            out["bf16"] = t.float().numpy().reshape(vl.rows, ld)[:, :vl.cols].reshape(vl.shape)
        return out

    def close(self) -> None:
        lib = L.lib()
        self.unregister_mirror()
        if self.mirror_base:
            try:
                lib.dtf_arena_free(self.mirror_base)
            except Exception:
                pass
            self.mirror_base = 0
        for p in self._imported:
            try:
                lib.dtf_ipc_close(p)
            except Exception:
                pass
        self._imported = []
