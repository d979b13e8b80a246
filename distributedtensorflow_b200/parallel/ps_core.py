"""Host-side state machine of ONE parameter-server task (CPU backends).

This is what lives behind ``server.join()`` in a ``--job_name ps`` process
(run_mnist_distributed.py:100-103) when the cluster runs on CPUs: the variables (slices) placed
on this task, their optimizer slots, Adam's beta powers, the shared ``global_step``, and — for
synchronous training — the ``ConditionalAccumulator`` per variable and the token ``FIFOQueue`` of
``SyncReplicasOptimizer`` (templates/00_mnist_replica.py:174-191). State survives worker exits and
restarts; it is lost when the PS dies unless a checkpoint was written (TF1 semantics, T2).

Concurrency: apply is intentionally unlocked (``use_locking=False`` Hogwild) unless the optimizer
spec asks for locking; accumulators and queues are properly synchronised because their TF
counterparts are.
"""
from __future__ import annotations

import threading
import time
from collections import deque
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np
import torch

from ..train.optimizers import Optimizer, OptimizerSpec

_TORCH_DTYPES = {"float32": torch.float32, "int32": torch.int32, "int64": torch.int64}


class AbortedError(RuntimeError):
    """A PS/worker task went away (tf.errors.AbortedError / UnavailableError equivalent)."""


class OutOfRangeError(RuntimeError):
    """A closed queue was dequeued (tf.errors.OutOfRangeError)."""


class ConditionalAccumulator:
    """tf.ConditionalAccumulator: sums gradients whose ``local_step >= global_step`` (stale ones
    are silently dropped); ``take_grad(n)`` blocks until ``n`` were applied, returns their MEAN,
    resets the sum and advances the accumulator's global_step by one."""

    def __init__(self, shape, dtype=torch.float32):
        self._sum = torch.zeros(shape, dtype=dtype)
        self._count = 0
        self._global_step = 0
        self._cv = threading.Condition()
        self.dropped = 0

    def apply_grad(self, grad: torch.Tensor, local_step: int) -> bool:
        with self._cv:
            if local_step < self._global_step:
                self.dropped += 1
                return False
            self._sum.add_(grad.reshape(self._sum.shape))
            self._count += 1
            self._cv.notify_all()
            return True

    def take_grad(self, num_required: int, timeout: Optional[float] = None) -> torch.Tensor:
        deadline = None if timeout is None else time.monotonic() + timeout
        with self._cv:
            while self._count < num_required:
                remaining = None if deadline is None else deadline - time.monotonic()
                if remaining is not None and remaining <= 0:
                    raise TimeoutError("take_grad timed out")
                self._cv.wait(remaining if remaining is not None else 1.0)
            avg = self._sum / float(self._count)
            self._sum.zero_()
            self._count = 0
            self._global_step += 1
            return avg

    def wait_ready(self, num_required: int, deadline: Optional[float]) -> None:
        """Block until ``num_required`` gradients are accumulated WITHOUT consuming them."""
        with self._cv:
            while self._count < num_required:
                remaining = None if deadline is None else deadline - time.monotonic()
                if remaining is not None and remaining <= 0:
                    raise TimeoutError("take_grad timed out")
                self._cv.wait(remaining if remaining is not None else 1.0)

    def set_global_step(self, new_step: int) -> None:
        with self._cv:
            if new_step > self._global_step:
                self._global_step = new_step

    def num_accumulated(self) -> int:
        with self._cv:
            return self._count

    @property
    def global_step(self) -> int:
        return self._global_step


class FIFOQueue:
    """Unbounded scalar FIFO with blocking dequeue (the ``sync_token_q``)."""

    def __init__(self):
        self._q = deque()
        self._cv = threading.Condition()
        self._closed = False

    def enqueue_many(self, values: Iterable[int]) -> None:
        with self._cv:
            self._q.extend(values)
            self._cv.notify_all()

    def dequeue(self, timeout: Optional[float] = None) -> int:
        deadline = None if timeout is None else time.monotonic() + timeout
        with self._cv:
            while not self._q:
                if self._closed:
                    raise OutOfRangeError("queue is closed and empty")
                remaining = None if deadline is None else deadline - time.monotonic()
                if remaining is not None and remaining <= 0:
                    raise TimeoutError("dequeue timed out")
                self._cv.wait(remaining if remaining is not None else 1.0)
            return self._q.popleft()

    def size(self) -> int:
        with self._cv:
            return len(self._q)

    def close(self) -> None:
        with self._cv:
            self._closed = True
            self._cv.notify_all()


class PsShard:
    """All state of PS task ``task_index``. Methods are the RPC surface."""

    def __init__(self, task_index: int):
        self.task_index = task_index
        self.vars: Dict[str, torch.Tensor] = {}          # slice tensors (flat 1-D)
        self.meta: Dict[str, dict] = {}                  # name -> {first, n, dtype}
        self.slots: Dict[str, List[torch.Tensor]] = {}
        self.slot_spec: Optional[OptimizerSpec] = None
        self.powers: Dict[str, torch.Tensor] = {}
        self.initialized = False
        self.accumulators: Dict[str, ConditionalAccumulator] = {}
        self.queues: Dict[str, FIFOQueue] = {}
        self.heartbeats: Dict[int, Tuple[int, float]] = {}
        self.apply_count = 0
        self.shutdown_requested = threading.Event()
        self._lock = threading.Lock()        # structure changes + locked applies
        self._step_lock = threading.Lock()   # AssignAdd on global_step is atomic in TF too
        self.incarnation = int(time.time() * 1e6)
        self.declared: set = set()           # names whose storage lives in a device arena
        self.device: Optional[str] = None    # "cuda:N" when this PS task owns a GPU arena
        self._arena: Optional[dict] = None
        # owner-applies data plane (device arenas): the apply daemon, where its flag block lives, the
        # workers' mirrors mapped into this process, the pinned staging buffer of the parallel save
        self._mbox_daemon = None
        self._mbox_sig: Optional[str] = None
        self._mbox_flags_off: Optional[int] = None
        self._mbox_backstop_us = 0
        self._mirrors: Dict[int, dict] = {}
        self._save_stage: Optional[int] = None

    # -- creation / init --------------------------------------------------------------------------
    def create(self, defs: Sequence[dict], spec: Optional[dict]) -> bool:
        """defs: [{name, first, n, dtype}] slices owned by this shard. Idempotent: re-creating an
        existing variable keeps its value (PS state outlives workers)."""
        with self._lock:
            for d in defs:
                name = d["name"]
                if name not in self.vars:
                    self.vars[name] = torch.zeros(int(d["n"]), dtype=_TORCH_DTYPES[d["dtype"]])
                    self.meta[name] = dict(d)
            if spec is not None:
                ospec = _spec_from(spec)
                self.slot_spec = ospec
                for d in defs:
                    name = d["name"]
                    if not d.get("trainable", True) or d["dtype"] != "float32":
                        continue
                    if name not in self.slots or len(self.slots[name]) != ospec.num_slots:
                        self.slots[name] = [
                            torch.full((int(d["n"]),), ospec.slot_init(i), dtype=torch.float32)
                            for i in range(ospec.num_slots)]
                if ospec.kind == "adam" and spec.get("_owns_powers"):
                    if "beta1_power" not in self.powers:
                        self.powers["beta1_power"] = torch.tensor(ospec.beta1, dtype=torch.float32)
                        self.powers["beta2_power"] = torch.tensor(ospec.beta2, dtype=torch.float32)
        return True

    def assign(self, name: str, value: np.ndarray) -> None:
        t = self.vars[name]
        with torch.no_grad():
            t.copy_(torch.from_numpy(np.ascontiguousarray(value)).reshape(-1).to(t.dtype))

    def assign_slot(self, name: str, slot: int, value: np.ndarray) -> None:
        with torch.no_grad():
            self.slots[name][slot].copy_(torch.from_numpy(np.ascontiguousarray(value)).reshape(-1))

    def assign_power(self, name: str, value: float) -> None:
        self.powers[name] = torch.tensor(float(value), dtype=torch.float32)

    def reset_slots(self) -> None:
        spec = self.slot_spec
        if spec is None:
            return
        for name, sl in self.slots.items():
            for i, s in enumerate(sl):
                s.fill_(spec.slot_init(i))
        if spec.kind == "adam" and self.powers:
            self.powers["beta1_power"].fill_(spec.beta1)
            self.powers["beta2_power"].fill_(spec.beta2)

    def mark_initialized(self, value: bool = True) -> None:
        self.initialized = value

    def is_initialized(self, names: Optional[Sequence[str]] = None) -> bool:
        if not self.initialized:
            return False
        return all(n in self.vars for n in (names or ()))

    def uninitialized(self, names: Sequence[str]) -> List[str]:
        """report_uninitialized_variables for the slices this shard should hold."""
        if not self.initialized:
            return list(names)
        return [n for n in names if n not in self.vars and n not in self.declared]

    # -- device arena (B200 path): the shard's variables live in GPU memory that workers map ----
    def declare(self, names: Sequence[str]) -> None:
        self.declared.update(names)

    def nvl_alloc(self, nbytes: int, device_index: int) -> dict:
        """Allocate (once) this shard's arena on ``cuda:device_index`` and return what a peer
        needs to map it: a 64-byte CUDA IPC handle, plus pid/pointer for same-process users."""
        import os
        from ..ops import _lib as L
        with self._lock:
            if self._arena is not None and self._arena["nbytes"] >= nbytes:
                return self._arena
            import ctypes as C
            lib = L.lib()
            L.check(lib.dtf_set_device(device_index), "set_device")
            ptr = C.c_void_p()
            L.check(lib.dtf_arena_alloc(C.byref(ptr), nbytes), "arena_alloc")
            handle = C.create_string_buffer(L.IPC_HANDLE_BYTES)
            L.check(lib.dtf_ipc_export(ptr, handle), "ipc_export")
            self._arena = dict(handle=bytes(handle.raw), pid=os.getpid(), ptr=int(ptr.value),
                               device=int(device_index), nbytes=int(nbytes))
            self.initialized = False
            return self._arena

    def nvl_info(self) -> Optional[dict]:
        return self._arena

    def save_arena_ops(self, ops: Sequence[Sequence]) -> List[int]:
        """Checkpoint, PS side: this task copies the byte ranges of ITS arena named by ``ops``
        (``[path, file_offset, arena_offset, nbytes]``) to the host and stores them straight into the
        (pre-sized) checkpoint data files, all shards doing so in parallel; returns the unmasked
        crc32c of every range, from which the chief - which never sees the bytes - derives each
        tensor's checksum (``dtf_crc32c_combine``) for the index. Reference: the sharded saver of
        templates/00_between_graph_replication_async_mnist.py:43-45, where every PS runs its own SaveV2."""
        import ctypes as C
        import os
        from ..ops import _lib as L
        from ..utils import hostlib
        if self._arena is None:
            raise RuntimeError("save_arena_ops before nvl_alloc")
        lib, hl = L.lib(), hostlib.lib()
        L.check(lib.dtf_set_device(self._arena["device"]), "set_device")
        base = self._arena["ptr"]
        chunk = 32 << 20
        stage = self._save_stage
        if stage is None:
            p = C.c_void_p()
            L.check(lib.dtf_host_alloc_pinned(C.byref(p), chunk), "pinned staging buffer")
            stage = self._save_stage = int(p.value)
        view = (C.c_ubyte * chunk).from_address(stage)
        crcs: List[int] = []
        fds: Dict[str, int] = {}
        try:
            for path, foff, aoff, nbytes in ops:
                fd = fds.get(path)
                if fd is None:
                    fd = fds[path] = os.open(path, os.O_WRONLY)
                crc, done = 0, 0
                while done < nbytes:
                    n = min(chunk, int(nbytes) - done)
                    L.check(lib.dtf_memcpy_d2h(stage, base + int(aoff) + done, n, None), "d2h")
                    L.check(lib.dtf_stream_sync(None), "sync")
                    crc = int(hl.dtf_crc32c_extend(crc, stage, n))
                    mv = memoryview(view)[:n]
                    wrote = 0
                    while wrote < n:
                        wrote += os.pwrite(fd, mv[wrote:], int(foff) + done + wrote)
                    done += n
                crcs.append(crc)
        finally:
            for fd in fds.values():
                os.close(fd)
        return crcs

    # -- apply daemon ("owner applies" push, csrc/runtime/mbox_daemon.cu) ---------------------------
    def mbox_start(self, desc: dict) -> dict:
        """Start (once) the native apply daemon on this shard's arena. ``desc`` describes the
        shard's trainable slices with byte offsets from the arena base — every task derives the
        same layout, so whichever worker calls first wins and later calls are checked no-ops.
        This is the PS doing what the reference's PS does: execute the optimizer's apply ops that
        replica_device_setter placed on /job:ps (run_mnist_distributed.py:107,116)."""
        import atexit
        import ctypes as C
        from ..ops import _lib as L
        from ..ops import kernels as K
        sig = repr(sorted((k, repr(v)) for k, v in desc.items()))
        with self._lock:
            if self._arena is None:
                raise RuntimeError("mbox_start before nvl_alloc")
            base = self._arena["ptr"]
            sig = sig + "@%x" % base
            old = self._mbox_daemon
            if old is not None:
                if sig == self._mbox_sig:
                    return {"started": False}
                # a new graph / optimizer / arena on this PS task (new session in the same process):
                # retire the daemon that serves the previous layout, and forget the mirrors that were
                # registered for it (their layout belongs to the old graph)
                self._mbox_daemon = None
                L.lib().dtf_mbox_daemon_stop(old)
                self._drop_mirrors()

            def at(off):
                return None if off is None or off < 0 else base + int(off)

            self._mbox_flags_off = int(desc["flags_off"])
            # a (re)started daemon begins from a clean flag block: after a layout change that
            # region of the arena held something else (a stale lock or mirror address would be fatal)
            L.check(L.lib().dtf_set_device(self._arena["device"]), "set_device")
            zeros = (C.c_ubyte * L.MBOX_FLAGS_BYTES)()
            L.check(L.lib().dtf_memcpy_h2d(self._flags_ptr(), zeros, L.MBOX_FLAGS_BYTES, None), "h2d")
            L.check(L.lib().dtf_stream_sync(None), "sync")
            self._mirrors.clear()
            mirror_items = None
            if desc.get("mirror_items") is not None:
                mirror_items = K.make_mirror_items([tuple(m) for m in desc["mirror_items"]])
            entries = [dict(var=at(it["var_off"]), slot0=at(it["slot0_off"]), slot1=at(it["slot1_off"]),
                            shadow=at(it["shadow_off"]), grad=at(it["grad_off"]), n=int(it["n"]),
                            first=0, cols=int(it["cols"]), shadow_ld=int(it["shadow_ld"]))
                       for it in desc["items"]]
            items = K.make_apply_items(entries)
            h = desc["hyper"]
            hyper = K.make_hyper(h["lr"], h["beta1"], h["beta2"], h["epsilon"], h["momentum"], 1.0,
                                 bool(h["use_nesterov"]), False)
            handle = C.c_void_p()
            lib = L.lib()
            rc = lib.dtf_mbox_daemon_start(C.byref(handle), self._arena["device"], int(desc["opt"]), items,
                                           len(entries), int(desc["slot_stride"]), int(desc["depth"]),
                                           int(desc["n_workers"]), C.byref(hyper), base,
                                           base if desc["owns_step"] else None,
                                           base + int(desc["flags_off"]), int(desc.get("poll_us", 20)),
                                           mirror_items, int(desc.get("shard", self.task_index)),
                                           int(self._mbox_backstop_us),
                                           int(desc.get("sync_R", 0) or 0))
            L.check(rc, "dtf_mbox_daemon_start")
            self._mbox_daemon = handle
            self._mbox_sig = sig
            atexit.register(self.mbox_stop)
            return {"started": True}

    def mbox_stats(self) -> dict:
        import ctypes as C
        from ..ops import _lib as L
        polls, drains, applied = C.c_ulonglong(), C.c_ulonglong(), C.c_ulonglong()
        with self._lock:                       # not while mbox_stop frees the handle
            d = self._mbox_daemon
            if d is None:
                return {"running": False}
            rc = L.lib().dtf_mbox_daemon_stats(d, C.byref(polls), C.byref(drains), C.byref(applied))
        L.check(rc, "apply daemon")
        return {"running": True, "polls": polls.value, "drains": drains.value, "applied": applied.value}

    def mbox_pause(self, pause: bool) -> None:
        """Quiesce the shard for a consistent read (checkpoint): parks the daemon, makes drains
        launched from a colocated worker's step graph no-ops (``flags.paused``) and waits for a
        drain in flight to retire."""
        from ..ops import _lib as L
        d = self._mbox_daemon
        if d is not None:
            L.check(L.lib().dtf_mbox_daemon_pause(d, 1 if pause else 0), "dtf_mbox_daemon_pause")
        if self._arena is not None and self._mbox_flags_off is not None:
            self._flags_write_u32(L.MBOX_OFF_PAUSED, 1 if pause else 0)
            if pause:
                self._wait_drain_idle()

    def mbox_attach(self, worker: int, attach: bool, backstop_us: int = 400) -> None:
        """A worker living in this process runs this shard's drains as nodes of its own step graph
        (device-driven: no host thread between a gradient's arrival and its apply). The daemon
        stays as the backstop for gradients that wait longer than ``backstop_us`` (that worker has
        stopped stepping); detaching returns the shard to the daemon alone."""
        from ..ops import _lib as L
        with self._lock:
            self._mbox_backstop_us = int(backstop_us) if attach else 0
            d = self._mbox_daemon
            if d is not None:
                L.check(L.lib().dtf_mbox_daemon_set_backstop(d, self._mbox_backstop_us), "set_backstop")

    # -- worker mirrors (csrc/kernels/mailbox.cu: the drain broadcasts fresh parameters) ------------
    def _flags_ptr(self) -> int:
        return self._arena["ptr"] + self._mbox_flags_off

    def _flags_write_u32(self, off: int, value: int) -> None:
        import ctypes as C
        from ..ops import _lib as L
        lib = L.lib()
        L.check(lib.dtf_set_device(self._arena["device"]), "set_device")
        buf = (C.c_uint * 1)(int(value))
        L.check(lib.dtf_memcpy_h2d(self._flags_ptr() + off, buf, 4, None), "h2d")
        L.check(lib.dtf_stream_sync(None), "sync")

    def _flags_read_u32(self, off: int) -> int:
        import ctypes as C
        from ..ops import _lib as L
        lib = L.lib()
        L.check(lib.dtf_set_device(self._arena["device"]), "set_device")
        buf = (C.c_uint * 1)(0)
        L.check(lib.dtf_memcpy_d2h(buf, self._flags_ptr() + off, 4, None), "d2h")
        L.check(lib.dtf_stream_sync(None), "sync")
        return int(buf[0])

    def _wait_drain_idle(self, timeout: float = 2.0) -> None:
        from ..ops import _lib as L
        t0 = time.monotonic()
        while self._flags_read_u32(L.MBOX_OFF_LOCK) != 0 and time.monotonic() - t0 < timeout:
            time.sleep(50e-6)

    def _drop_mirrors(self) -> None:
        """Clear every registered mirror address (caller holds the lock, no drain is running)."""
        import ctypes as C
        from ..ops import _lib as L
        mirrors = self._mirrors
        if self._arena is None or self._mbox_flags_off is None:
            mirrors.clear()
            return
        lib = L.lib()
        L.check(lib.dtf_set_device(self._arena["device"]), "set_device")
        zeros = (C.c_ulonglong * L.MAX_WORKERS)()
        L.check(lib.dtf_memcpy_h2d(self._flags_ptr() + L.MBOX_OFF_MIRROR_BASE, zeros, 8 * L.MAX_WORKERS, None), "h2d")
        L.check(lib.dtf_stream_sync(None), "sync")
        for ent in mirrors.values():
            if ent["imported"]:
                lib.dtf_ipc_close(ent["ptr"])
        mirrors.clear()

    def mirror_register(self, worker: int, info: dict) -> dict:
        """Map worker ``worker``'s mirror into this (the owner's) process and publish its address in
        ``flags.mirror_base[worker]``; every later drain stores the fresh parameters there."""
        import ctypes as C
        import os
        from ..ops import _lib as L
        if self._arena is None or self._mbox_flags_off is None:
            raise RuntimeError("mirror_register before mbox_start")
        if not 0 <= int(worker) < L.MAX_WORKERS:
            raise ValueError("mirror_register: worker %r out of range" % (worker,))
        lib = L.lib()
        with self._lock:
            mirrors = self._mirrors
            L.check(lib.dtf_set_device(self._arena["device"]), "set_device")
            imported = False
            if int(info["pid"]) == os.getpid():
                ptr = int(info["ptr"])
                if int(info["device"]) != self._arena["device"]:
                    L.check(lib.dtf_enable_peer_access(int(info["device"])), "enable_peer_access")
            else:
                p = C.c_void_p()
                L.check(lib.dtf_ipc_import(bytes(info["handle"]), C.byref(p)), "ipc_import(mirror)")
                ptr, imported = int(p.value), True
            old = mirrors.get(int(worker))
            buf = (C.c_ulonglong * 1)(ptr)
            L.check(lib.dtf_memcpy_h2d(self._flags_ptr() + L.MBOX_OFF_MIRROR_BASE + 8 * int(worker), buf, 8, None), "h2d")
            L.check(lib.dtf_stream_sync(None), "sync")
            mirrors[int(worker)] = dict(ptr=ptr, imported=imported)
        if old is not None and old["imported"] and old["ptr"] != ptr:
            self._wait_drain_idle()
            lib.dtf_ipc_close(old["ptr"])
        return {"ok": True}

    def mirror_unregister(self, worker: int) -> None:
        import ctypes as C
        from ..ops import _lib as L
        lib = L.lib()
        with self._lock:
            ent = self._mirrors.pop(int(worker), None)
            if ent is None or self._arena is None:
                return
            L.check(lib.dtf_set_device(self._arena["device"]), "set_device")
            buf = (C.c_ulonglong * 1)(0)
            L.check(lib.dtf_memcpy_h2d(self._flags_ptr() + L.MBOX_OFF_MIRROR_BASE + 8 * int(worker), buf, 8, None), "h2d")
            L.check(lib.dtf_stream_sync(None), "sync")
        # a drain that read the old address may still be storing: let it retire before unmapping
        time.sleep(200e-6)
        self._wait_drain_idle()
        if ent["imported"]:
            lib.dtf_ipc_close(ent["ptr"])

    def mbox_stop(self) -> None:
        from ..ops import _lib as L
        with self._lock:
            d = self._mbox_daemon
            self._mbox_daemon = None
        if d is not None:
            L.lib().dtf_mbox_daemon_stop(d)

    # -- data plane ---------------------------------------------------------------------------------
    def pull(self, names: Sequence[str]) -> List[np.ndarray]:
        # Unlocked reads: a pull can observe a half-applied update, exactly like TF's PS.
        return [self.vars[n].numpy().copy() for n in names]

    def read_powers(self) -> Dict[str, float]:
        return {k: float(v) for k, v in self.powers.items()}

    def push_apply(self, spec: dict, grads: Dict[str, np.ndarray],
                   powers: Optional[Dict[str, float]]) -> None:
        ospec = _spec_from(spec)
        pw = None
        if ospec.kind == "adam":
            pw = powers if powers is not None else self.powers
        ctx = self._lock if ospec.use_locking else _NullCtx()
        with ctx:
            for name, g in grads.items():
                gt = torch.from_numpy(np.ascontiguousarray(g)).reshape(-1)
                Optimizer.apply_dense(ospec, self.vars[name], gt, self.slots.get(name, ()), pw)
        self.apply_count += 1

    def finish(self, spec: dict) -> None:
        ospec = _spec_from(spec)
        if self.powers:
            Optimizer.finish(ospec, self.powers)

    def step_add(self, name: str, delta: int = 1) -> int:
        with self._step_lock:
            t = self.vars[name]
            t.add_(delta)
            return int(t.item())

    def read_scalar(self, name: str) -> int:
        return int(self.vars[name].item())

    def ping(self) -> dict:
        return {"task": self.task_index, "incarnation": self.incarnation}

    def heartbeat(self, worker: int, step: int) -> None:
        self.heartbeats[worker] = (step, time.time())

    # -- sync-replicas objects --------------------------------------------------------------------
    def accum_create(self, names_shapes: Sequence[Tuple[str, int]]) -> None:
        with self._lock:
            for name, n in names_shapes:
                if name not in self.accumulators:
                    self.accumulators[name] = ConditionalAccumulator((int(n),))

    def accum_apply(self, grads: Dict[str, np.ndarray], local_step: int) -> int:
        ok = 0
        for name, g in grads.items():
            ok += bool(self.accumulators[name].apply_grad(torch.from_numpy(np.ascontiguousarray(g)), local_step))
        return ok

    def accum_take(self, names: Sequence[str], num_required: int,
                   timeout: Optional[float] = None) -> List[np.ndarray]:
        return [self.accumulators[n].take_grad(num_required, timeout).numpy() for n in names]

    def accum_take_apply(self, names: Sequence[str], num_required: int, spec: dict,
                         powers: Optional[Dict[str, float]], timeout: Optional[float] = None) -> int:
        """The chief's sync_op as it executes ON this PS: block until ``num_required`` fresh
        gradients were accumulated for every listed variable, apply their mean once."""
        # all-or-nothing for the shard: first wait (consuming nothing) until EVERY listed accumulator
        # holds enough gradients, then take them. A timeout half-way can therefore never discard the
        # means already taken and advance only some accumulators' step.
        deadline = None if timeout is None else time.monotonic() + timeout
        for n in names:
            self.accumulators[n].wait_ready(num_required, deadline)
        grads = {n: self.accumulators[n].take_grad(num_required, None).numpy() for n in names}
        self.push_apply(spec, grads, powers)
        return len(grads)

    def accum_set_global_step(self, step: int) -> None:
        for a in self.accumulators.values():
            a.set_global_step(step)

    def accum_stats(self) -> Dict[str, Tuple[int, int, int]]:
        return {n: (a.num_accumulated(), a.global_step, a.dropped) for n, a in self.accumulators.items()}

    def queue_create(self, name: str) -> None:
        with self._lock:
            self.queues.setdefault(name, FIFOQueue())

    def queue_enqueue_many(self, name: str, values: Sequence[int]) -> None:
        self.queues[name].enqueue_many(values)

    def queue_dequeue(self, name: str, timeout: Optional[float] = None) -> int:
        return self.queues[name].dequeue(timeout)

    def queue_size(self, name: str) -> int:
        return self.queues[name].size()

    def queue_close(self, name: str) -> None:
        if name in self.queues:
            self.queues[name].close()

    # -- checkpoint support ---------------------------------------------------------------------------
    def snapshot(self) -> Dict[str, np.ndarray]:
        """Every tensor this shard would write into its checkpoint data shard."""
        out = {n: t.numpy().copy() for n, t in self.vars.items()}
        spec = self.slot_spec
        if spec is not None:
            for n, sl in self.slots.items():
                for sname, s in zip(spec.slot_names(), sl):
                    out[f"{n}/{sname}"] = s.numpy().copy()
        for k, v in self.powers.items():
            out[k] = np.asarray(float(v), np.float32)
        return out

    def request_shutdown(self) -> None:
        # the apply daemon keeps running while the server drains (workers may be mid-step); it is
        # stopped by PsRpcServer.stop() / LocalCluster.close() / atexit
        self.shutdown_requested.set()
        for q in self.queues.values():
            q.close()


def _spec_from(d: dict) -> OptimizerSpec:
    return OptimizerSpec(**{k: v for k, v in d.items() if not k.startswith("_")})


class _NullCtx:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False
