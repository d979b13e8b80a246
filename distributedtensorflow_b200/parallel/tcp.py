"""TCP control/data plane: one OS process per task, msgpack-RPC at the config.json addresses.

This is the B200 build's stand-in for the reference's gRPC runtime (``tf.train.Server`` at
run_mnist_distributed.py:96-98; ``server.join()`` :102; ``server.target`` :128): a PS task binds
``host:port`` of its own ``(job, task_index)`` and serves the :class:`PsShard` methods; workers
dial lazily and reconnect when a PS restarts (a dead peer surfaces as ``AbortedError``, which the
recoverable session layer handles like TF's ``_RecoverableSession``). On the CPU path tensors
travel through these sockets too (BASELINE config 1); on the B200 path only handles, readiness
and shutdown do — tensors move over NVLink from inside kernels.

Wire format: 4-byte big-endian length + msgpack map ``{m: method, a: args}`` ->
``{ok: bool, r: result | e: message, t: error type}``. numpy arrays are msgpack ext type 1
(dtype string, shape, raw bytes): no pickle, so a peer cannot make this process run code.
"""
from __future__ import annotations

import socket
import socketserver
import struct
import threading
import time
from typing import Any, Optional, Sequence

import msgpack
import numpy as np

from .ps_core import AbortedError, OutOfRangeError, PsShard
from .transport import PsClient

_EXT_NDARRAY = 1
_MAX_MSG = 1 << 31


def _default(obj):
    if isinstance(obj, np.ndarray):
        a = np.ascontiguousarray(obj)
        head = msgpack.packb((a.dtype.str, list(a.shape)), use_bin_type=True)
        return msgpack.ExtType(_EXT_NDARRAY, struct.pack(">I", len(head)) + head + a.tobytes())
    if isinstance(obj, (np.integer,)):
        return int(obj)
    if isinstance(obj, (np.floating,)):
        return float(obj)
    if isinstance(obj, tuple):
        return list(obj)
    raise TypeError(f"cannot serialise {type(obj)}")


def _ext_hook(code, data):
    if code == _EXT_NDARRAY:
        (hl,) = struct.unpack(">I", data[:4])
        dtype, shape = msgpack.unpackb(data[4:4 + hl], raw=False)
        return np.frombuffer(data[4 + hl:], dtype=np.dtype(dtype)).reshape(shape).copy()
    return msgpack.ExtType(code, data)


def pack(obj) -> bytes:
    return msgpack.packb(obj, default=_default, use_bin_type=True)


def unpack(data: bytes):
    return msgpack.unpackb(data, ext_hook=_ext_hook, raw=False, strict_map_key=False,
                           max_bin_len=_MAX_MSG, max_ext_len=_MAX_MSG, max_str_len=_MAX_MSG,
                           max_array_len=1 << 24, max_map_len=1 << 24)


def _recv_exact(sock: socket.socket, n: int) -> bytes:
    buf = bytearray(n)
    view = memoryview(buf)
    got = 0
    while got < n:
        k = sock.recv_into(view[got:], n - got)
        if k == 0:
            raise ConnectionError("peer closed the connection")
        got += k
    return bytes(buf)


def send_msg(sock: socket.socket, obj) -> None:
    data = pack(obj)
    sock.sendall(struct.pack(">I", len(data)) + data)


def recv_msg(sock: socket.socket):
    (n,) = struct.unpack(">I", _recv_exact(sock, 4))
    return unpack(_recv_exact(sock, n))


# Methods a remote peer may invoke on a PsShard (everything else is refused).
_ALLOWED = {
    "create", "assign", "assign_slot", "assign_power", "reset_slots", "mark_initialized",
    "is_initialized", "uninitialized", "pull", "read_powers", "push_apply", "finish", "step_add",
    "read_scalar", "heartbeat", "accum_create", "accum_apply", "accum_take", "accum_take_apply",
    "accum_set_global_step", "accum_stats", "queue_create", "queue_enqueue_many", "queue_dequeue",
    "queue_size", "queue_close", "snapshot", "declare", "nvl_alloc", "nvl_info", "request_shutdown", "ping", "kv_set", "kv_get",
    "mbox_start", "mbox_stats", "mbox_pause", "mbox_attach", "mirror_register", "mirror_unregister", "save_arena_ops",
}


class _Handler(socketserver.BaseRequestHandler):
    def handle(self):
        srv: "PsRpcServer" = self.server.owner  # type: ignore[attr-defined]
        sock = self.request
        sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
        while True:
            try:
                req = recv_msg(sock)
            except (ConnectionError, OSError, struct.error):
                return
            method, args = req.get("m"), req.get("a", [])
            try:
                if method not in _ALLOWED:
                    raise PermissionError(f"method {method!r} is not exported")
                result = srv.dispatch(method, args)
                resp = {"ok": True, "r": result}
            except Exception as e:  # the error travels back to the caller
                resp = {"ok": False, "e": str(e), "t": type(e).__name__}
            try:
                send_msg(sock, resp)
            except (ConnectionError, OSError):
                return


class _TcpServer(socketserver.ThreadingTCPServer):
    allow_reuse_address = True
    daemon_threads = True
    request_queue_size = 128


class PsRpcServer:
    """Serves one :class:`PsShard` (or any object) on ``host:port`` in background threads."""

    def __init__(self, shard: PsShard, host: str, port: int):
        self.shard = shard
        self.kv: dict = {}
        self.last_activity = time.monotonic()
        self.drain_idle, self.drain_max = 1.0, 10.0
        self._kv_cv = threading.Condition()
        self._srv = _TcpServer((host, port), _Handler)
        self._srv.owner = self  # type: ignore[attr-defined]
        self.host, self.port = self._srv.server_address[:2]
        self._thread = threading.Thread(target=self._srv.serve_forever, kwargs={"poll_interval": 0.05},
                                        name=f"ps-rpc-{self.port}", daemon=True)
        self._thread.start()

    def dispatch(self, method: str, args: Sequence[Any]):
        self.last_activity = time.monotonic()
        if method == "ping":
            return {"task": self.shard.task_index, "incarnation": self.shard.incarnation}
        if method == "kv_set":
            with self._kv_cv:
                self.kv[args[0]] = args[1]
                self._kv_cv.notify_all()
            return True
        if method == "kv_get":
            key, timeout = args[0], (args[1] if len(args) > 1 else 0.0)
            deadline = time.monotonic() + (timeout or 0.0)
            with self._kv_cv:
                while key not in self.kv:
                    rem = deadline - time.monotonic()
                    if rem <= 0:
                        return None
                    self._kv_cv.wait(min(rem, 0.5))
                return self.kv[key]
        return getattr(self.shard, method)(*args)

    def join(self, poll: float = 0.1) -> None:
        """Block until a shutdown request arrives (``server.join()``), then drain: keep serving
        until the other tasks have been quiet for ``drain_idle`` seconds (bounded), so workers
        that are mid-step when the chief says "done" can finish that step and stop by themselves."""
        while not self.shard.shutdown_requested.wait(poll):
            if getattr(self.shard, "signal_pending", False):     # SIGINT / SIGTERM seen by Server.join
                self.shard.signal_pending = False
                self.shard.request_shutdown()
        deadline = time.monotonic() + self.drain_max
        while time.monotonic() < deadline and time.monotonic() - self.last_activity < self.drain_idle:
            time.sleep(poll)

    def stop(self) -> None:
        self._srv.shutdown()
        self._srv.server_close()
        if hasattr(self.shard, "mbox_stop"):
            self.shard.mbox_stop()


class TcpClient(PsClient):
    """Dials the PS tasks of a cluster lazily; one persistent connection per (thread, shard)."""

    def __init__(self, ps_endpoints: Sequence[str], connect_timeout: float = 30.0,
                 call_timeout: Optional[float] = None):
        self.endpoints = list(ps_endpoints)
        self.num_shards = len(self.endpoints)
        self.connect_timeout = connect_timeout
        self.call_timeout = call_timeout
        self._local = threading.local()

    def _conn(self, shard: int) -> socket.socket:
        conns = getattr(self._local, "conns", None)
        if conns is None:
            conns = self._local.conns = {}
        s = conns.get(shard)
        if s is not None:
            return s
        host, port = self.endpoints[shard].rsplit(":", 1)
        deadline = time.monotonic() + self.connect_timeout
        last: Optional[Exception] = None
        while time.monotonic() < deadline:
            try:
                s = socket.create_connection((host, int(port)), timeout=5.0)
                s.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
                s.settimeout(self.call_timeout)
                conns[shard] = s
                return s
            except OSError as e:
                last = e
                time.sleep(0.1)
        raise AbortedError(f"/job:ps/task:{shard} at {self.endpoints[shard]} is unreachable: {last}")

    def _drop(self, shard: int) -> None:
        conns = getattr(self._local, "conns", {})
        s = conns.pop(shard, None)
        if s is not None:
            try:
                s.close()
            except OSError:
                pass

    def call(self, shard: int, method: str, *args):
        s = self._conn(shard)
        try:
            send_msg(s, {"m": method, "a": list(args)})
            resp = recv_msg(s)
        except (ConnectionError, OSError, struct.error) as e:
            self._drop(shard)
            raise AbortedError(f"/job:ps/task:{shard} went away during {method}: {e}") from e
        if resp["ok"]:
            return resp["r"]
        et = resp.get("t", "RuntimeError")
        if et == "TimeoutError":
            raise TimeoutError(resp["e"])
        if et == "OutOfRangeError":
            raise OutOfRangeError(resp["e"])
        if et == "KeyError":
            # a restarted PS has lost its variables: same signal TF gives (FailedPrecondition ->
            # the recoverable session re-initialises)
            raise AbortedError(f"/job:ps/task:{shard} lost state ({resp['e']})")
        raise RuntimeError(f"PS {shard} {method} failed: {et}: {resp['e']}")

    def close(self) -> None:
        for shard in list(getattr(self._local, "conns", {}).keys()):
            self._drop(shard)
