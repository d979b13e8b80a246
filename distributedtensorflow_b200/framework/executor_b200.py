"""B200 executor: compiles the reference's model graphs into the hand-written sm_100a kernels and
replays one CUDA graph per training step.

What ``mon_sess.run(train_op)`` costs in the reference: a Python -> gRPC RunStep,
graph partitions on worker and PS, 8 RecvTensor RPCs down and 8 up through host memory, CPU conv /
matmul, ApplyAdam on the PS. Here the same step is ONE graph launch on the worker's GPU:

    H2D(batch) -> clear scratch -> [pull_bulk from remote shards] -> forward (tcgen05 GEMMs /
    implicit-GEMM conv, first layer reading its weights straight from the PS arena) -> fused
    softmax-xent fwd+bwd -> backward GEMMs -> push+apply on the owning shards (+ global_step
    atomic) -> D2H(loss, step)

The graph is pattern-compiled from the TF1-style IR (framework/graph.py). Supported families:
dense stacks (the MLP of templates/00_mnist_replica.py:138-164 with its clipped-sum loss, the wide
MLP) and the CNN of run_mnist_distributed.py:46-70 with the mean sparse softmax loss. Anything
else raises and can run with ``ConfigProto(backend="cpu")``.
"""
from __future__ import annotations

import ctypes as C
import os
import time
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from ..ops import _lib as L
from ..ops import kernels as K
from ..parallel.nvl import NvlStore, SliceRegion, pad8
from ..utils.tracing import nvtx_range
from . import graph as G


# ------------------------------------------------------------------------------------------------
# graph -> plan
# ------------------------------------------------------------------------------------------------
@dataclass
class DenseLayer:
    w: G.Variable
    b: Optional[G.Variable]
    relu: bool
    out_node: G.Tensor


@dataclass
class ConvLayer:
    k: G.Variable
    b: Optional[G.Variable]
    relu: bool
    padding: str


@dataclass
class Plan:
    kind: str                       # "mlp" | "cnn"
    input_node: G.Tensor
    labels_node: G.Tensor
    logits_node: G.Tensor
    loss_node: G.Tensor
    loss_kind: str                  # "sparse_mean" | "dense_clipped_sum"
    dense: List[DenseLayer]
    convs: List[ConvLayer]
    in_features: int
    num_classes: int


class UnsupportedGraphError(NotImplementedError):
    pass


def opts_flag(config, key: str, default: bool) -> bool:
    opts = getattr(config, "b200_options", {}) or {}
    return bool(opts.get(key, default))


def _is_input(n: G.Tensor) -> bool:
    return n.op in ("Placeholder", "IteratorGetNext")


def parse_loss(loss: G.Tensor) -> Tuple[str, G.Tensor, G.Tensor]:
    if loss.op == "Mean" and loss.attrs.get("axis") is None and loss.inputs[0].op == "SparseSoftmaxXent":
        x = loss.inputs[0]
        return "sparse_mean", x.inputs[0], x.inputs[1]
    if loss.op == "Neg" and loss.inputs[0].op == "Sum" and loss.inputs[0].attrs.get("axis") is None:
        mul = loss.inputs[0].inputs[0]
        if mul.op == "Mul":
            for a, b in ((mul.inputs[0], mul.inputs[1]), (mul.inputs[1], mul.inputs[0])):
                if a.op == "Log" and a.inputs[0].op == "ClipByValue":
                    clip = a.inputs[0]
                    if (clip.inputs[0].op == "Softmax" and abs(clip.attrs["min"] - 1e-10) < 1e-16
                            and clip.attrs["max"] == 1.0 and _is_input(b)):
                        return "dense_clipped_sum", clip.inputs[0].inputs[0], b
    raise UnsupportedGraphError("loss expression is not one of the B200-compiled forms "
                                "(mean sparse softmax xent / clipped sum dense xent)")


def parse_plan(loss: G.Tensor) -> Plan:
    loss_kind, logits, labels = parse_loss(loss)
    layers: List[Any] = []
    node = logits
    while True:
        if _is_input(node):
            input_node = node
            break
        if node.op == "Reshape":
            layers.append(("reshape", node.attrs["shape"]))
            node = node.inputs[0]
            continue
        if node.op == "MaxPool":
            if node.attrs["pool_size"] != (2, 2) or node.attrs["strides"] != (2, 2):
                raise UnsupportedGraphError("only 2x2/2 max-pool is compiled for B200")
            layers.append(("pool",))
            node = node.inputs[0]
            continue
        out_node = node
        relu = False
        if node.op == "Relu":
            relu, node = True, node.inputs[0]
        bias = None
        if node.op == "BiasAdd":
            bias, node = node.inputs[1], node.inputs[0]
        if node.op == "MatMul":
            layers.append(DenseLayer(node.inputs[1], bias, relu, out_node))
            node = node.inputs[0]
        elif node.op == "Conv2D":
            layers.append(ConvLayer(node.inputs[1], bias, relu, node.attrs["padding"]))
            node = node.inputs[0]
        else:
            raise UnsupportedGraphError(f"op {node.op} ({node.name}) is not compiled for B200")
    layers.reverse()
    seq = [l for l in layers if not (isinstance(l, tuple) and l[0] == "reshape")]
    dense = [l for l in seq if isinstance(l, DenseLayer)]
    convs = [l for l in seq if isinstance(l, ConvLayer)]
    for l in dense + convs:
        if not isinstance(l.w if isinstance(l, DenseLayer) else l.k, G.Variable):
            raise UnsupportedGraphError("layer weights must be Variables")
    if not dense or dense[-1].relu:
        raise UnsupportedGraphError("the graph must end in a linear (logits) dense layer")
    if convs:
        kinds = ["conv" if isinstance(l, ConvLayer) else "pool" if isinstance(l, tuple) else "dense" for l in seq]
        ok = (kinds[:4] == ["conv", "pool", "conv", "pool"] and all(k == "dense" for k in kinds[4:])
              and tuple(convs[0].k.shape) == (5, 5, 1, 32) and tuple(convs[1].k.shape) == (5, 5, 32, 64)
              and all(c.relu and c.padding == "SAME" and c.b is not None for c in convs)
              and dense[0].w.shape[0] == 3136)
        if not ok:
            raise UnsupportedGraphError("only the reference CNN topology (conv5x5x32-pool-conv5x5x64-"
                                        "pool-dense...) is compiled for B200")
        kind, in_features = "cnn", 784
    else:
        if any(isinstance(l, tuple) for l in seq):
            raise UnsupportedGraphError("pooling without convolutions is not compiled for B200")
        kind, in_features = "mlp", int(dense[0].w.shape[0])
    for a, b in zip(dense[:-1], dense[1:]):
        if a.w.shape[1] != b.w.shape[0] or not a.relu:
            raise UnsupportedGraphError("hidden dense layers must chain and use ReLU")
    return Plan(kind, input_node, labels, logits, loss, loss_kind, dense, convs, in_features,
                int(dense[-1].w.shape[1]))


# ------------------------------------------------------------------------------------------------
# per-batch-size device state
# ------------------------------------------------------------------------------------------------
class _Zero:
    """One contiguous fp32 region holding every atomically accumulated buffer of a step, cleared
    by a single fill kernel."""

    def __init__(self):
        self.sizes: List[Tuple[str, int]] = []

    def add(self, name: str, n: int) -> None:
        self.sizes.append((name, (n + 63) // 64 * 64))

    def build(self, device) -> Dict[str, torch.Tensor]:
        total = sum(n for _, n in self.sizes)
        self.buf = torch.zeros(max(total, 64), dtype=torch.float32, device=device)
        out, o = {}, 0
        self.ws_off = None                 # split-K workspaces / counters are added last: [ws_off, end)
        for name, n in self.sizes:
            if self.ws_off is None and (name.startswith("ws_") or name.startswith("ct_")):
                self.ws_off = o
            out[name] = self.buf[o:o + n]
            o += n
        return out


SPLITK_MAX_TILES = int(os.environ.get("DTF_B200_SPLITK_MAX_TILES", "40"))
SPLITK_MAX_SPLITS = int(os.environ.get("DTF_B200_SPLITK_MAX_SPLITS", "64"))


def plan_gemm(M: int, N: int, K: int) -> Tuple[int, int]:
    """(block_n, splits) for the skinny MNIST GEMMs: prefer more, narrower tiles while the grid is
    far below 148 SMs, and split K (with the in-kernel fix-up) until the grid roughly fills them."""
    if -(-M // 256) * -(-N // 256) >= 74 and K >= 256 and os.environ.get("DTF_B200_NO_PAIR_GEMM") != "1":
        return 256, 1   # large GEMM: persistent CTA-pair kernel (256 x 256 tiles, cta_group::2)
    tiles128 = -(-M // 128) * -(-N // 128)
    bn = 64 if (N <= 64 or tiles128 < 74) else 128
    tiles = -(-M // 128) * -(-N // bn)
    kb = -(-K // 64)
    splits = 1
    if tiles < SPLITK_MAX_TILES and kb >= 4:
        splits = max(1, min(kb // 2, 148 // tiles, SPLITK_MAX_SPLITS))
    return bn, splits


class StepState:
    def __init__(self, runner: "B200StepRunner", B: int, train: bool, x_dtype, y_dtype):
        self.B, self.train = B, train
        self.gemm_plans: Dict[str, dict] = {}
        p, dev = runner.plan, runner.device
        bf, f32 = torch.bfloat16, torch.float32
        ysh = (B,) if p.loss_kind == "sparse_mean" else (B, p.num_classes)
        # inputs and labels share ONE pinned staging buffer and ONE device buffer: a step's H2D is a
        # single copy node (two nodes cost an extra ~4 us of graph dependency latency)
        xb = B * p.in_features * torch.empty((), dtype=x_dtype).element_size()
        yb = int(np.prod(ysh)) * torch.empty((), dtype=y_dtype).element_size()
        yoff = (xb + 255) // 256 * 256
        self.io_pin = torch.zeros(yoff + (yb + 255) // 256 * 256, dtype=torch.uint8).pin_memory()
        self.io_dev = torch.zeros(self.io_pin.numel(), dtype=torch.uint8, device=dev)
        self.x_pin = self.io_pin[:xb].view(x_dtype).view(B, p.in_features)
        self.y_pin = self.io_pin[yoff:yoff + yb].view(y_dtype).view(ysh)
        self.y_bytes16 = (yb + 15) // 16 * 16
        self.x_dev = self.io_dev[:xb].view(x_dtype).view(B, p.in_features)
        self.y_dev = self.io_dev[yoff:yoff + yb].view(y_dtype).view(ysh)
        self.out_pin = torch.zeros(4, dtype=torch.float32).pin_memory()      # loss
        self.step_pin = torch.zeros(4, dtype=torch.int64).pin_memory()       # global_step, pulled, min consumed
        self.step_dev = torch.zeros(4, dtype=torch.int64, device=dev)
        # numpy views of the pinned buffers: the per-step host path reads / writes them directly
        self.x_np, self.y_np = self.x_pin.numpy(), self.y_pin.numpy()
        self.out_np, self.step_np = self.out_pin.numpy(), self.step_pin.numpy()
        # result sink: the step's last kernel stores {loss, seq, global_step} here (pinned, mapped)
        self.res_pin = torch.zeros(4, dtype=torch.int32).pin_memory()
        self.res_u32 = self.res_pin.numpy().view(np.uint32)
        self.res_f32 = self.res_pin.numpy().view(np.float32)
        self.res_u64 = self.res_pin.numpy().view(np.uint64)
        self.res_seq_dev = torch.zeros(4, dtype=torch.int32, device=dev)
        self.res_expected = 0
        self.zero_copy = False
        if not train:
            self.head_dw = False
        self.counter = torch.zeros(4, dtype=torch.int32, device=dev)
        dims = [int(d.w.shape[0]) for d in p.dense] + [p.num_classes]
        self.dims = dims
        nl = len(p.dense)
        # activations: act[i] is the input of dense layer i (act[0] = network / conv output)
        self.act = [torch.zeros((B, pad8(dims[i])), dtype=bf, device=dev) for i in range(nl)]
        self.logits = torch.zeros((B, p.num_classes), dtype=f32, device=dev)
        zero = _Zero()
        zero.add("loss", 1)
        if train:
            self.dlogits = torch.zeros((B, pad8(p.num_classes)), dtype=bf, device=dev)
            self.dact = [None] + [torch.zeros((B, pad8(dims[i])), dtype=bf, device=dev) for i in range(1, nl)]
            self.gw = [torch.zeros((dims[i], dims[i + 1]), dtype=f32, device=dev) for i in range(nl)]
            for i in range(nl):
                zero.add(f"gb{i}", dims[i + 1])
            # the fused head kernel also accumulates the logits layer's weight gradient (atomics):
            # that buffer lives in the zeroed region
            self.head_dw = bool(runner._head_is_fused() and runner.head_dw)
            if self.head_dw:
                zero.add("gw_last", dims[nl - 1] * dims[nl])
        if p.kind == "cnn":
            self.pool1p = torch.zeros((B, 18, 18, 32), dtype=bf, device=dev)
            self.amax1 = torch.zeros((B, 14, 14, 32), dtype=torch.uint8, device=dev)
            self.amax2 = torch.zeros((B, 7, 7, 64), dtype=torch.uint8, device=dev)
            if train:
                self.dpool2 = torch.zeros((B, 3136), dtype=bf, device=dev)
                self.dconv2p = torch.zeros((B, 18, 18, 64), dtype=bf, device=dev)
                self.dpool1 = torch.zeros((B, 14, 14, 32), dtype=bf, device=dev)
                zero.add("gk1", 25 * 32)
                zero.add("gkb1", 32)
                zero.add("gk2", 800 * 64)
                zero.add("gkb2", 64)
        # split-K workspaces + tile counters of the skinny GEMMs live in the zeroed region too
        nl_ = len(p.dense)
        wants = [(f"fwd{i}", B, dims[i + 1], dims[i]) for i in range(nl_)]
        if train:
            wants += [(f"dx{i}", B, dims[i], dims[i + 1]) for i in range(1 if p.kind == "mlp" else 0, nl_)]
        for key, M_, N_, K_ in wants:
            bn, splits = plan_gemm(M_, N_, K_)
            plan = dict(bn=bn, splits=splits)
            if splits > 1:
                ldws = (N_ + 3) // 4 * 4
                zero.add("ws_" + key, M_ * ldws)
                zero.add("ct_" + key, -(-M_ // 128) * -(-N_ // bn))
                plan["ldws"] = ldws
            self.gemm_plans[key] = plan
        self.z = zero.build(dev)
        if train and self.head_dw:
            self.gw[nl - 1] = self.z["gw_last"][:dims[nl - 1] * dims[nl]].view(dims[nl - 1], dims[nl])
        for key, plan in self.gemm_plans.items():
            if plan["splits"] > 1:
                M_ = B
                plan["ws"] = self.z["ws_" + key][:M_ * plan["ldws"]].view(M_, plan["ldws"])
                plan["counters"] = self.z["ct_" + key]
        self.zero = zero
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.graph_exec = 0           # raw cudaGraphExec_t of `graph`
        self.launches_per_step = 0


# ------------------------------------------------------------------------------------------------
# the runner
# ------------------------------------------------------------------------------------------------
class B200StepRunner:
    def __init__(self, graph: G.Graph, client, layout, worker_id: int, config, cluster=None):
        if not torch.cuda.is_available():
            raise RuntimeError("backend 'b200' needs a CUDA device")
        L.lib()  # fail loudly when the extension is missing
        self.graph, self.client, self.layout = graph, client, layout
        self.worker_id, self.config, self.cluster = worker_id, config, cluster
        opts = getattr(config, "b200_options", {}) or {}
        self.use_cuda_graph = bool(opts.get("cuda_graph", True))
        self.fused_first_pull = bool(opts.get("fused_first_pull", True))
        self.atomic_sgd = bool(opts.get("atomic_sgd", False))
        self.health_every = int(opts.get("health_every", 256))   # control-plane ping period (steps)
        self._runs = 0
        # host-side breakdown of run() (seconds): set to a dict(stage=0, admit=0, launch=0, wait=0, n=0) to collect
        self.host_prof: Optional[Dict[str, float]] = None
        self._incarnations: Dict[int, int] = {}
        # dW GEMM whose epilogue is the push+apply (no gradient buffer). Measured on the reference CNN it
        # is slower than GEMM + the coalesced apply kernel (180 vs 158 us/step), so it is opt-in.
        self.fused_apply = bool(opts.get("fused_apply", False))
        self.clear_at_end = bool(opts.get("clear_at_end", True))
        # logits GEMM + softmax-xent + dX GEMM of the last layer + bias column-sum as ONE kernel
        # (csrc/kernels/head_fused.cu) when the head has <= 16 classes
        self.fused_head = bool(opts.get("fused_head", True))
        # ... and the logits layer's weight gradient too (327 K float atomics for the CNN head: measured
        # +7 us on the critical path for a 10 us GEMM saved on the side stream, so off by default)
        self.head_dw = bool(opts.get("head_dw", False))
        # the step's last kernel writes loss / global_step into pinned host memory and the host
        # spins on a sequence word: no D2H copy nodes, no stream synchronise per step
        self.host_sink = bool(opts.get("host_sink", True))
        # report loss / global_step to the host right after the loss kernel, while the backward pass
        # and the applies are still running: the host's per-step work overlaps the rest of the step
        # (the next step is stream-ordered behind it; host reads of shard memory sync the stream)
        self.early_report = bool(opts.get("early_report", True))
        # the step's first kernel reads the input batch in place from pinned host memory
        self.zero_copy_input = bool(opts.get("zero_copy_input", True))
        # accumulators clean themselves (the apply zeroes the gradients it consumed, the split-K
        # fix-up its workspace, the finishing thread the loss): no fill kernel at the end of the step
        self.self_clean = bool(opts.get("self_clean", True))     # measured +1-2% with PDL off
        # push(+apply) of each dense kernel as soon as its gradient exists (overlaps NVLink traffic
        # with the rest of the backward pass). With every shard local the applies are pure HBM work
        # that only fights the compute kernels for SMs: one apply kernel at the end is faster.
        self.early_apply = opts.get("early_apply", None)
        # owner-applies mode: the dense kernels' weight-gradient GEMMs store straight into the owners'
        # mailboxes from their epilogue (no gradient buffer, no copy kernel). Off until it has been
        # measured on more than one GPU.
        self.fused_push = bool(opts.get("fused_push", False))
        self.early_apply_ctas = int(opts.get("early_apply_ctas", 0))   # cap the grid of overlapped applies
        # owner-applies mode: only gradients at least this large get their own early push kernel (its
        # NVLink time hides behind the backward pass); smaller ones would just serialise the push
        # stream in front of the big one and ride in the final push instead
        self.early_push_min_bytes = int(opts.get("early_push_min_bytes", 256 * 1024))
        # asynchronous gradient push: "rmw" = the worker's SMs read-modify-write the owner's state
        # over NVLink (dtf_push_apply); "mailbox" = posted gradient stores + the owner's native apply
        # daemon (csrc/kernels/mailbox.cu, csrc/runtime/mbox_daemon.cu); "auto" = mailbox whenever
        # there is more than one worker (a lone worker keeps the exact in-order apply)
        self.push_mode = str(opts.get("push_mode", os.environ.get("DTF_B200_PUSH_MODE", "auto")))
        self.mailbox_depth = int(opts.get("mailbox_depth", 4))      # = max pushes in flight per worker
        self.mailbox_timeout = float(opts.get("mailbox_timeout", 60.0))
        self.mailbox_poll_us = int(opts.get("mailbox_poll_us", 20))
        # owner-applies mode, "owner pushes, worker never pulls": every worker keeps a mirror of the
        # parameters in its own HBM that the owners' drains rewrite (csrc/kernels/mailbox.cu); the
        # step then contains no NVLink read at all (no pull, no flow-control / global_step round trip)
        self.mirror = bool(opts.get("mirror", os.environ.get("DTF_B200_MIRROR", "1") != "0"))
        # a shard that lives in this worker's process is drained by a node of the worker's own step
        # graph (device-driven apply); the shard's daemon stays as the backstop for an idle owner
        self.graph_drain = bool(opts.get("graph_drain", os.environ.get("DTF_B200_GRAPH_DRAIN", "1") != "0"))
        self.drain_at = str(opts.get("drain_at", os.environ.get("DTF_B200_DRAIN_AT", "start")))
        self.drain_backstop_us = int(opts.get("drain_backstop_us", 400))
        # grid cap of the in-graph drain: it runs beside the forward pass, where a flood of CTAs
        # doubled the convolutions' time; a few dozen CTAs over a longer window are free
        # (-1 = auto: a short drain (<= 1024 chunks of 4096 elements, the MNIST models) is capped at 148 CTAs;
        # a long one (the wide MLP: thousands of chunks) runs uncapped at low priority - capped, its few
        # resident CTAs slowed the persistent GEMM's CTAs on the SMs they shared 3x, and the static tile
        # schedule made the whole GEMM wait for them)
        self.drain_ctas = int(opts.get("drain_ctas", os.environ.get("DTF_B200_DRAIN_CTAS", "-1")))
        self.use_mirror = False
        self.sync_generic = False
        self.sync_path = str(opts.get("sync_path", os.environ.get("DTF_B200_SYNC_PATH", "auto")))   # auto | mailbox
        self.sync_token_timeout = float(opts.get("sync_token_timeout", 20.0))
        self.device_flow = False
        self.max_queue = int(opts.get("max_queue", 64))     # steps the host may run ahead of the device
        self.my_shards: List[int] = []
        self.use_mailbox = False
        self._pushes = 0
        self._min_consumed = 0
        self.split_k = bool(opts.get("split_k", True))
        # conv2's pool/ReLU gradient as the epilogue of the dX GEMM in front of it (one kernel less on the
        # backward critical path)
        self.fused_pool_bwd = bool(opts.get("fused_pool_bwd", os.environ.get("DTF_B200_FUSED_POOL_BWD", "1") != "0"))
        # programmatic dependent launch: inside the captured multi-stream graph it measured 2-3% SLOWER
        # than plain kernel edges (the programmatic edges constrain cross-stream overlap), so off
        self.pdl = bool(opts.get("pdl", False))
        self.pdl_chain = bool(opts.get("pdl_chain", os.environ.get("DTF_B200_PDL_CHAIN", "0") == "1"))
        self.pdl_forward = bool(opts.get("pdl_forward", os.environ.get("DTF_B200_PDL_FORWARD", "1") == "1"))
        L.lib().dtf_set_pdl(1 if self.pdl else 0)
        self.device = torch.device(self._pick_device())
        torch.cuda.set_device(self.device)
        self.store: Optional[NvlStore] = None
        self.plan: Optional[Plan] = None
        self.spec = None
        self.train_op = None
        self.states: Dict[Tuple[int, bool], StepState] = {}
        self.local: Dict[str, torch.Tensor] = {}     # pulled copies of remote tensors
        # the activation-gradient chain is the critical path: its stream outranks the side stream
        # that carries weight-gradient GEMMs and the bandwidth-heavy applies
        # priorities: activation-gradient chain > weight-gradient GEMMs > bandwidth-heavy applies, so a
        # pending weight-gradient CTA is dispatched before the apply's hundreds of CTAs
        def mk(prio):
            try:
                return torch.cuda.Stream(device=self.device, priority=prio)
            except Exception:
                return torch.cuda.Stream(device=self.device, priority=max(prio, -1))
        self.stream = mk(-3)
        self.side_stream = mk(-2)
        # pushes: the mailbox push is a light copy kernel that should start as soon as its gradient
        # exists (NVLink time overlaps compute), so it gets the side stream's priority; the remote
        # read-modify-write apply is heavy and yields to the compute kernels
        self.apply_stream = mk(int(opts.get("apply_stream_priority", 0 if self.push_mode == "rmw" else -2)))
        self.overlap = bool(opts.get("overlap", True))
        self._stream_handle = self.stream.cuda_stream
        self._graph_launch = L.lib().dtf_graph_launch

    def _pick_device(self) -> str:
        if getattr(self.config, "device", None):
            return self.config.device
        if self.cluster is not None and "worker" in self.cluster.jobs:
            try:
                d = self.cluster.parsed_address("worker", self.worker_id).device
                if d and d.startswith("cuda"):
                    return d
            except Exception:
                pass
        return "cuda:%d" % torch.cuda.current_device()

    def _shard_devices(self) -> List[Optional[int]]:
        out: List[Optional[int]] = []
        for s in range(self.client.num_shards):
            dev = None
            if self.cluster is not None and "ps" in self.cluster.jobs:
                try:
                    d = self.cluster.parsed_address("ps", s).device
                    if d and d.startswith("cuda:"):
                        dev = int(d.split(":")[1])
                except Exception:
                    dev = None
            out.append(dev)
        return out

    # -- variable lifecycle (called by Session) ---------------------------------------------------
    handles_sync = True   # SyncReplicasOptimizer rounds run in csrc/kernels/sync_step.cu / mailbox.cu

    def _num_workers(self) -> int:
        if self.cluster is not None and "worker" in self.cluster.jobs:
            return self.cluster.num_tasks("worker")
        if self.sync_opt is not None:        # in-process cluster: the optimizer knows how many replicas there are
            return int(self.sync_opt.total_num_replicas)
        return 1

    @property
    def coop_sync(self) -> bool:
        """Synchronous mode through the one-kernel cooperative round (symmetric topologies)."""
        return self.sync_opt is not None and not self.sync_generic

    def create_variables(self, spec, trainable_names: Sequence[str]) -> None:
        self.spec = spec
        self.trainable_names = list(trainable_names)
        top = next((n for n in self.graph.nodes if type(n).__name__ == "TrainOp"), None)
        self.sync_opt = top.sync if top is not None else None
        sync_workers = 0
        self.sync_generic = False
        if self.sync_opt is not None:
            world = self.client.num_shards
            nw = self._num_workers()
            # (several worker threads of ONE process share a GPU: two all-SM cooperative kernels cannot
            # be resident together, so an in-process cluster with more than one worker is not symmetric)
            symmetric = (nw == world and self.sync_opt.total_num_replicas == world
                         and self.sync_opt.replicas_to_aggregate == world
                         and (self.cluster is not None or nw == 1))
            # The one-kernel cooperative round needs a worker next to every shard and all of them in
            # every round. Anything else - backup workers (replicas_to_aggregate < total_num_replicas),
            # PS-only tasks, "1 PS + N workers" - runs the round on the mailboxes: tagged pushes, the
            # owner's drain averages the fresh gradients and drops stale ones, tokens are epoch words
            # in the workers' mirror headers (csrc/kernels/mailbox.cu, dtf_mailbox_drain_sync).
            self.sync_generic = (not symmetric) or self.sync_path == "mailbox"
            if self.sync_generic:
                if nw > L.MAX_WORKERS or not self.mirror:
                    raise NotImplementedError("synchronous mode on the mailboxes supports at most %d workers "
                                              "and needs worker mirrors" % L.MAX_WORKERS)
                if self.sync_opt.replicas_to_aggregate > nw:
                    raise NotImplementedError("replicas_to_aggregate (%d) > number of workers (%d): gradient "
                                              "accumulation over several steps of one worker is only on the "
                                              "CPU backend" % (self.sync_opt.replicas_to_aggregate, nw))
            else:
                sync_workers = world
        depth, mirrors = 1, True
        self.use_mailbox = False
        if self.sync_generic:
            nw = self._num_workers()
            self.use_mailbox = True
            sync_workers, depth, mirrors = nw, 2, False
        elif self.sync_opt is None and spec is not None and self.trainable_names and self.push_mode != "rmw":
            if self.push_mode not in ("auto", "mailbox"):
                raise ValueError("push_mode must be 'auto', 'rmw' or 'mailbox' (got %r)" % self.push_mode)
            nw = self.cluster.num_tasks("worker") if (self.cluster is not None and "worker" in self.cluster.jobs) else 1
            fits = nw <= L.MAX_WORKERS and self.worker_id < nw
            if self.push_mode == "mailbox" and not fits:
                raise NotImplementedError("push_mode='mailbox' supports at most %d workers" % L.MAX_WORKERS)
            # a lone worker keeps the exact in-order apply; with several workers the applies are
            # interleaved anyway and the write-only push is what scales
            # use_locking=True asks for serialised applies (TF takes the variable's mutex): the owner's
            # daemon applies one gradient at a time per shard, which is exactly that
            locked = bool(getattr(spec, "use_locking", False))
            self.use_mailbox = fits and (self.push_mode == "mailbox" or nw > 1 or locked)
            if self.use_mailbox:
                sync_workers, depth, mirrors = nw, max(1, self.mailbox_depth), False
        self.store = NvlStore(self.client, self.layout, spec, trainable_names, self.device.index or 0,
                              self._shard_devices(), sync_workers=sync_workers, mailbox_depth=depth,
                              mirrors=mirrors)
        self.sync_round_no: Optional[int] = None
        self.mirror_filled = False
        self.store.sync_hook = self.stream.synchronize
        self.use_mirror = bool(self.use_mailbox and (self.mirror or self.sync_generic)
                               and self.store.num_shards <= L.MIRROR_MAX_SHARDS)
        if self.use_mirror:
            self.store.alloc_mirror()
        # flow control on the device needs the mirror header (and pushes made by the push kernel, not
        # by a GEMM epilogue)
        self.device_flow = bool(self.use_mirror and not self.fused_push
                                and opts_flag(self.config, "device_flow", True))
        if self.use_mailbox:
            self._mbox_setup()

    def init_variables(self, values: Dict[str, np.ndarray]) -> None:
        st = self.store
        for name, arr in values.items():
            if name in st.arena.regions:
                st.write_variable(name, arr)
        if self.spec is not None:
            for name in self.trainable_names:
                for i in range(self.spec.num_slots):
                    st.write_slot(name, i, self.spec.slot_init(i))
        step0 = int(values[st.arena.step_var]) if st.arena.step_var in values else 0
        for s in range(st.num_shards):
            b1 = self.spec.beta1 if (self.spec is not None and self.spec.kind == "adam") else 0.0
            b2 = self.spec.beta2 if (self.spec is not None and self.spec.kind == "adam") else 0.0
            st.write_ctrl(s, global_step=step0 if s == st.arena.step_shard else 0, beta1_power=b1,
                          beta2_power=b2, initialized=1)
        self.client.mark_initialized(True)
        self.mirror_filled = False
        self.sync_round_no = None

    def uninitialized_variables(self) -> List[str]:
        return self.client.uninitialized_variables(self.layout)

    def snapshot_shards(self) -> List[Dict[str, np.ndarray]]:
        """Per-shard tensors for the Saver. In owner-applies mode the shards' apply daemons are parked
        while their memory is read, so each shard's tensors are a consistent cut (no half-applied
        gradient); workers that keep pushing meanwhile just fill their mailboxes (bounded by depth)."""
        self.stream.synchronize()
        if not self.use_mailbox:
            return self._snapshot_shards()
        self.flush_pushes()
        paused = []
        try:
            for sh in range(self.client.num_shards):
                self.client.call(sh, "mbox_pause", True)
                paused.append(sh)
            return self._snapshot_shards()
        finally:
            for sh in paused:
                try:
                    self.client.call(sh, "mbox_pause", False)
                except Exception:
                    pass

    def save_parallel(self, prefix: str, var_names) -> int:
        """Sharded checkpoint written by the shards themselves. The chief plans where every tensor
        lives (data file = the variable's home shard, tensors in key order), pre-sizes the data
        files, and sends every PS task the list of byte ranges of ITS arena to store
        (``PsShard.save_arena_ops``); all tasks copy and write in parallel and return per-range
        checksums, which the chief combines into the index it writes last. A flat-sliced variable
        becomes ONE contiguous tensor of the checkpoint although N tasks wrote it. Returns bytes
        written. (Reference: templates/00_between_graph_replication_async_mnist.py:43-45.)"""
        import threading
        from ..ckpt import bundle as B
        st, al = self.store, self.store.arena
        self.stream.synchronize()
        self.flush_pushes()
        slot_names = self.spec.slot_names() if self.spec is not None else []
        n_files = st.num_shards
        # -- plan ---------------------------------------------------------------------------------
        tensors: Dict[str, dict] = {}      # key -> dict(shape, nbytes, file, pieces=[(shard, arena_off, first, n)])
        for name, regs in al.regions.items():
            if name not in var_names or not regs:
                continue
            vl = self.layout[name]
            regs = sorted(regs, key=lambda r: r.sl.first)
            keys = [(name, None)] + [(f"{name}/{sn}", i) for i, sn in enumerate(slot_names) if i < len(regs[0].slot_off)]
            for key, slot in keys:
                pieces = [(r.shard, r.var_off if slot is None else r.slot_off[slot], r.sl.first, r.sl.n) for r in regs]
                tensors[key] = dict(shape=tuple(vl.shape), nbytes=vl.num_elements * 4, file=regs[0].shard,
                                    pieces=pieces, dtype=B.DT_FLOAT)
        small: Dict[str, np.ndarray] = {}  # scalars that live in control blocks: written by the chief
        if al.step_var:
            vl = self.layout[al.step_var]
            npdt = np.int64 if vl.dtype == "int64" else np.int32
            small[al.step_var] = (np.asarray(st.read_global_step(), npdt).reshape(vl.shape), al.step_shard)
        if self.spec is not None and self.spec.kind == "adam":
            c = st.read_ctrl(al.powers_shard)
            small["beta1_power"] = (np.asarray(c["beta1_power"], np.float32).reshape(()), al.powers_shard)
            small["beta2_power"] = (np.asarray(c["beta2_power"], np.float32).reshape(()), al.powers_shard)
        for key, (arr, shard) in small.items():
            tensors[key] = dict(shape=tuple(arr.shape), nbytes=arr.nbytes, file=shard, pieces=None, value=arr,
                                dtype=B._NP_TO_DT[arr.dtype])
        offsets = [0] * n_files
        for key in sorted(tensors):
            t = tensors[key]
            t["offset"] = offsets[t["file"]]
            offsets[t["file"]] += t["nbytes"]
        idx = B.BundleIndexWriter(prefix, n_files)
        finals = [idx.data_path(i) for i in range(n_files)]
        tag = ".tmp%d_%x" % (os.getpid(), int(time.time() * 1e6) & 0xffffff)
        tmps = [p + tag for p in finals]
        for p_, size in zip(tmps, offsets):
            with open(p_, "wb") as f:
                f.truncate(size)
        ops: List[List[list]] = [[] for _ in range(n_files)]
        op_keys: List[List[Tuple[str, int]]] = [[] for _ in range(n_files)]
        for key, t in tensors.items():
            if t["pieces"] is None:
                continue
            for k, (shard, aoff, first, n) in enumerate(t["pieces"]):
                ops[shard].append([tmps[t["file"]], t["offset"] + first * 4, int(aoff), int(n) * 4])
                op_keys[shard].append((key, k))
        # -- every PS task stores its own bytes, all at once ---------------------------------------
        paused = []
        results: List[Optional[List[int]]] = [None] * n_files
        errors: List[BaseException] = []
        try:
            if self.use_mailbox:
                for sh in range(n_files):
                    self.client.call(sh, "mbox_pause", True)
                    paused.append(sh)

            def work(sh):
                try:
                    results[sh] = self.client.call(sh, "save_arena_ops", ops[sh]) if ops[sh] else []
                except BaseException as e:      # surfaced below
                    errors.append(e)
            threads = [threading.Thread(target=work, args=(sh,), daemon=True) for sh in range(n_files)]
            [t.start() for t in threads]
            [t.join() for t in threads]
        finally:
            for sh in paused:
                try:
                    self.client.call(sh, "mbox_pause", False)
                except Exception:
                    pass
        if errors:
            for p_ in tmps:
                try:
                    os.remove(p_)
                except OSError:
                    pass
            raise errors[0]
        piece_crc: Dict[Tuple[str, int], int] = {}
        for sh in range(n_files):
            for (key, k), crc in zip(op_keys[sh], results[sh]):
                piece_crc[(key, k)] = int(crc)
        total = 0
        for key in sorted(tensors):
            t = tensors[key]
            if t["pieces"] is None:
                raw = np.ascontiguousarray(t["value"]).tobytes()
                with open(tmps[t["file"]], "r+b") as f:
                    f.seek(t["offset"])
                    f.write(raw)
                crc = B.crc32c(np.frombuffer(raw, np.uint8))
            else:
                crc = 0
                for k, (_shard, _aoff, _first, n) in enumerate(t["pieces"]):
                    crc = piece_crc[(key, k)] if k == 0 else B.crc32c_combine(crc, piece_crc[(key, k)], n * 4)
            idx.add(key, t["dtype"], t["shape"], t["file"], t["offset"], t["nbytes"], crc)
            total += t["nbytes"]
        for a, b in zip(tmps, finals):
            os.replace(a, b)
        idx.finish()
        return total

    def _snapshot_shards(self) -> List[Dict[str, np.ndarray]]:
        st = self.store
        shards: List[Dict[str, np.ndarray]] = [dict() for _ in range(st.num_shards)]
        slot_names = self.spec.slot_names() if self.spec is not None else []
        for name, regs in st.arena.regions.items():
            for r in regs:
                shards[r.shard][name] = st._d2h(st.var_ptr(r), r.sl.n * 4, np.float32)
                for i, sn in enumerate(slot_names):
                    if i < len(r.slot_off):
                        shards[r.shard][f"{name}/{sn}"] = st._d2h(st.slot_ptr(r, i), r.sl.n * 4, np.float32)
        if st.arena.step_var:
            vl = self.layout[st.arena.step_var]
            npdt = np.int64 if vl.dtype == "int64" else np.int32
            shards[st.arena.step_shard][st.arena.step_var] = np.asarray([st.read_global_step()], npdt)
        if self.spec is not None and self.spec.kind == "adam":
            c = st.read_ctrl(st.arena.powers_shard)
            shards[st.arena.powers_shard]["beta1_power"] = np.asarray(c["beta1_power"], np.float32)
            shards[st.arena.powers_shard]["beta2_power"] = np.asarray(c["beta2_power"], np.float32)
        return shards

    def restore_tensors(self, tensors: Dict[str, np.ndarray], var_names, slot_names: Sequence[str]) -> None:
        st = self.store
        missing = [n for n in var_names if n in st.arena.regions and n not in tensors]
        if missing:
            raise KeyError("Tensors not found in checkpoint: %s" % missing)
        for name in st.arena.regions:
            if name in tensors:
                st.write_variable(name, tensors[name])
            for i, sn in enumerate(slot_names):
                key = f"{name}/{sn}"
                if key in tensors:
                    st.write_slot(name, i, tensors[key])
        step = int(np.asarray(tensors.get(st.arena.step_var, 0)).reshape(-1)[0]) if st.arena.step_var else 0
        adam = self.spec is not None and self.spec.kind == "adam"
        for s in range(st.num_shards):
            b1 = float(tensors.get("beta1_power", self.spec.beta1 if adam else 0.0)) if adam else 0.0
            b2 = float(tensors.get("beta2_power", self.spec.beta2 if adam else 0.0)) if adam else 0.0
            st.write_ctrl(s, global_step=step if s == st.arena.step_shard else 0, beta1_power=b1,
                          beta2_power=b2, initialized=1)
        self.client.mark_initialized(True)

    # -- parameter access ---------------------------------------------------------------------------
    def _regions(self, var: G.Variable) -> List[SliceRegion]:
        return self.store.arena.regions[var.name]

    def _regions_rotated(self, var: G.Variable) -> List[SliceRegion]:
        """Slices in the order THIS worker should touch them: starting at its own shard and going
        round the ring, so that N workers sweeping their slice lists in lock-step hit N different
        shards (and NVLink ports) at any moment instead of all hammering shard 0 first."""
        n = max(1, self.client.num_shards)
        return sorted(self._regions(var), key=lambda r: (r.shard - self.worker_id) % n)

    def _single_local(self, var: G.Variable) -> Optional[SliceRegion]:
        regs = self._regions(var)
        if len(regs) == 1 and self.store.is_local[regs[0].shard]:
            return regs[0]
        return None

    def _setup_params(self) -> None:
        """Decide, per tensor the kernels read, whether to alias the arena (local shard), read it
        in place over NVLink (fused pull) or copy it with pull_bulk. Builds ``self.copies``."""
        st, p = self.store, self.plan
        self.copies: List[Tuple[int, int, int]] = []
        self.wptr: Dict[str, Tuple[int, int]] = {}    # bf16 serving copy: name -> (ptr, ld)
        self.fptr: Dict[str, int] = {}                # fp32 tensor: name -> ptr
        if self.coop_sync:
            self._setup_params_sync()
            return
        if self.use_mirror:
            self._setup_params_mirror()
            return

        def want_shadow(var: G.Variable, allow_in_place: bool) -> None:
            regs = self._regions(var)
            ld = regs[0].shadow_ld
            if len(regs) == 1 and (st.is_local[regs[0].shard] or (allow_in_place and self.fused_first_pull)):
                self.wptr[var.name] = (st.shadow_ptr(regs[0]), ld)
                return
            rows = sum(r.rows for r in regs)
            buf = torch.zeros((rows, ld), dtype=torch.bfloat16, device=self.device)
            self.local["w:" + var.name] = buf
            for r in self._regions_rotated(var):
                row = r.sl.first // r.cols
                self.copies.append((st.shadow_ptr(r), buf.data_ptr() + row * ld * 2, r.rows * ld * 2))
            self.wptr[var.name] = (buf.data_ptr(), ld)

        def want_f32(var: G.Variable, allow_in_place: bool) -> None:
            regs = self._regions(var)
            if len(regs) == 1 and (st.is_local[regs[0].shard] or allow_in_place):
                self.fptr[var.name] = st.var_ptr(regs[0])
                return
            n = self.layout[var.name].num_elements
            buf = torch.zeros(((n + 63) // 64 * 64,), dtype=torch.float32, device=self.device)
            self.local["f:" + var.name] = buf
            for r in self._regions_rotated(var):
                nbytes = (r.sl.n * 4 + 15) // 16 * 16
                self.copies.append((st.var_ptr(r), buf.data_ptr() + r.sl.first * 4, nbytes))
            self.fptr[var.name] = buf.data_ptr()

        if p.kind == "cnn":
            want_f32(p.convs[0].k, True)       # conv1 reads w1/b1 in place (3.3 KB)
            want_f32(p.convs[0].b, True)
            want_shadow(p.convs[1].k, False)   # w2 is read by fwd AND dgrad: needs one snapshot
            want_f32(p.convs[1].b, False)
        # copies[:late_mark] are needed by the step's first kernels; the rest (the dense stack behind
        # the convolutions / behind an MLP's first layer — the bulk of the bytes) is pulled on the
        # side stream while those kernels run, and layer `late_layer` waits for it
        self.late_mark = len(self.copies) if p.kind == "cnn" else None
        self.late_layer = 0 if p.kind == "cnn" else 1
        for i, d in enumerate(p.dense):
            if self.late_mark is None and i == 1:
                self.late_mark = len(self.copies)
            # first GEMM layer of an MLP: weights are used by the forward only -> fused pull
            want_shadow(d.w, allow_in_place=(p.kind == "mlp" and i == 0))
            if d.b is not None:
                want_f32(d.b, False)
        if self.late_mark is None:
            self.late_mark = len(self.copies)

    def _setup_params_sync(self) -> None:
        """Synchronous mode: kernels read the worker's MIRROR inside its own arena; the owners'
        all-gather keeps it current, so there is no per-step pull. ``copies`` fills the mirror once
        (after init / restore) from the owners' serving copies."""
        st, p = self.store, self.plan
        al = st.arena
        me = st.base[self.worker_id]

        def use_bf16(var):
            regs = self._regions(var)
            ld = regs[0].shadow_ld
            self.wptr[var.name] = (me + al.mirror_bf16[var.name], ld)
            for r in self._regions_rotated(var):
                row = r.sl.first // r.cols
                self.copies.append((st.shadow_ptr(r), me + al.mirror_bf16[var.name] + row * ld * 2, r.rows * ld * 2))

        def use_f32(var):
            self.fptr[var.name] = me + al.mirror_f32[var.name]
            for r in self._regions_rotated(var):
                self.copies.append((st.var_ptr(r), me + al.mirror_f32[var.name] + r.sl.first * 4,
                                    (r.sl.n * 4 + 15) // 16 * 16))

        if p.kind == "cnn":
            use_f32(p.convs[0].k)
            use_f32(p.convs[0].b)
            use_bf16(p.convs[1].k)
            use_f32(p.convs[1].b)
        for d in p.dense:
            use_bf16(d.w)
            if d.b is not None:
                use_f32(d.b)

    def _setup_params_mirror(self) -> None:
        """Owner-pushes mode: every tensor the kernels read is this worker's MIRROR (local HBM),
        rewritten by the owners' drains. ``copies`` fills it once after init / restore."""
        st, p = self.store, self.plan
        ml, me = st.mirror, st.mirror_base
        self.late_mark, self.late_layer = None, 0

        def use_bf16(var):
            regs = self._regions(var)
            ld = regs[0].shadow_ld
            self.wptr[var.name] = (me + ml.bf16[var.name], ld)
            for r in self._regions_rotated(var):
                row = r.sl.first // r.cols
                self.copies.append((st.shadow_ptr(r), me + ml.bf16[var.name] + row * ld * 2, r.rows * ld * 2))

        def use_f32(var):
            if var.name not in ml.f32:
                raise UnsupportedGraphError("variable %s (%d elements) has no fp32 mirror" %
                                            (var.name, self.layout[var.name].num_elements))
            self.fptr[var.name] = me + ml.f32[var.name]
            for r in self._regions_rotated(var):
                self.copies.append((st.var_ptr(r), me + ml.f32[var.name] + r.sl.first * 4,
                                    (r.sl.n * 4 + 15) // 16 * 16))

        if p.kind == "cnn":
            use_f32(p.convs[0].k)
            use_f32(p.convs[0].b)
            use_bf16(p.convs[1].k)
            use_f32(p.convs[1].b)
        for d in p.dense:
            use_bf16(d.w)
            if d.b is not None:
                use_f32(d.b)

    def _sync_segments(self, s: StepState):
        """Device copy of the dtf_sync_segment table for this state's gradient buffers."""
        if getattr(s, "sync_table", None) is not None:
            return s.sync_table, s.sync_n
        import ctypes as C
        st, al = self.store, self.store.arena
        segs = []
        for var in self.train_op.var_list:
            gptr = self._grad_ptr(s, var)
            for r in self._regions_rotated(var):
                sg = L.SyncSegment()
                sg.owner, sg.cols, sg.shadow_ld = r.shard, r.cols, r.shadow_ld
                sg.n, sg.first, sg.first_row = r.sl.n, r.sl.first, r.sl.first // r.cols
                sg.grad_local = gptr + r.sl.first * 4
                sg.mailbox_off = al.mailbox_off[(var.name, r.shard)]
                sg.mailbox_stride = al.mailbox_stride[r.shard]
                sg.var_off = r.var_off
                sg.slot0_off = r.slot_off[0] if len(r.slot_off) > 0 else 0
                sg.slot1_off = r.slot_off[1] if len(r.slot_off) > 1 else 0
                sg.shadow_off = r.shadow_off or 0
                sg.mirror_f32_off = al.mirror_f32[var.name] if var.name in self.fptr else 0
                sg.mirror_bf16_off = al.mirror_bf16[var.name] if var.name in self.wptr else 0
                segs.append(sg)
        arr = (L.SyncSegment * len(segs))(*segs)
        raw = bytes(memoryview(arr))
        s.sync_table = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(self.device)
        s.sync_n = len(segs)
        return s.sync_table, s.sync_n

    def _sync_round(self, s: StepState) -> None:
        import ctypes as C
        st, spec = self.store, self.spec
        if self.sync_round_no is None:
            self.sync_round_no = int(st.read_ctrl(self.worker_id)["epoch"])
        self.sync_round_no += 1
        table, n = self._sync_segments(s)
        hyper = K.make_hyper(spec.learning_rate, spec.beta1, spec.beta2, spec.epsilon, spec.momentum,
                             1.0, spec.use_nesterov, False)
        bases = (C.c_ulonglong * st.num_shards)(*st.base)
        rc = L.lib().dtf_sync_round(spec.code, table.data_ptr(), n, self.worker_id, st.num_shards,
                                    self.sync_round_no, C.byref(hyper), bases, s.counter.data_ptr() + 8,
                                    s.step_dev.data_ptr(), st.arena.step_shard, L.current_stream())
        L.check(rc, "dtf_sync_round")
        K._count()

    # -- kernels of one step ------------------------------------------------------------------------
    def _forward(self, s: StepState) -> None:
        p = self.plan
        if p.kind == "cnn":
            scale = 1.0 / 255.0 if s.x_dev.dtype == torch.uint8 else 1.0
            c1, c2 = p.convs
            if s.zero_copy:
                # the first kernel reads the batch straight from pinned host memory (and mirrors it
                # into x_dev for conv1's backward): no H2D copy node in front of the step
                K_conv1_fwd(s.x_pin, self.fptr[c1.k.name], self.fptr[c1.b.name], s.pool1p, s.amax1, scale,
                            x_copy=s.x_dev, ingest=True, y_src=s.y_pin.data_ptr(), y_dst=s.y_dev.data_ptr(),
                            y_bytes=s.y_bytes16)
            else:
                K_conv1_fwd(s.x_dev, self.fptr[c1.k.name], self.fptr[c1.b.name], s.pool1p, s.amax1, scale)
            # programmatic dependent launch for the FORWARD chain only (one stream, kernel after kernel):
            # each kernel's prologue (TMEM allocation, mbarrier init, descriptor prefetch) overlaps its
            # predecessor's tail. Across the multi-stream backward it measured slower, so it ends with
            # the loss kernel.
            # (not beside an in-graph drain: the early-launched kernels' resident CTAs delayed the drain
            # from the head of the step into the GEMM / head phase: 0.139 vs 0.127 ms/step at 8 GPUs)
            if self.pdl_forward and not self.pdl and not (self.use_mailbox and self.my_shards):
                L.lib().dtf_set_pdl(1)
            K_conv2_fwd(s.pool1p, self.wptr[c2.k.name][0], self.fptr[c2.b.name],
                        s.act[0].view(s.B, 7, 7, 64), s.amax2)
        else:
            src = s.x_pin if s.zero_copy else s.x_dev     # pinned host memory is device-addressable
            if s.x_dev.dtype == torch.uint8:
                K.u8_to_bf16(src, s.act[0])
            else:
                K.f32_to_bf16(src, s.act[0])
        nl = len(p.dense)
        fused_head = self._head_is_fused()
        for i, d in enumerate(p.dense):
            if getattr(s, "pull_late", None) is not None and i == self.late_layer:
                torch.cuda.current_stream().wait_event(s.pull_late)   # the dense stack's weights have landed
                s.pull_late = None
            if fused_head and i == nl - 1:
                break                      # the logits layer is part of the fused head kernel (_loss)
            wp, ld = self.wptr[d.w.name]
            out = s.act[i + 1] if i + 1 < nl else s.logits
            bias = _PtrTensor(self.fptr[d.b.name]) if d.b is not None else None
            K.gemm(s.act[i], None, out, s.B, s.dims[i + 1], s.dims[i], b_mn=True, bias=bias,
                   relu=d.relu, b_ptr=wp, ldb=ld, **self._plan_kw(s, f"fwd{i}"))

    def _head_is_fused(self) -> bool:
        """The one-kernel head is a latency optimisation for the reference models' small heads
        (128 x 1024 -> 10, 100 x 100 -> 10). A wide hidden layer (the 8192-unit MLP) keeps the
        tensor-core GEMMs: the CUDA-core head would need B/4 CTAs x H atomics for the bias gradient."""
        p = self.plan
        nl = len(p.dense)
        if not self.fused_head or nl < 2 or p.num_classes > 16:
            return False
        if int(p.dense[-1].w.shape[0]) > 2048:
            return False
        return self.wptr[p.dense[-1].w.name][1] == 16

    def _plan_kw(self, s: StepState, key: str) -> dict:
        plan = s.gemm_plans.get(key)
        if plan is None or not self.split_k:
            return {}
        kw = dict(block_n=plan["bn"])
        if plan["splits"] > 1:
            kw.update(splits=plan["splits"], ws=plan["ws"], counters=plan["counters"])
        return kw

    def _loss(self, s: StepState) -> None:
        try:
            self._loss_body(s)
        finally:
            if self.pdl_forward and not self.pdl:
                L.lib().dtf_set_pdl(0)

    def _loss_body(self, s: StepState) -> None:
        p = self.plan
        nl = len(p.dense)
        if getattr(s, "y_ready", None) is not None:
            torch.cuda.current_stream().wait_event(s.y_ready)
        if self._head_is_fused():
            d = p.dense[-1]
            wp, ld = self.wptr[d.w.name]
            K.head_fused(s.act[nl - 1], s.dims[nl - 1], wp, ld,
                         self.fptr[d.b.name] if d.b is not None else None, s.y_dev, s.B, p.num_classes,
                         s.logits, s.z["loss"],
                         dlogits=s.dlogits if s.train else None,
                         dbias_accum=s.z[f"gb{nl - 1}"] if s.train else None,
                         dact=s.dact[nl - 1] if s.train else None,
                         dprev_bias_accum=s.z[f"gb{nl - 2}"] if (s.train and p.dense[nl - 2].b is not None) else None,
                         dw_accum=s.gw[nl - 1] if (s.train and s.head_dw) else None)
            return
        dl = s.dlogits if s.train else self._scratch_dlogits(s)
        gb = s.z[f"gb{len(p.dense) - 1}"] if s.train else None
        if p.loss_kind == "sparse_mean":
            K.softmax_xent_sparse(s.logits, s.y_dev, s.z["loss"], dl, gb, p.num_classes)
        else:
            K.softmax_xent_dense(s.logits, s.y_dev, s.z["loss"], dl, gb, p.num_classes)

    def _scratch_dlogits(self, s: StepState) -> torch.Tensor:
        if not hasattr(s, "_dl"):
            s._dl = torch.zeros((s.B, pad8(self.plan.num_classes)), dtype=torch.bfloat16, device=self.device)
        return s._dl

    def _backward(self, s: StepState) -> None:
        """Backward pass on three streams inside the captured graph (fork / join through events):

        * main   — the activation-gradient chain (dX GEMMs, pooling / conv gradients): the critical
                   path, highest priority;
        * side   — weight-gradient GEMMs and conv2's filter gradient: they depend only on tensors
                   that already exist, so they run beside the chain;
        * apply  — the push(+apply) of each dense kernel as soon as its gradient exists (and the
                   dX GEMM that reads the same weights has been ordered before it). The big one
                   (12.8 MB of Adam state for dense/kernel) hides behind the convolution backward
                   instead of trailing the step, and never delays a weight-gradient GEMM."""
        p = self.plan
        nl = len(p.dense)
        main = torch.cuda.current_stream()
        sync = self.coop_sync
        par = self.overlap and not sync
        side = self.side_stream if par else main
        early = self.early_apply
        if early is None:
            early = self.use_mailbox or self.fused_apply or not all(self.store.is_local)
        apl = self.apply_stream if (par and early) else main
        applied: List[G.Variable] = []
        late_dw: List[Any] = []

        def after(stream):
            ev = torch.cuda.Event()
            ev.record(stream)
            return ev

        def fork(dst, ev=None):
            if dst is not main:
                dst.wait_event(ev if ev is not None else after(main))

        dz = s.dlogits
        head_fused = self._head_is_fused()
        pre_dw: Dict[int, Any] = {}
        fork(side)                               # dlogits (and every activation) is ready
        s.reported = bool(self.early_report and self._uses_sink(s))
        if s.reported:
            with torch.cuda.stream(side):
                # the loss is final: hand loss + global_step to the host now
                # (owner-pushes mode: global_step is read from my mirror's header, offset 0, which the
                # step-owning shard's drain keeps fresh - no NVLink round trip)
                K.step_report(self.store.mirror_base if self.use_mirror else self.store.step_ctrl,
                              self.worker_id, not self.use_mailbox, s.step_dev.data_ptr(), self._sink(s))
        for i in range(nl - 1, -1, -1):
            d = p.dense[i]
            din, dout = s.dims[i], s.dims[i + 1]
            fused = self._fused_apply_spec(d.w) if (par and early and self.fused_apply
                                                    and not self.use_mailbox) else None
            push_spec = self._fused_push_spec(d.w) if (par and early and self.fused_push
                                                       and self.use_mailbox) else None
            dz_i = dz
            dw_done = pre_dw.pop(i, None)
            if fused is None and push_spec is None and dw_done is None:
                with torch.cuda.stream(side):
                    if head_fused and i == nl - 1 and par and not (self.fused_apply and not self.use_mailbox) \
                            and not (self.fused_push and self.use_mailbox):
                        # the fused head already produced dZ_{i-1}: issue the LARGER weight-gradient
                        # GEMM of the layer below first, so its (big) apply can start earliest
                        K.gemm(s.act[i - 1], s.dact[i], s.gw[i - 1], s.dims[i - 1], s.dims[i], s.B,
                               a_mn=True, b_mn=True)
                        pre_dw[i - 1] = after(side)
                    if head_fused and i == nl - 1 and s.head_dw:
                        dw_done = after(side) if par else None   # dW_i came out of the fused head kernel
                    else:
                        # dW_i = act_i^T dZ_i
                        K.gemm(s.act[i], dz, s.gw[i], din, dout, s.B, a_mn=True, b_mn=True)
                        if par:
                            dw_done = after(side)
            wp, ld = self.wptr[d.w.name]
            if i > 0 and i == nl - 1 and head_fused:
                dz = s.dact[i]                   # dZ_{i-1} and db_{i-1} came out of the fused head kernel
            elif i > 0:
                # dZ_{i-1} = (dZ_i W_i^T) * relu'(act_i);  db_{i-1} = colsum
                K.gemm(dz, None, s.dact[i], s.B, din, dout, mask=s.act[i], b_ptr=wp, ldb=ld,
                       **self._plan_kw(s, f"dx{i}"))
                K.colsum_bf16(s.dact[i], s.B, din, s.z[f"gb{i - 1}"])
                dz = s.dact[i]
                fork(side)                       # dZ_{i-1} is ready
            elif p.kind == "cnn" and self.fused_pool_bwd:
                # dX GEMM whose epilogue IS conv2's ReLU + max-pool gradient: the pooled-map gradient goes
                # straight to its argmax positions in the zero-bordered conv2 output gradient (and the
                # conv2 bias gradient is summed) - no dpool2 round trip, no pool2_bwd kernel on the chain
                kw = self._plan_kw(s, "dx0")
                if os.environ.get("DTF_B200_POOL_SPLITS") == "1":
                    kw = dict(block_n=64)
                K.gemm(dz, None, None, s.B, din, dout, b_ptr=wp, ldb=ld, mask=s.act[0],
                       pool=dict(out=s.dconv2p, amax=s.amax2, dbias=s.z["gkb2"]), **kw)
            elif p.kind == "cnn":
                K.gemm(dz, None, s.dpool2, s.B, din, dout, b_ptr=wp, ldb=ld, **self._plan_kw(s, "dx0"))
            small_push = (self.use_mailbox and fused is None and push_spec is None
                          and self.layout[d.w.name].num_elements * 4 < self.early_push_min_bytes)
            if small_push and dw_done is not None:
                late_dw.append(dw_done)      # a small gradient rides in the step's final, publishing push
            if par and early and not small_push:
                # W_i may be read in place from the local shard by the dX GEMM above: order the
                # apply after it, and after the gradient it consumes
                fork(apl)
                if dw_done is not None:
                    apl.wait_event(dw_done)
                with torch.cuda.stream(apl):
                    if fused is not None:
                        # dW_i GEMM whose epilogue IS the push+apply: the gradient never exists in
                        # memory, the TMEM accumulator goes straight into the owner's var / m / v
                        K.gemm(s.act[i], dz_i, None, din, dout, s.B, a_mn=True, b_mn=True, apply=fused)
                    elif push_spec is not None:
                        # dW_i GEMM whose epilogue IS the push: rows are stored straight into this
                        # worker's mailbox buffer on each owning shard (published at the end of the step)
                        K.gemm(s.act[i], dz_i, None, din, dout, s.B, a_mn=True, b_mn=True, scatter=push_spec)
                    else:
                        self._apply_vars(s, [d.w], finish=False)
                applied.append(d.w)
        clean_ws = self._self_cleans(s) and s.zero.ws_off is not None
        if p.kind == "cnn":
            c1, c2 = p.convs
            scale = 1.0 / 255.0 if s.x_dev.dtype == torch.uint8 else 1.0
            if not self.fused_pool_bwd:
                K.pool2_bwd(s.dpool2, s.act[0].view(s.B, 7, 7, 64), s.amax2, s.dconv2p, s.z["gkb2"])
            fork(side)                           # dconv2p is ready
            with torch.cuda.stream(side):
                K.conv2_wgrad(s.pool1p, s.dconv2p, s.z["gk2"])
                if clean_ws:                     # every split-K GEMM of the step is done: re-zero
                    K.fill_u32(s.zero.buf[s.zero.ws_off:], 0)   # their workspaces beside the conv backward
                    clean_ws = False
            K.conv2_dgrad(s.dconv2p, self.wptr[c2.k.name][0], s.dpool1)
            K.conv1_bwd(s.x_dev, s.dpool1, s.pool1p, s.amax1, s.z["gk1"], s.z["gkb1"], scale)
        if clean_ws:
            fork(side)
            with torch.cuda.stream(side):
                K.fill_u32(s.zero.buf[s.zero.ws_off:], 0)
        if par:                                  # join
            main.wait_event(after(side))
            if apl is not main:
                main.wait_event(after(apl))
        if sync:
            return                   # the round kernel is launched after the (captured) compute part
        rest = [v for v in self.train_op.var_list if v not in applied]
        self._apply_vars(s, rest, finish=True)

    def _grad_ptr(self, s: StepState, var: G.Variable) -> int:
        p = self.plan
        if p.kind == "cnn":
            c1, c2 = p.convs
            m = {c1.k.name: "gk1", c1.b.name: "gkb1", c2.k.name: "gk2", c2.b.name: "gkb2"}
            if var.name in m:
                return s.z[m[var.name]].data_ptr()
        for i, d in enumerate(p.dense):
            if var is d.w:
                return s.gw[i].data_ptr()
            if d.b is not None and var is d.b:
                return s.z[f"gb{i}"].data_ptr()
        raise KeyError(var.name)

    def _fused_apply_spec(self, var: G.Variable) -> Optional[dict]:
        """Arguments of the GEMM's fused push+apply epilogue for a 2-D variable, or None when the
        variable's slicing cannot be expressed as uniform row slices."""
        regs = sorted(self._regions(var), key=lambda r: r.sl.first)
        if len(var.shape) != 2 or not regs or len(regs) > 8:
            return None
        M, N = int(var.shape[0]), int(var.shape[1])
        per = regs[0].rows if len(regs) > 1 else M
        st, spec = self.store, self.spec
        slices: List[dict] = []
        for k, r in enumerate(regs):
            if r.sl.first != k * per * N or (k + 1 < len(regs) and r.rows != per):
                return None
            slices.append(dict(var=st.var_ptr(r), slot0=st.slot_ptr(r, 0), slot1=st.slot_ptr(r, 1),
                               shadow=st.shadow_ptr(r)))
        hyper = K.make_hyper(spec.learning_rate, spec.beta1, spec.beta2, spec.epsilon, spec.momentum,
                             1.0, spec.use_nesterov, False)
        return dict(opt=spec.code, hyper=hyper, rows_per_slice=per, shadow_ld=regs[0].shadow_ld,
                    ctrl_powers=st.powers_ctrl, slices=slices)

    # -- "owner applies" push (mailboxes + the PS tasks' apply daemons) --------------------------------
    def _mbox_setup(self) -> None:
        """Start every shard's apply daemon (idempotent: the first worker to ask wins) and reset
        this worker's sequence numbers on every owner."""
        st, al, me, spec = self.store, self.store.arena, self.worker_id, self.spec
        hyper = dict(lr=float(spec.learning_rate), beta1=float(spec.beta1), beta2=float(spec.beta2),
                     epsilon=float(spec.epsilon), momentum=float(spec.momentum),
                     use_nesterov=bool(spec.use_nesterov))
        ml = st.mirror if self.use_mirror else None
        self._drain_desc: Dict[int, dict] = {}
        for sh in range(st.num_shards):
            items = []
            mitems = []
            for name in self.trainable_names:
                for r in al.regions.get(name, []):
                    if r.shard != sh:
                        continue
                    if ml is not None:
                        mitems.append((ml.bf16.get(name, -1), ml.f32.get(name, -1), int(r.sl.first)))
                    items.append(dict(var_off=r.var_off,
                                      slot0_off=r.slot_off[0] if len(r.slot_off) > 0 else -1,
                                      slot1_off=r.slot_off[1] if len(r.slot_off) > 1 else -1,
                                      shadow_off=r.shadow_off if r.shadow_off is not None else -1,
                                      grad_off=al.mailbox_off[(name, sh)], n=int(r.sl.n), cols=int(r.cols),
                                      shadow_ld=int(r.shadow_ld)))
            desc = dict(opt=int(spec.code), items=items, slot_stride=int(al.mailbox_stride[sh]),
                        depth=int(al.mailbox_depth), n_workers=int(al.sync_workers), hyper=hyper,
                        owns_step=bool(al.step_shard == sh), flags_off=int(al.mbox_flags_off[sh]),
                        poll_us=int(self.mailbox_poll_us), shard=int(sh),
                        sync_R=int(self.sync_opt.replicas_to_aggregate) if self.sync_generic else 0,
                        mirror_items=[list(m) for m in mitems] if ml is not None else None)
            self._drain_desc[sh] = desc
            self.client.call(sh, "mbox_start", desc)
        zero = np.zeros(1, np.uint32)
        for sh in range(st.num_shards):
            f = st.mbox_flags_ptr(sh)
            # produced first: "nothing pending" holds at every instant of the reset
            st._h2d(f + L.MBOX_OFF_PRODUCED + 4 * me, zero)
            st._h2d(f + L.MBOX_OFF_CONSUMED + 4 * me, zero)
            st._h2d(f + L.MBOX_OFF_DONE + 4 * me, zero)
        self._pushes = 0
        self._min_consumed = 0
        self.mbox_seq = torch.zeros(4, dtype=torch.int32, device=self.device)
        # the push kernel's last CTA stores "how many of my pushes every owner has applied" straight
        # into this pinned host word (UVA: pinned memory is device-addressable), so admitting the
        # next step costs the host one memory read, not a stream sync or a copy
        self.mbox_host = torch.zeros(16, dtype=torch.int32).pin_memory()
        self._mbox_host_np = self.mbox_host.numpy()
        torch.cuda.current_stream(self.device).synchronize()   # (never device-wide: see dtf_arena_alloc)
        if self.use_mirror:
            st.register_mirror(me)
        # shards whose arena lives in THIS process on THIS GPU: their drains become nodes of my step graph
        self.my_shards = []
        if self.sync_generic:
            # two device words: the round this worker last contributed to (none yet) and its local_step
            self.sync_words = torch.tensor([-1, 0, 0, 0], dtype=torch.int32, device=self.device)
        if self.graph_drain and not self.sync_generic:
            self.my_shards = [sh for sh in range(st.num_shards) if st.same_process[sh] and st.is_local[sh]]
            for sh in self.my_shards:
                self.client.call(sh, "mbox_attach", me, True, self.drain_backstop_us)
            self._drain_scratch = torch.zeros(64 * max(1, len(self.my_shards)), dtype=torch.int32,
                                              device=self.device)
            torch.cuda.current_stream(self.device).synchronize()   # (never device-wide: see dtf_arena_alloc)

    def _push_vars(self, s: StepState, variables: Sequence[G.Variable], publish: bool) -> None:
        """Posted stores of the gradients into this worker's mailbox on every owning shard;
        ``publish`` releases the push (one sequence number per owner) and must be the step's last."""
        st, al, me = self.store, self.store.arena, self.worker_id
        items = []
        zb = s.zero.buf.data_ptr()
        ze = zb + s.zero.buf.numel() * 4
        clean = self._self_cleans(s)
        for var in variables:
            gptr = self._grad_ptr(s, var)
            for r in self._regions_rotated(var):
                dst0 = st.mbox_slot_ptr(r.shard, me, 0) + st.mbox_slice_off(var.name, r.shard)
                # an accumulated gradient (bias / conv filter sums) is cleared by the kernel that consumes it
                items.append((gptr + r.sl.first * 4, dst0, (r.sl.n * 4 + 15) // 16 * 16,
                              al.mailbox_stride[r.shard], int(clean and zb <= gptr < ze),
                              r.shard if self.device_flow else -1))
        flags = [st.mbox_flags_ptr(sh) for sh in range(st.num_shards)]
        K.mailbox_push(items, al.mailbox_depth, me, flags, publish, self.mbox_seq.data_ptr(),
                       s.counter.data_ptr(), self.mbox_host.data_ptr() if publish else None,
                       st.step_ctrl if publish else None, s.step_dev.data_ptr() if publish else None,
                       sink=self._sink(s) if (publish and self._uses_sink(s)
                                              and not getattr(s, "reported", False)) else None,
                       my_hdr=st.mirror_base if self.use_mirror else None,
                       sync_tag=(self.sync_words.data_ptr() + 4) if (publish and self.sync_generic) else None)

    def _drain_local(self, s: StepState) -> None:
        """Device-driven owner-side apply: drain the mailboxes of the shard(s) living in this process
        (claim lock + snapshot, then apply from local HBM and broadcast into the workers' mirrors).
        A node of the step graph; whatever has been published when it runs gets applied."""
        st, spec = self.store, self.spec
        hyper = K.make_hyper(spec.learning_rate, spec.beta1, spec.beta2, spec.epsilon, spec.momentum,
                             1.0, spec.use_nesterov, False)
        for k, sh in enumerate(self.my_shards):
            desc = self._drain_desc[sh]
            base = st.base[sh]

            def at(off):
                return None if off is None or off < 0 else base + int(off)

            entries = [dict(var=at(it["var_off"]), slot0=at(it["slot0_off"]), slot1=at(it["slot1_off"]),
                            shadow=at(it["shadow_off"]), grad=at(it["grad_off"]), n=int(it["n"]), first=0,
                            cols=int(it["cols"]), shadow_ld=int(it["shadow_ld"])) for it in desc["items"]]
            mitems = K.make_mirror_items([tuple(m) for m in desc["mirror_items"]]) \
                if desc.get("mirror_items") is not None else None
            scratch = self._drain_scratch.data_ptr() + 256 * k
            cap = self.drain_ctas
            if cap < 0:
                chunks = sum(-(-int(it["n"]) // 4096) for it in desc["items"])
                # a short drain is compute-heavy (one Adam update with IEEE-grade division and sqrt per
                # pending gradient per element): one CTA slot per SM gets it out of the way of the forward
                # pass soonest (2 GPUs, measured: cap 64 0.139, cap 148 0.129, uncapped 0.130 ms/step)
                cap = 148 if chunks <= 1024 else 0
            K.mailbox_drain(spec.code, K.make_apply_items(entries), len(entries), int(desc["slot_stride"]),
                            int(desc["depth"]), int(desc["n_workers"]), hyper, base,
                            base if desc["owns_step"] else None, base + int(desc["flags_off"]),
                            scratch, scratch + 128, None, mirror_items=mitems, shard_index=sh,
                            max_ctas=cap)

    def _mbox_consumed(self) -> int:
        """min over owners of how many of MY gradients they have applied (host-side read)."""
        st, me = self.store, self.worker_id
        if self.use_mirror:     # one local copy: the owners keep these words fresh in my mirror's header
            return int(st.mirror_hdr_u32(L.MIRROR_OFF_CONSUMED, st.num_shards).min())
        return min(int(st._d2h(st.mbox_flags_ptr(sh) + L.MBOX_OFF_CONSUMED + 4 * me, 4, np.uint32)[0])
                   for sh in range(st.num_shards))

    def _mbox_wait(self, needed: int) -> None:
        """Flow control / flush: block until every owner has applied ``needed`` of my pushes.
        Fast path: the host word the push kernels keep fresh. Slow path (nothing of mine is in
        flight, so nothing will refresh that word): read the owners' counters over the mapping."""
        if self._min_consumed >= needed:
            return
        from ..parallel.ps_core import AbortedError
        t0 = time.monotonic()
        while True:
            v = int(self._mbox_host_np[0])
            if v > self._min_consumed:
                self._min_consumed = v
            if self._min_consumed >= needed:
                return
            if self.stream.query():
                self._min_consumed = max(self._min_consumed, self._mbox_consumed())
                if self._min_consumed >= needed:
                    return
                time.sleep(20e-6)
            if time.monotonic() - t0 > self.mailbox_timeout:
                raise AbortedError("a parameter server stopped applying gradients: %d of %d pushes of "
                                   "worker %d consumed after %.0f s" % (self._min_consumed, needed,
                                                                         self.worker_id, self.mailbox_timeout))

    def _mbox_admit(self) -> None:
        """Called right before a training step is enqueued: push k reuses the mailbox buffer of
        push k - depth, which the owners must have applied. Bounds staleness to ``depth`` pushes.
        With mirrors the push kernels enforce this themselves on the device (they poll the
        consumed[] words the owners keep fresh in this worker's mirror header), so the host just
        counts; it only holds back when it is a whole ``max_queue`` of steps ahead of the device."""
        if self.device_flow:
            if self._pushes - self._min_consumed > self.max_queue:
                self._mbox_wait(self._pushes + 1 - self.max_queue)
        else:
            self._mbox_wait(self._pushes + 1 - self.store.arena.mailbox_depth)
        self._pushes += 1

    def flush_pushes(self) -> None:
        """Wait until every gradient this worker has pushed is applied (exact global_step reads)."""
        if self.use_mailbox and not self.sync_generic:   # (a round may never complete once replicas stop)
            self._mbox_wait(self._pushes)

    def _fused_push_spec(self, var: G.Variable) -> Optional[dict]:
        """Arguments of the GEMM's scatter epilogue (rows -> this worker's mailbox buffer on the slice's
        owner) for a 2-D variable, or None when its slicing is not uniform row slices."""
        regs = sorted(self._regions(var), key=lambda r: r.sl.first)
        if len(var.shape) != 2 or not regs or len(regs) > 8:
            return None
        M, N = int(var.shape[0]), int(var.shape[1])
        per = regs[0].rows if len(regs) > 1 else M
        st, al, me = self.store, self.store.arena, self.worker_id
        slices = []
        for k, r in enumerate(regs):
            if r.sl.first != k * per * N or (k + 1 < len(regs) and r.rows != per):
                return None
            slices.append((st.mbox_slot_ptr(r.shard, me, 0) + st.mbox_slice_off(var.name, r.shard),
                           al.mailbox_stride[r.shard]))
        return dict(rows_per_slice=per, depth=al.mailbox_depth, seq=self.mbox_seq.data_ptr(), slices=slices)

    def _sink(self, s: StepState):
        return K.make_sink(s.z["loss"].data_ptr(), s.res_pin.data_ptr(), s.res_seq_dev.data_ptr(),
                           clear_loss=self._self_cleans(s))

    def _apply_vars(self, s: StepState, variables: Sequence[G.Variable], finish: bool) -> None:
        """push+apply of ``variables`` on the current stream; ``finish`` adds TF's _finish (beta
        powers) and the global_step atomic and must be the step's last apply."""
        if self.use_mailbox:
            self._push_vars(s, variables, publish=finish)
            return
        st, spec = self.store, self.spec
        reported = bool(getattr(s, "reported", False))
        sink = self._sink(s) if (finish and self._uses_sink(s) and not reported) else None
        entries = []
        zb = s.zero.buf.data_ptr()
        ze = zb + s.zero.buf.numel() * 4
        clean = self._self_cleans(s)
        for var in variables:
            gptr = self._grad_ptr(s, var)
            for r in self._regions_rotated(var):
                entries.append(dict(var=st.var_ptr(r), slot0=st.slot_ptr(r, 0), slot1=st.slot_ptr(r, 1),
                                    shadow=st.shadow_ptr(r), grad=gptr + r.sl.first * 4, n=r.sl.n,
                                    first=0, cols=r.cols, shadow_ld=r.shadow_ld,
                                    zero_grad=int(clean and zb <= gptr < ze)))
        hyper = K.make_hyper(spec.learning_rate, spec.beta1, spec.beta2, spec.epsilon, spec.momentum,
                             1.0, spec.use_nesterov, self.atomic_sgd and spec.kind == "sgd")
        chunks = [entries[i:i + L.MAX_APPLY_ITEMS] for i in range(0, len(entries), L.MAX_APPLY_ITEMS)]
        if not chunks and finish:
            chunks = [[]]
        for ci, ch in enumerate(chunks):
            last = finish and ci == len(chunks) - 1
            items = K.make_apply_items(ch)
            K.push_apply(spec.code, items, len(ch), hyper, st.powers_ctrl, st.step_ctrl,
                         (2 if reported else 1) if last else 0,
                         self.worker_id, s.step_dev.data_ptr() if (last and not reported) else None,
                         s.counter.data_ptr(), sink=sink if last else None,
                         max_ctas=0 if finish else self.early_apply_ctas)

    def _pull(self, s: StepState) -> None:
        if self.coop_sync or self.use_mirror:
            if self.mirror_filled:
                return
            self.mirror_filled = True
            if self.use_mirror:
                self._mirror_seed()
        s.pull_late = None
        mark = getattr(self, "late_mark", None)
        if not self.coop_sync and self.overlap and mark is not None and len(self.copies) > mark:
            main = torch.cuda.current_stream()
            ev = torch.cuda.Event()
            ev.record(main)
            self.side_stream.wait_event(ev)
            with torch.cuda.stream(self.side_stream):
                K.pull_bulk(self.copies[mark:])
                s.pull_late = torch.cuda.Event()
                s.pull_late.record(self.side_stream)
            if mark > 0:
                K.pull_bulk(self.copies[:mark], self.store.step_ctrl, s.step_dev.data_ptr() + 8)
        elif self.copies:
            K.pull_bulk(self.copies, self.store.step_ctrl, s.step_dev.data_ptr() + 8)

    def _mirror_seed(self) -> None:
        """First use of the mirror (after init / restore): global_step into its header; the
        parameters follow through ``copies`` on the stream."""
        st = self.store
        gs = np.asarray([st.read_global_step()], np.uint64)
        st._h2d(st.mirror_base + L.MIRROR_OFF_GLOBAL_STEP, gs)

    def _self_cleans(self, s: StepState) -> bool:
        """True when every buffer of the zeroed region is reset by its consumer (worker-side apply
        with a result sink, every dense layer has a bias so each column sum is consumed)."""
        if self.use_mailbox and not (self.early_report and not self.fused_push):
            return False      # the loss is cleared by the early report; the scatter epilogue has no source to clear
        return (self.self_clean and self._uses_sink(s) and self.clear_at_end
                and not (self.fused_apply) and all(d.b is not None for d in self.plan.dense))

    def _uses_sink(self, s: StepState) -> bool:
        return self.host_sink and s.train and not self.coop_sync

    def _enqueue_step(self, s: StepState, with_io: bool = True) -> None:
        """Everything one ``sess.run(train_op)`` does on the device, in stream order."""
        if self.pdl_chain and not self.pdl:
            # PDL only along the critical chain (the step's main stream): prologues overlap the
            # predecessor's tail there; the side / apply streams keep full kernel edges
            L.pdl_main_stream = torch.cuda.current_stream().cuda_stream
            try:
                self._enqueue_step_body(s, with_io)
            finally:
                L.pdl_main_stream = None
                L.lib().dtf_set_pdl(0)
            return
        self._enqueue_step_body(s, with_io)

    def _enqueue_step_body(self, s: StepState, with_io: bool = True) -> None:
        s.zero_copy = bool(with_io and self.zero_copy_input)
        s.y_ready = None
        if s.zero_copy and self.plan.kind != "cnn":
            # labels: a copy node on the side stream, parallel to the first kernel; the loss waits for
            # it (the CNN's conv1 kernel mirrors the labels itself)
            main = torch.cuda.current_stream()
            ev = torch.cuda.Event()
            ev.record(main)
            self.side_stream.wait_event(ev)
            with torch.cuda.stream(self.side_stream):
                s.y_dev.copy_(s.y_pin, non_blocking=True)
                s.y_ready = torch.cuda.Event()
                s.y_ready.record(self.side_stream)
        elif with_io and not s.zero_copy:
            s.io_dev.copy_(s.io_pin, non_blocking=True)
        clear_end = s.train and self.clear_at_end and not self.coop_sync
        if not clear_end:
            K.fill_u32(s.zero.buf, 0)
        if s.train and self.sync_generic:
            # the token: wait (on the device, polling my own mirror header) until the round I last
            # contributed to is complete on every shard, and adopt the current round as local_step
            K.sync_token_wait(self.store.mirror_base, self.store.num_shards, self.sync_words.data_ptr(),
                              self.sync_words.data_ptr() + 4, None)
        drain_ev = None
        if s.train and self.my_shards and self.use_mailbox and self.drain_at == "start":
            # the owner-side apply of whatever the workers have published, beside the forward pass
            main = torch.cuda.current_stream()
            ev = torch.cuda.Event()
            ev.record(main)
            self.apply_stream.wait_event(ev)
            with torch.cuda.stream(self.apply_stream):
                self._drain_local(s)
                drain_ev = torch.cuda.Event()
                drain_ev.record(self.apply_stream)
        with nvtx_range("dtf/pull"):
            self._pull(s)
        with nvtx_range("dtf/forward"):
            self._forward(s)
            self._loss(s)
        if s.train:
            with nvtx_range("dtf/backward_push_apply"):
                self._backward(s)      # includes the push+apply of every variable
            if drain_ev is not None:
                torch.cuda.current_stream().wait_event(drain_ev)     # join (the apply stream is ordered anyway)
            if self.my_shards and self.use_mailbox and self.drain_at == "end":
                self._drain_local(s)
        if with_io and not self._uses_sink(s):
            s.out_pin[:1].copy_(s.z["loss"][:1], non_blocking=True)
            if not self.coop_sync or not s.train:
                s.step_pin.copy_(s.step_dev, non_blocking=True)
        if clear_end and not self._self_cleans(s):
            # the accumulators were consumed by the applies above: clearing them now (instead of at
            # the start of the next step) takes the fill kernel off the next step's critical path
            K.fill_u32(s.zero.buf, 0)

    def _enqueue_sync_tail(self, s: StepState) -> None:
        """Synchronous mode: the cross-GPU round (cooperative kernel, launched outside the captured
        graph) and the read-back of the new global step."""
        self._sync_round(s)      # reduce-scatter + mean + apply + all-gather, one kernel
        s.step_pin.copy_(s.step_dev, non_blocking=True)

    # -- state management ---------------------------------------------------------------------------
    def _ensure_plan(self, loss: G.Tensor, train_op=None) -> None:
        if self.plan is None:
            self.plan = parse_plan(loss)
            if train_op is not None:
                self.train_op = train_op
            else:
                for n in self.graph.nodes:
                    if type(n).__name__ == "TrainOp":
                        self.train_op = n
            self._setup_params()
        elif train_op is not None and self.train_op is None:
            self.train_op = train_op

    def _state(self, B: int, train: bool, x: np.ndarray, y: np.ndarray) -> StepState:
        key = (B, train, x.dtype.char, y.dtype.char)
        s = self.states.get(key)
        if s is None:
            xdt = torch.uint8 if x.dtype == np.uint8 else torch.float32
            ydt = torch.int32 if self.plan.loss_kind == "sparse_mean" else torch.float32
            with torch.cuda.stream(self.stream):
                s = StepState(self, B, train, xdt, ydt)
            self.stream.synchronize()
            self.states[key] = s
        return s

    def _stage(self, s: StepState, x: np.ndarray, y: np.ndarray) -> None:
        xn, yn = s.x_np, s.y_np
        np.copyto(xn, x.reshape(xn.shape), casting="same_kind" if x.dtype != xn.dtype else "no")
        np.copyto(yn, y.reshape(yn.shape), casting="unsafe")

    def _launch(self, s: StepState) -> None:
        sync = self.coop_sync and s.train
        if s.graph_exec and not sync:
            # hot path: one cudaGraphLaunch on our stream, no Python stream-context bookkeeping
            if self._graph_launch(s.graph_exec, self._stream_handle) != 0:
                L.check(-1, "dtf_graph_launch")
            return
        with torch.cuda.stream(self.stream):
            if not self.use_cuda_graph:
                self._enqueue_step(s)
                if sync:
                    self._enqueue_sync_tail(s)
                return
            if s.graph is None:
                # one eager step first (sets kernel attributes, warms the allocator), then capture
                before = K.LAUNCHES
                self._enqueue_step(s)
                s.launches_per_step = K.LAUNCHES - before
                self.stream.synchronize()
                g = torch.cuda.CUDAGraph()
                # thread-local capture mode: the PS apply daemon (a native host thread of this
                # process) keeps polling and launching on its own stream while we capture
                with torch.cuda.graph(g, stream=self.stream, capture_error_mode="thread_local"):
                    self._enqueue_step(s)
                s.graph = g
                try:
                    s.graph_exec = int(g.raw_cuda_graph_exec())
                except Exception:
                    s.graph_exec = 0          # older torch: fall back to CUDAGraph.replay()
                if sync:
                    self._enqueue_sync_tail(s)
                return  # the eager step already did this run's work
            s.graph.replay()
            if sync:
                self._enqueue_sync_tail(s)

    # -- Session.run entry ---------------------------------------------------------------------------
    def run(self, fetches: Sequence[G.Tensor], feeds: Dict[G.Tensor, Any]) -> List[Any]:
        from ..train.optimizers import TrainOp
        train_ops = [f for f in fetches if isinstance(f, TrainOp)]
        plain = [f for f in fetches if not isinstance(f, TrainOp)]
        gs_var = G.get_global_step(self.graph)
        if not train_ops and all(f is gs_var for f in plain):
            self.flush_pushes()
            step = self.store.read_global_step()
            return [self._as_step(step) for _ in fetches]
        top = train_ops[0] if train_ops else None
        self._runs += 1
        if self.health_every > 0 and self._runs % self.health_every == 0:
            self.check_health()
        loss_node = top.loss if top is not None else self._find_loss(plain)
        self._ensure_plan(loss_node, top)
        p = self.plan
        if p.input_node not in feeds or p.labels_node not in feeds:
            raise ValueError("feed the model input and labels (or initialise the iterator)")
        x, y = np.asarray(feeds[p.input_node]), np.asarray(feeds[p.labels_node])
        B = int(x.shape[0])
        if x.dtype not in (np.uint8, np.float32):
            x = x.astype(np.float32)
        s = self._state(B, top is not None, x.reshape(B, -1), y)
        prof = self.host_prof
        t_a = time.perf_counter() if prof is not None else 0.0
        with nvtx_range("dtf/stage_inputs"):
            self._stage(s, x.reshape(B, -1), y)
        t_b = time.perf_counter() if prof is not None else 0.0
        with nvtx_range("dtf/launch_step"):
            if top is not None:
                if self.use_mailbox:
                    self._mbox_admit()
                t_c = time.perf_counter() if prof is not None else 0.0
                s.res_expected += 1
                self._launch(s)
            else:
                t_c = t_b
                with torch.cuda.stream(self.stream):
                    self._enqueue_step(s)
        t_d = time.perf_counter() if prof is not None else 0.0
        sink = top is not None and self._uses_sink(s)
        with nvtx_range("dtf/wait_step"):
            if sink:
                self._wait_result(s)
            else:
                self.stream.synchronize()
        if prof is not None:
            t_e = time.perf_counter()
            prof["stage"] += t_b - t_a
            prof["admit"] += t_c - t_b
            prof["launch"] += t_d - t_c
            prof["wait"] += t_e - t_d
            prof["n"] += 1
        if sink:
            loss_val = np.float32(s.res_f32[0])
            step_val = int(s.res_u64[1])
        else:
            loss_val = np.float32(s.out_np[0])
            step_val = int(s.step_np[0]) if top is not None else None
        s.last_loss = loss_val

        out: List[Any] = []
        extra: Optional[Dict[int, Any]] = None
        for f in fetches:
            if isinstance(f, TrainOp):
                out.append(None)
            elif f is p.loss_node:
                out.append(loss_val)
            elif f is gs_var:
                out.append(self._as_step(step_val if step_val is not None else self.store.read_global_step()))
            elif f is p.logits_node:
                out.append(s.logits.cpu().numpy())
            else:
                if extra is None:
                    extra = self._eval_rest([g for g in plain if g not in (p.loss_node, gs_var, p.logits_node)],
                                            feeds, s)
                out.append(extra[id(f)])
        return out

    def _wait_result(self, s: StepState) -> None:
        """Spin (natively, GIL released) on the sequence word the step's last kernel writes."""
        exp = s.res_expected & 0xffffffff
        wait = L.lib().dtf_host_wait_u32
        ptr = s.res_pin.data_ptr() + 4          # word 1 = sequence number
        waited = 0
        while wait(ptr, exp, 200000) != 0:
            if self.stream.query():                # drained without reporting: surface the error
                self.stream.synchronize()
                if int(s.res_u32[1]) != exp:
                    raise RuntimeError("training step finished without reporting its results")
            waited += 1
            if self.sync_generic and waited * 0.2 >= self.sync_token_timeout and waited % 5 == 0:
                self._sync_token_expired(waited * 0.2)
            if waited % 10 == 0 and self.health_every > 0:
                # two seconds without a result: a step may be parked in its device-side flow control
                # behind an owner that died; ask the control plane (raises AbortedError if so)
                self.check_health()

    def _sync_token_expired(self, waited_s: float) -> None:
        """A synchronous step has been waiting for its token for a long time. If every other replica
        has finished training the round can never complete: end this replica's training the way TF
        does when the token queue is closed (OutOfRangeError = clean stop for MonitoredSession).
        Otherwise keep waiting up to ``mailbox_timeout`` and then give the session a chance to recover."""
        from ..parallel.ps_core import AbortedError, OutOfRangeError
        st, lib = self.store, L.lib()
        done = np.zeros(L.MAX_WORKERS, np.uint32)
        L.check(lib.dtf_memcpy_d2h(done.ctypes.data_as(C.c_void_p),
                                   st.mbox_flags_ptr(st.arena.step_shard) + L.MBOX_OFF_DONE, done.nbytes, None), "d2h")
        L.check(lib.dtf_stream_sync(None), "sync")
        nw = int(st.arena.sync_workers)
        others_done = all(done[w] != 0 for w in range(nw) if w != self.worker_id)
        if not others_done and waited_s < self.mailbox_timeout:
            return
        one = (C.c_uint * 1)(1)     # release my parked kernels, then let the stream drain
        L.check(lib.dtf_memcpy_h2d(st.mirror_base + L.MIRROR_OFF_ABORT, one, 4, None), "h2d")
        L.check(lib.dtf_stream_sync(None), "sync")
        self.stream.synchronize()
        if others_done:
            raise OutOfRangeError("synchronization token queue closed: every other replica has finished")
        raise AbortedError("no synchronization token for %.0f s (worker %d)" % (waited_s, self.worker_id))

    def _as_step(self, step: int):
        gs = G.get_global_step(self.graph)
        return np.int64(step) if gs is None or gs.dtype == "int64" else np.int32(step)

    def _find_loss(self, fetches: Sequence[G.Tensor]) -> G.Tensor:
        if self.plan is not None:
            return self.plan.loss_node
        for f in fetches:
            try:
                parse_loss(f)
                return f
            except UnsupportedGraphError:
                continue
        for n in self.graph.nodes:
            if type(n).__name__ == "TrainOp":
                return n.loss
        raise UnsupportedGraphError("cannot find a compiled loss among the fetches")

    def _eval_rest(self, nodes: Sequence[G.Tensor], feeds, s: StepState) -> Dict[int, Any]:
        """Fetches that hang off the logits (accuracy, softmax ...) are finished on the host."""
        from .executor_cpu import evaluate, _to_numpy
        f2 = dict(feeds)
        f2[self.plan.logits_node] = s.logits.cpu().numpy()
        f2[self.plan.loss_node] = np.float32(getattr(s, "last_loss", s.out_np[0]))
        vals = evaluate(list(nodes), f2, {})
        return {id(n): _to_numpy(v) for n, v in zip(nodes, vals)}

    def check_health(self) -> None:
        """Failure detection for the device data plane: a mapped arena stays readable after its
        owner died, so liveness is checked over the control plane. A PS that is gone, or that came
        back as a new incarnation (its arena and state were lost), surfaces as AbortedError — the
        recoverable session then rebuilds the session and the chief restores / re-initialises
        (reference: run_mnist_distributed.py:146)."""
        from ..parallel.ps_core import AbortedError
        for s in range(self.client.num_shards):
            info = self.client.call(s, "ping")       # raises AbortedError when unreachable
            inc = int(info.get("incarnation", 0))
            if self._incarnations.setdefault(s, inc) != inc:
                raise AbortedError(f"/job:ps/task:{s} restarted (state lost)")

    def traffic_per_step(self) -> Dict[str, Any]:
        """Bytes one training step moves, from the layout (no measurement): what is pulled from /
        pushed to REMOTE shards (crosses NVLink) and in total. Used by DataPlaneStatsHook."""
        st = self.store
        pull_remote = pull_total = 0
        for src, _dst, nbytes in ([] if self.use_mirror else getattr(self, "copies", [])):
            pull_total += nbytes
            sh = next((i for i in range(st.num_shards)
                       if st.base[i] <= src < st.base[i] + st.arena.shard_bytes[i]), None)
            if sh is not None and not st.is_local[sh]:
                pull_remote += nbytes
        push_remote = push_total = 0
        per_elem = 4 if self.use_mailbox else {0: 8, 1: 16, 2: 16, 3: 24}.get(getattr(self.spec, "code", 3), 24)
        if self.train_op is not None:
            for var in self.train_op.var_list:
                for r in self._regions(var):
                    b = r.sl.n * per_elem + (0 if self.use_mailbox or r.shadow_off is None else r.sl.n * 2)
                    push_total += b
                    if not st.is_local[r.shard]:
                        push_remote += b
        return dict(pull_remote_bytes=pull_remote, pull_total_bytes=pull_total,
                    push_remote_bytes=push_remote, push_total_bytes=push_total,
                    push_mode="mailbox" if self.use_mailbox else "rmw")

    def worker_heartbeats(self) -> List[int]:
        """Last global step each worker published in the step-owning shard's control block."""
        return self.store.read_ctrl(self.store.arena.step_shard)["heartbeat"]

    # -- benchmarking hooks ---------------------------------------------------------------------------
    def train_state(self) -> StepState:
        for (B, train, *_), s in self.states.items():
            if train:
                return s
        raise RuntimeError("no training step has run yet")

    def replay(self, n: int) -> None:
        """n back-to-back device steps on the current staged batch (no host round trip)."""
        s = self.train_state()
        with torch.cuda.stream(self.stream):
            for _ in range(n):
                s.res_expected += 1
                if self.use_mailbox:
                    self._mbox_admit()
                if s.graph_exec:
                    L.check(self._graph_launch(s.graph_exec, self._stream_handle), "dtf_graph_launch")
                elif s.graph is not None:
                    s.graph.replay()
                else:
                    self._enqueue_step(s)

    def mark_done(self) -> None:
        """Tell every owner that this worker has finished training (``done[worker]`` in its mailbox
        flags). In synchronous mode a replica still waiting for a token learns from these words that
        the round can never complete and stops cleanly (``_sync_token_expired``)."""
        if self.use_mailbox and self.store is not None:
            one = np.ones(1, np.uint32)
            for sh in range(self.store.num_shards):
                self.store._h2d(self.store.mbox_flags_ptr(sh) + L.MBOX_OFF_DONE + 4 * self.worker_id, one)

    def close(self) -> None:
        if self.use_mirror and self.store is not None and self.store.mirror_base:
            # release kernels of mine that are parked in device-side flow control (dead owner)
            try:
                lib = L.lib()
                one = (C.c_uint * 1)(1)
                L.check(lib.dtf_memcpy_h2d(self.store.mirror_base + L.MIRROR_OFF_ABORT, one, 4, None), "h2d")
                L.check(lib.dtf_stream_sync(None), "sync")
            except Exception:
                pass
        try:
            self.stream.synchronize()
        except Exception:
            pass
        if self.use_mailbox and self.store is not None:
            try:
                self.flush_pushes()
                self.mark_done()
            except Exception:
                pass
        if self.use_mailbox and self.store is not None:
            for sh in self.my_shards:
                try:
                    self.client.call(sh, "mbox_attach", self.worker_id, False, 0)
                except Exception:
                    pass
            self.my_shards = []
        self.states.clear()
        if self.store is not None:
            self.store.close()


class _PtrTensor:
    """Minimal stand-in so raw arena pointers can be passed where kernels.py expects a tensor."""

    def __init__(self, ptr: int):
        self._ptr = ptr

    def data_ptr(self) -> int:
        return self._ptr


def K_conv1_fwd(x, w1_ptr, b1_ptr, pool1p, amax1, scale, x_copy=None, ingest=False, **kw):
    K.conv1_fwd(x, _PtrTensor(w1_ptr), _PtrTensor(b1_ptr), pool1p, amax1, scale, x_copy=x_copy, ingest=ingest, **kw)


def K_conv2_fwd(pool1p, w2_ptr, b2_ptr, pool2, amax2):
    K.conv2_fwd(pool1p, w2_ptr, _PtrTensor(b2_ptr), pool2, amax2)
