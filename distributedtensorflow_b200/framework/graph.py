"""A small deferred-execution graph: the TF1-style surface the reference's scripts are written in.

The reference builds a TF1 graph (run_mnist_distributed.py:107-116,
templates/00_mnist_replica.py:133-180) and then drives it with ``session.run``. This module is the
equivalent: symbolic ``Tensor`` nodes, ``Variable`` nodes that are *placed on PS shards by the
active device function*, and a handful of ops. It is deliberately tiny — enough to express the
reference's CNN, its MLP and their losses (plus arbitrary compositions of the same ops) — because
execution is done by backend executors:

* ``executor_cpu``  evaluates any graph with PyTorch ops + autograd (the oracle and the CPU path),
* ``executor_b200`` pattern-compiles sequential conv/dense graphs into the hand-written sm_100a
  kernels and a CUDA graph.

Every process builds the same graph independently ("between-graph replication"); they share state
only because variable names and PS placements coincide.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np

from .. import placement as _placement

float32 = "float32"
int32 = "int32"
int64 = "int64"
uint8 = "uint8"
_DTYPE_BYTES = {"float32": 4, "int32": 4, "int64": 8, "uint8": 1, "bfloat16": 2}


class GraphKeys:
    GLOBAL_VARIABLES = "variables"
    TRAINABLE_VARIABLES = "trainable_variables"
    LOCAL_VARIABLES = "local_variables"
    GLOBAL_STEP = "global_step"
    SUMMARIES = "summaries"
    INIT_OP = "init_op"
    READY_OP = "ready_op"
    SAVERS = "savers"


class Graph:
    def __init__(self):
        self.nodes: List["Tensor"] = []
        self.collections: Dict[str, list] = {}
        self._names: Dict[str, int] = {}
        self.seed: Optional[int] = None
        self._finalized = False

    def unique_name(self, base: str) -> str:
        """TF naming: first use is ``base``, later uses ``base_1``, ``base_2`` ..."""
        n = self._names.get(base, 0)
        self._names[base] = n + 1
        return base if n == 0 else f"{base}_{n}"

    def add_to_collection(self, key: str, value) -> None:
        self.collections.setdefault(key, []).append(value)

    def get_collection(self, key: str) -> list:
        return list(self.collections.get(key, []))

    def finalize(self) -> None:
        self._finalized = True

    @property
    def finalized(self) -> bool:
        return self._finalized

    def variables(self) -> List["Variable"]:
        return self.get_collection(GraphKeys.GLOBAL_VARIABLES)

    def trainable_variables(self) -> List["Variable"]:
        return self.get_collection(GraphKeys.TRAINABLE_VARIABLES)


class _Default:
    # process-wide, like TF's global default graph: helper threads (queue runners, in-process
    # worker threads of the fake cluster) see the graph the main thread built
    graph: Optional[Graph] = None


def get_default_graph() -> Graph:
    if _Default.graph is None:
        _Default.graph = Graph()
    return _Default.graph


def reset_default_graph() -> Graph:
    _Default.graph = Graph()
    return _Default.graph


def set_random_seed(seed: int) -> None:
    get_default_graph().seed = seed


class Tensor:
    """A symbolic value: the output of ``op`` applied to ``inputs``."""

    def __init__(self, op: str, inputs: Sequence["Tensor"] = (), attrs: Optional[dict] = None,
                 shape: Optional[Sequence[Optional[int]]] = None, dtype: str = float32,
                 name: Optional[str] = None, graph: Optional[Graph] = None):
        self.graph = graph or get_default_graph()
        if self.graph.finalized:
            raise RuntimeError("Graph is finalized and cannot be modified.")
        self.op = op
        self.inputs = list(inputs)
        self.attrs = dict(attrs or {})
        self.shape = tuple(shape) if shape is not None else None
        self.dtype = dtype
        self.name = self.graph.unique_name(name or op)
        info = _placement.OpInfo(name=self.name, type=op)
        self.device = _placement.apply_device_functions(info) if op not in ("Variable", "VariableV2") else ""
        self.graph.nodes.append(self)

    def __repr__(self) -> str:
        return f"<Tensor {self.name!r} op={self.op} shape={self.shape} dtype={self.dtype}>"

    # operator sugar used by the MLP template's loss expression
    def __neg__(self):
        return neg(self)

    def __mul__(self, other):
        return multiply(self, other)

    __rmul__ = __mul__

    def __add__(self, other):
        return add(self, other)

    __radd__ = __add__

    def __sub__(self, other):
        return add(self, neg(_as_tensor(other)))

    def get_shape(self):
        return self.shape


# ------------------------------------------------------------------------------------------------
# initializers
# ------------------------------------------------------------------------------------------------
@dataclass
class Initializer:
    kind: str                       # zeros | ones | constant | truncated_normal | glorot_uniform | random_normal | value
    shape: Tuple[int, ...] = ()
    dtype: str = float32
    mean: float = 0.0
    stddev: float = 1.0
    value: Any = None
    seed: Optional[int] = None

    def numpy(self, name: str, graph_seed: Optional[int]) -> np.ndarray:
        """Deterministic host-side materialisation (the chief runs this for init_op)."""
        import zlib
        np_dtype = {"float32": np.float32, "int32": np.int32, "int64": np.int64}[self.dtype]
        base = self.seed if self.seed is not None else (graph_seed if graph_seed is not None else 87654321)
        rng = np.random.default_rng([base, zlib.crc32(name.encode())])
        shp = self.shape
        if self.kind == "zeros":
            return np.zeros(shp, np_dtype)
        if self.kind == "ones":
            return np.ones(shp, np_dtype)
        if self.kind == "constant":
            return np.full(shp, self.value, np_dtype)
        if self.kind == "value":
            return np.asarray(self.value, np_dtype).reshape(shp)
        if self.kind == "random_normal":
            return (rng.standard_normal(shp) * self.stddev + self.mean).astype(np_dtype)
        if self.kind == "truncated_normal":
            # values more than two standard deviations from the mean are re-drawn
            out = rng.standard_normal(shp)
            bad = np.abs(out) > 2.0
            while bad.any():
                out[bad] = rng.standard_normal(int(bad.sum()))
                bad = np.abs(out) > 2.0
            return (out * self.stddev + self.mean).astype(np_dtype)
        if self.kind == "glorot_uniform":
            if len(shp) < 1:
                fan_in = fan_out = 1
            elif len(shp) == 1:
                fan_in = fan_out = shp[0]
            elif len(shp) == 2:
                fan_in, fan_out = shp
            else:
                rf = int(math.prod(shp[:-2]))
                fan_in, fan_out = shp[-2] * rf, shp[-1] * rf
            limit = math.sqrt(6.0 / (fan_in + fan_out))
            return rng.uniform(-limit, limit, shp).astype(np_dtype)
        raise ValueError(f"unknown initializer {self.kind}")


def zeros(shape, dtype=float32) -> Initializer:
    return Initializer("zeros", tuple(shape), dtype)


def ones(shape, dtype=float32) -> Initializer:
    return Initializer("ones", tuple(shape), dtype)


def truncated_normal(shape, mean=0.0, stddev=1.0, dtype=float32, seed=None) -> Initializer:
    return Initializer("truncated_normal", tuple(shape), dtype, mean, stddev, seed=seed)


def random_normal(shape, mean=0.0, stddev=1.0, dtype=float32, seed=None) -> Initializer:
    return Initializer("random_normal", tuple(shape), dtype, mean, stddev, seed=seed)


def glorot_uniform_initializer(seed=None):
    def make(shape, dtype=float32):
        return Initializer("glorot_uniform", tuple(shape), dtype, seed=seed)
    return make


def zeros_initializer():
    def make(shape, dtype=float32):
        return Initializer("zeros", tuple(shape), dtype)
    return make


def constant_initializer(value):
    def make(shape, dtype=float32):
        return Initializer("constant", tuple(shape), dtype, value=value)
    return make


# ------------------------------------------------------------------------------------------------
# variables
# ------------------------------------------------------------------------------------------------
class Variable(Tensor):
    """A PS-resident tensor. Creation consults the device-function stack exactly like TF's
    ``VariableV2`` op, so ``replica_device_setter`` decides which PS task owns it."""

    def __init__(self, initial_value, name: Optional[str] = None, trainable: bool = True,
                 dtype: Optional[str] = None, collections: Optional[Sequence[str]] = None,
                 colocate_with: Optional["Variable"] = None):
        if isinstance(initial_value, Initializer):
            init = initial_value
        else:
            arr = np.asarray(initial_value)
            if dtype is None:
                # tf.Variable(0) is int32 (templates/00_mnist_replica.py:138), python floats fp32
                if isinstance(initial_value, (int, float)) or arr.dtype.kind == "f":
                    dtype = float32 if arr.dtype.kind == "f" else int32
                else:
                    dtype = int64 if arr.dtype == np.int64 else int32
            init = Initializer("value", tuple(arr.shape), dtype, value=arr)
        if dtype is not None and dtype != init.dtype:
            init = Initializer(init.kind, init.shape, dtype, init.mean, init.stddev, init.value, init.seed)
        super().__init__("VariableV2", (), {}, init.shape, init.dtype, name or "Variable")
        self.initializer = init
        self.trainable = trainable
        info = _placement.OpInfo(name=self.name, type="VariableV2", shape=init.shape,
                                 dtype_bytes=_DTYPE_BYTES[init.dtype])
        if colocate_with is not None:
            self.device = colocate_with.device
        else:
            self.device = _placement.apply_device_functions(info)
        spec = _placement.DeviceSpec.from_string(self.device)
        self.ps_task: Optional[int] = spec.task if spec.job == "ps" else None
        cols = list(collections) if collections is not None else [GraphKeys.GLOBAL_VARIABLES]
        if trainable and GraphKeys.TRAINABLE_VARIABLES not in cols and collections is None:
            cols.append(GraphKeys.TRAINABLE_VARIABLES)
        for c in cols:
            self.graph.add_to_collection(c, self)

    @property
    def num_elements(self) -> int:
        return int(math.prod(self.shape)) if self.shape else 1

    @property
    def is_sliced(self) -> bool:
        return self.ps_task == _placement.SLICED

    def initial_numpy(self) -> np.ndarray:
        return self.initializer.numpy(self.name, self.graph.seed)


def get_variable(name: str, shape, dtype=float32, initializer=None, trainable=True) -> Variable:
    make = initializer or glorot_uniform_initializer()
    return Variable(make(tuple(shape), dtype), name=name, trainable=trainable)


def global_variables() -> List[Variable]:
    return get_default_graph().variables()


def trainable_variables() -> List[Variable]:
    return get_default_graph().trainable_variables()


# ------------------------------------------------------------------------------------------------
# ops
# ------------------------------------------------------------------------------------------------
def _as_tensor(x) -> Tensor:
    if isinstance(x, Tensor):
        return x
    arr = np.asarray(x, dtype=np.float32)
    return Tensor("Const", (), {"value": arr}, arr.shape, float32, "Const")


def constant(value, dtype=float32, name="Const") -> Tensor:
    arr = np.asarray(value)
    return Tensor("Const", (), {"value": arr}, arr.shape, dtype, name)


def placeholder(dtype, shape=None, name="Placeholder") -> Tensor:
    return Tensor("Placeholder", (), {}, shape, dtype, name)


def _resolve_shape(in_shape, new_shape):
    new_shape = list(new_shape)
    if in_shape is None or any(d is None for d in in_shape):
        known = None
    else:
        known = int(math.prod(in_shape))
    if -1 in new_shape:
        i = new_shape.index(-1)
        rest = int(math.prod([d for d in new_shape if d != -1]))
        new_shape[i] = None if known is None else known // rest
    return tuple(new_shape)


def reshape(tensor: Tensor, shape, name="Reshape") -> Tensor:
    return Tensor("Reshape", [tensor], {"shape": tuple(shape)}, _resolve_shape(tensor.shape, shape),
                  tensor.dtype, name)


def relu(x: Tensor, name="Relu") -> Tensor:
    return Tensor("Relu", [x], {}, x.shape, x.dtype, name)


def softmax(x: Tensor, name="Softmax") -> Tensor:
    return Tensor("Softmax", [x], {}, x.shape, x.dtype, name)


def log(x: Tensor, name="Log") -> Tensor:
    return Tensor("Log", [x], {}, x.shape, x.dtype, name)


def clip_by_value(x: Tensor, clip_value_min: float, clip_value_max: float, name="clip_by_value") -> Tensor:
    return Tensor("ClipByValue", [x], {"min": float(clip_value_min), "max": float(clip_value_max)},
                  x.shape, x.dtype, name)


def multiply(a, b, name="mul") -> Tensor:
    a, b = _as_tensor(a), _as_tensor(b)
    return Tensor("Mul", [a, b], {}, a.shape if a.shape else b.shape, a.dtype, name)


def add(a, b, name="add") -> Tensor:
    a, b = _as_tensor(a), _as_tensor(b)
    return Tensor("Add", [a, b], {}, a.shape if a.shape else b.shape, a.dtype, name)


def neg(x: Tensor, name="Neg") -> Tensor:
    return Tensor("Neg", [x], {}, x.shape, x.dtype, name)


def reduce_sum(x: Tensor, axis=None, name="Sum") -> Tensor:
    return Tensor("Sum", [x], {"axis": axis}, () if axis is None else None, x.dtype, name)


def reduce_mean(x: Tensor, axis=None, name="Mean") -> Tensor:
    return Tensor("Mean", [x], {"axis": axis}, () if axis is None else None, x.dtype, name)


def matmul(a: Tensor, b: Tensor, name="MatMul") -> Tensor:
    shp = (a.shape[0] if a.shape else None, b.shape[-1] if b.shape else None)
    return Tensor("MatMul", [a, b], {}, shp, a.dtype, name)


def bias_add(x: Tensor, b: Tensor, name="BiasAdd") -> Tensor:
    return Tensor("BiasAdd", [x, b], {}, x.shape, x.dtype, name)


def xw_plus_b(x: Tensor, weights: Tensor, biases: Tensor, name="xw_plus_b") -> Tensor:
    """tf.nn.xw_plus_b (templates/00_mnist_replica.py:160,163)."""
    return bias_add(matmul(x, weights, name=name + "/MatMul"), biases, name=name)


def conv2d_op(x: Tensor, kernel: Tensor, padding="SAME", strides=(1, 1), name="Conv2D") -> Tensor:
    n, h, w, _ = x.shape
    kh, kw, _, co = kernel.shape
    if padding.upper() == "SAME":
        oh, ow = (None if h is None else -(-h // strides[0])), (None if w is None else -(-w // strides[1]))
    else:
        oh = None if h is None else (h - kh) // strides[0] + 1
        ow = None if w is None else (w - kw) // strides[1] + 1
    return Tensor("Conv2D", [x, kernel], {"padding": padding.upper(), "strides": tuple(strides)},
                  (n, oh, ow, co), x.dtype, name)


def max_pool(x: Tensor, pool_size=(2, 2), strides=(2, 2), padding="VALID", name="MaxPool") -> Tensor:
    n, h, w, c = x.shape
    oh = None if h is None else (h - pool_size[0]) // strides[0] + 1
    ow = None if w is None else (w - pool_size[1]) // strides[1] + 1
    return Tensor("MaxPool", [x], {"pool_size": tuple(pool_size), "strides": tuple(strides),
                                   "padding": padding.upper()}, (n, oh, ow, c), x.dtype, name)


def cast(x: Tensor, dtype, name="Cast") -> Tensor:
    return Tensor("Cast", [x], {"dtype": dtype}, x.shape, dtype, name)


def argmax(x: Tensor, axis=-1, name="ArgMax") -> Tensor:
    return Tensor("ArgMax", [x], {"axis": axis}, x.shape[:-1] if x.shape else None, int64, name)


def equal(a: Tensor, b: Tensor, name="Equal") -> Tensor:
    return Tensor("Equal", [a, b], {}, a.shape, "bool", name)


def sparse_softmax_cross_entropy_with_logits(labels: Tensor, logits: Tensor,
                                             name="SparseSoftmaxCrossEntropyWithLogits") -> Tensor:
    return Tensor("SparseSoftmaxXent", [logits, labels], {}, (logits.shape[0],) if logits.shape else None,
                  float32, name)


def sparse_softmax_cross_entropy(labels: Tensor, logits: Tensor, scope=None) -> Tensor:
    """tf.losses.sparse_softmax_cross_entropy: mean over the batch
    (run_mnist_distributed.py:113)."""
    per_ex = sparse_softmax_cross_entropy_with_logits(labels, logits,
                                                      name="sparse_softmax_cross_entropy_loss/xentropy")
    loss = reduce_mean(per_ex, name="sparse_softmax_cross_entropy_loss/value")
    loss.graph.add_to_collection("losses", loss)
    return loss


# ------------------------------------------------------------------------------------------------
# tf.layers equivalents (variable names follow TF's defaults so checkpoints line up)
# ------------------------------------------------------------------------------------------------
def conv2d(inputs: Tensor, filters: int, kernel_size, padding="valid", activation=None,
           use_bias=True, kernel_initializer=None, bias_initializer=None, name=None) -> Tensor:
    """tf.layers.conv2d (run_mnist_distributed.py:52-57,59-64): NHWC, Glorot-uniform kernel,
    zero bias. Variables are created kernel first, then bias (the round-robin order)."""
    g = inputs.graph
    scope = g.unique_name(name or "conv2d")
    kh, kw = (kernel_size, kernel_size) if isinstance(kernel_size, int) else tuple(kernel_size)
    cin = inputs.shape[-1]
    kinit = kernel_initializer or glorot_uniform_initializer()
    kernel = Variable(kinit((kh, kw, cin, filters), float32), name=f"{scope}/kernel")
    out = conv2d_op(inputs, kernel, padding=padding, name=f"{scope}/Conv2D")
    if use_bias:
        binit = bias_initializer or zeros_initializer()
        bias = Variable(binit((filters,), float32), name=f"{scope}/bias")
        out = bias_add(out, bias, name=f"{scope}/BiasAdd")
    if activation is not None:
        out = activation(out, name=f"{scope}/Relu") if activation is relu else activation(out)
    return out


def max_pooling2d(inputs: Tensor, pool_size, strides, padding="valid", name=None) -> Tensor:
    """tf.layers.max_pooling2d (run_mnist_distributed.py:58,65)."""
    ps = (pool_size, pool_size) if isinstance(pool_size, int) else tuple(pool_size)
    st = (strides, strides) if isinstance(strides, int) else tuple(strides)
    scope = inputs.graph.unique_name(name or "max_pooling2d")
    return max_pool(inputs, ps, st, padding, name=f"{scope}/MaxPool")


def dense(inputs: Tensor, units: int, activation=None, use_bias=True, kernel_initializer=None,
          bias_initializer=None, name=None) -> Tensor:
    """tf.layers.dense (run_mnist_distributed.py:67,69)."""
    g = inputs.graph
    scope = g.unique_name(name or "dense")
    kinit = kernel_initializer or glorot_uniform_initializer()
    kernel = Variable(kinit((inputs.shape[-1], units), float32), name=f"{scope}/kernel")
    out = matmul(inputs, kernel, name=f"{scope}/MatMul")
    if use_bias:
        binit = bias_initializer or zeros_initializer()
        bias = Variable(binit((units,), float32), name=f"{scope}/bias")
        out = bias_add(out, bias, name=f"{scope}/BiasAdd")
    if activation is not None:
        out = activation(out, name=f"{scope}/Relu") if activation is relu else activation(out)
    return out


# ------------------------------------------------------------------------------------------------
# global step (tf.train.get_or_create_global_step, run_mnist_distributed.py:114)
# ------------------------------------------------------------------------------------------------
def get_global_step(graph: Optional[Graph] = None) -> Optional[Variable]:
    g = graph or get_default_graph()
    col = g.get_collection(GraphKeys.GLOBAL_STEP)
    if col:
        return col[0]
    for v in g.variables():
        if v.name == "global_step":
            return v
    return None


def create_global_step(graph: Optional[Graph] = None) -> Variable:
    g = graph or get_default_graph()
    if get_global_step(g) is not None:
        raise ValueError('"global_step" already exists.')
    v = Variable(Initializer("zeros", (), int64), name="global_step", trainable=False,
                 collections=[GraphKeys.GLOBAL_VARIABLES, GraphKeys.GLOBAL_STEP])
    return v


def get_or_create_global_step(graph: Optional[Graph] = None) -> Variable:
    g = graph or get_default_graph()
    gs = get_global_step(g)
    return gs if gs is not None else create_global_step(g)
