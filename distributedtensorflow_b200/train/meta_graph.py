"""``<prefix>.meta``: the MetaGraphDef TF1's ``Saver.save`` writes next to every checkpoint.

The reference's checkpoints come from ``MonitoredTrainingSession(checkpoint_dir=...)``
(templates/00_between_graph_replication_async_mnist.py:43-45), whose Saver exports the meta graph
with each save. There is no TensorFlow graph here to export, but the part of the meta graph that tools
read without running anything is cheap to state: one ``VariableV2`` node per variable with its dtype,
shape and the device ``replica_device_setter`` gave it (``/job:ps/task:k`` - the between-graph
placement, run_mnist_distributed.py:107), the ``variables`` / ``trainable_variables`` collections as
``VariableDef`` records, and the ``SaverDef`` (V2 format, sharded). Field numbers are those of
tensorflow/core/protobuf/meta_graph.proto, saver.proto, framework/{graph,node_def,attr_value,
tensor_shape,variable,versions}.proto; encoded by hand (``utils/protowire``).
"""
from __future__ import annotations

import os
from typing import Dict, List, Sequence

from ..utils import protowire as pw

DT = {"float32": 1, "float64": 2, "int32": 3, "uint8": 4, "int64": 9, "bool": 10, "bfloat16": 14}
GRAPH_DEF_PRODUCER = 27          # TF 1.12's graph version


def _shape_proto(shape: Sequence[int]) -> bytes:
    return b"".join(pw.field_bytes(2, pw.field_varint(1, int(d))) for d in shape)


def _attr(key: str, value: bytes) -> bytes:
    """One entry of NodeDef.attr (map<string, AttrValue>: key = 1, value = 2)."""
    return pw.field_bytes(5, pw.field_string(1, key) + pw.field_bytes(2, value))


def _variable_node(name: str, dtype: str, shape: Sequence[int], device: str) -> bytes:
    body = pw.field_string(1, name) + pw.field_string(2, "VariableV2")
    if device:
        body += pw.field_string(4, device)
    body += _attr("container", pw.field_bytes(2, b""))
    body += _attr("dtype", pw.field_varint(6, DT[dtype]))
    body += _attr("shape", pw.field_bytes(7, _shape_proto(shape)))
    body += _attr("shared_name", pw.field_bytes(2, b""))
    return body


def _variable_def(name: str, trainable: bool) -> bytes:
    body = pw.field_string(1, name + ":0") + pw.field_string(2, name + "/Assign") + pw.field_string(3, name + "/read:0")
    if trainable:
        body += pw.field_varint(7, 1)
    return body


def _collection(key: str, values: List[bytes]) -> bytes:
    """One entry of MetaGraphDef.collection_def with a bytes_list (CollectionDef field 2)."""
    bytes_list = b"".join(pw.field_bytes(1, v) for v in values)
    return pw.field_bytes(4, pw.field_string(1, key) + pw.field_bytes(2, pw.field_bytes(2, bytes_list)))


def encode_meta_graph(variables: Sequence, sharded: bool = True, max_to_keep: int = 5,
                      keep_checkpoint_every_n_hours: float = 10000.0) -> bytes:
    """``variables``: objects with ``name``, ``dtype``, ``shape``, ``device``, ``trainable``."""
    meta_info = pw.field_string(4, "train") + pw.field_string(5, "1.12.0") + \
        pw.field_string(6, "distributedtensorflow_b200")
    graph = b"".join(pw.field_bytes(1, _variable_node(v.name, v.dtype, tuple(v.shape), getattr(v, "device", "") or ""))
                     for v in variables)
    graph += pw.field_bytes(4, pw.field_varint(1, GRAPH_DEF_PRODUCER))
    saver = pw.field_string(1, "save/Const:0") + pw.field_string(2, "save/Identity:0" if sharded else "save/control_dependency:0") + \
        pw.field_string(3, "save/restore_all") + pw.field_varint(4, int(max_to_keep)) + \
        (pw.field_varint(5, 1) if sharded else b"") + pw.field_float(6, float(keep_checkpoint_every_n_hours)) + \
        pw.field_varint(7, 2)                                                  # CheckpointFormatVersion.V2
    out = pw.field_bytes(1, meta_info) + pw.field_bytes(2, graph) + pw.field_bytes(3, saver)
    out += _collection("variables", [_variable_def(v.name, bool(getattr(v, "trainable", False))) for v in variables])
    out += _collection("trainable_variables", [_variable_def(v.name, True) for v in variables
                                               if getattr(v, "trainable", False)])
    return out


def write_meta_graph(prefix: str, variables: Sequence, **saver_kw) -> str:
    path = prefix + ".meta"
    tmp = path + ".tmp%d" % os.getpid()
    with open(tmp, "wb") as f:
        f.write(encode_meta_graph(variables, **saver_kw))
    os.replace(tmp, path)
    return path


def read_meta_graph(path: str) -> Dict[str, object]:
    """Decode what ``encode_meta_graph`` writes (tests, tooling): variables with dtype / shape /
    device, the collections, the saver settings."""
    top = pw.parse(open(path, "rb").read())
    inv = {v: k for k, v in DT.items()}
    nodes = []
    graph = pw.parse(top[2][0])
    for raw in graph.get(1, []):
        nd = pw.parse(raw)
        attrs = {}
        for a in nd.get(5, []):
            kv = pw.parse(a)
            attrs[kv[1][0].decode()] = pw.parse(kv[2][0])
        shape = [int(pw.parse(d)[1][0]) if 1 in pw.parse(d) else 0 for d in pw.parse(attrs["shape"][7][0]).get(2, [])] \
            if attrs["shape"].get(7) else []
        nodes.append(dict(name=nd[1][0].decode(), op=nd[2][0].decode(),
                          device=nd[4][0].decode() if 4 in nd else "", dtype=inv[int(attrs["dtype"][6][0])],
                          shape=tuple(shape)))
    saver = pw.parse(top[3][0])
    collections = {}
    for raw in top.get(4, []):
        kv = pw.parse(raw)
        vals = pw.parse(pw.parse(kv[2][0])[2][0]).get(1, []) if 2 in pw.parse(kv[2][0]) else []
        collections[kv[1][0].decode()] = [pw.parse(v)[1][0].decode() for v in vals]
    return dict(nodes=nodes, producer=int(pw.parse(graph[4][0])[1][0]), collections=collections,
                saver=dict(max_to_keep=int(saver[4][0]), sharded=bool(saver.get(5, [0])[0]),
                           version=int(saver[7][0]), filename_tensor_name=saver[1][0].decode()),
                tags=[t.decode() for t in pw.parse(top[1][0]).get(4, [])])
