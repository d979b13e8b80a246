"""Config loader, ClusterSpec and variable placement."""
import json

import pytest

import distributedtensorflow_b200 as tf
from distributedtensorflow_b200 import placement as P
from distributedtensorflow_b200.cluster import Config, ClusterSpec, cluster_from_flags, parse_address
from distributedtensorflow_b200.models import mnist_cnn, mnist_mlp
from distributedtensorflow_b200.parallel.layout import build_layout, bytes_per_shard, powers_shard


def _write(tmp_path, obj):
    p = tmp_path / "config.json"
    p.write_text(json.dumps(obj))
    return str(p)


def test_config_order_is_task_index(tmp_path):
    # labels are never parsed: position in the JSON object is the task index
    path = _write(tmp_path, {"ps": {"ps:1": "10.0.0.1:8001", "ps:0": "10.0.0.2:8002"},
                             "workers": {"zzz": "10.0.0.3:6001", "aaa": "10.0.0.4:6002"}})
    cfg = Config(path)
    ps_hosts, worker_hosts = cfg.get_ps_and_worker_hosts()
    assert ps_hosts == ("10.0.0.1:8001", "10.0.0.2:8002")
    assert worker_hosts == ("10.0.0.3:6001", "10.0.0.4:6002")
    assert cfg.get_workers_with_addresses()[0] == ("zzz", "aaa")
    spec = cfg.cluster_spec()
    assert spec.task_address("worker", 1) == "10.0.0.4:6002"  # json key "workers", job "worker"


@pytest.mark.parametrize("missing", ["ps", "workers"])
def test_config_missing_key_error_text(tmp_path, missing):
    obj = {"ps": {"ps:0": "h:1"}, "workers": {"worker:0": "h:2"}}
    del obj[missing]
    with pytest.raises(AttributeError) as e:
        Config(_write(tmp_path, obj))
    assert str(e.value) == ('Please provide a value for "%s" configuration key in the config.json file!' % missing)


def test_reference_sample_config_loads():
    cfg = Config("config.json")
    assert len(cfg.get_ps_and_worker_hosts()[0]) >= 1


def test_cluster_spec_api():
    c = ClusterSpec({"ps": ["a:1", "b:2"], "worker": {0: "c:3", 2: "d:4"}})
    assert sorted(c.jobs) == ["ps", "worker"]
    assert c.num_tasks("ps") == 2 and c.task_indices("worker") == [0, 2]
    assert c.job_tasks("worker") == ["c:3", None, "d:4"]
    assert c.as_dict() == {"ps": ["a:1", "b:2"], "worker": {0: "c:3", 2: "d:4"}}
    assert ClusterSpec(c) == c and ClusterSpec({"ps": ["a:1"]}) != c
    with pytest.raises(ValueError):
        c.task_address("worker", 1)
    with pytest.raises(ValueError):
        c.num_tasks("chief")
    assert len(c.as_cluster_def()) > 0
    assert cluster_from_flags("a:1,b:2", "c:3").num_tasks("ps") == 2


def test_address_grammar():
    a = parse_address("127.0.0.1:2222/gpu:3")
    assert (a.host, a.port, a.device, a.endpoint) == ("127.0.0.1", 2222, "cuda:3", "127.0.0.1:2222")
    assert parse_address("host:1/cuda:0").device == "cuda:0"
    assert parse_address("host:1/cpu").device == "cpu"
    assert parse_address("host:1").device is None
    with pytest.raises(ValueError):
        parse_address("nohostport")


def _build_cnn(num_ps, strategy=None):
    tf.reset_default_graph()
    cluster = ClusterSpec({"ps": ["h:%d" % i for i in range(num_ps)], "worker": ["w:1"]})
    with tf.device(tf.train.replica_device_setter(worker_device="/job:worker/task:0", cluster=cluster,
                                                  ps_strategy=strategy)):
        x = tf.placeholder(tf.float32, [None, 784])
        y = tf.placeholder(tf.int32, [None])
        mnist_cnn.build_train_graph(x, y)
    return tf.global_variables()


# PS task of each CNN variable at 1 / 2 / 4 / 8 PS tasks
EXPECTED = {
    "conv2d/kernel": (0, 0, 0, 0), "conv2d/bias": (0, 1, 1, 1), "conv2d_1/kernel": (0, 0, 2, 2),
    "conv2d_1/bias": (0, 1, 3, 3), "dense/kernel": (0, 0, 0, 4), "dense/bias": (0, 1, 1, 5),
    "dense_1/kernel": (0, 0, 2, 6), "dense_1/bias": (0, 1, 3, 7), "global_step": (0, 0, 0, 0),
}


@pytest.mark.parametrize("col,num_ps", [(0, 1), (1, 2), (2, 4), (3, 8)])
def test_round_robin_matches_tf_table(col, num_ps):
    vs = _build_cnn(num_ps)
    assert [v.name for v in vs] == list(EXPECTED)
    for v in vs:
        assert v.ps_task == EXPECTED[v.name][col], v.name
        assert v.device == "/job:ps/task:%d" % EXPECTED[v.name][col]


def test_two_ps_imbalance_is_reproduced():
    vs = _build_cnn(2)
    loads = bytes_per_shard(build_layout(vs, 2), 2)
    assert loads == [13094016 + 8, 4520]  # all kernels (+ int64 global_step) on ps:0, biases on ps:1
    assert sum(v.num_elements for v in tf.trainable_variables()) == mnist_cnn.NUM_PARAMS == 3274634


def test_greedy_and_flat_sliced_balance():
    vs = _build_cnn(4, P.GreedyLoadBalancingStrategy(4))
    loads = bytes_per_shard(build_layout(vs, 4), 4)
    # the 12.8 MB tensor goes to the then least-loaded task and nothing big joins it afterwards
    assert sorted(loads)[-1] <= 12845056 + 4096 and sorted(loads)[-2] <= 204800 + 45000
    vs = _build_cnn(8, P.FlatSlicedStrategy(8))
    layout = build_layout(vs, 8)
    dk = layout["dense/kernel"]
    assert [s.row_end - s.row_begin for s in dk.slices] == [448] * 7 + [0]
    assert all(s.row_begin % 64 == 0 for s in dk.slices)
    assert sum(s.n for s in dk.slices) == 3136 * 1024
    loads = bytes_per_shard(layout, 8)
    assert max(loads) < 0.2 * sum(loads)  # no shard holds 98 % of the bytes any more
    # slots / beta powers follow their variable: the shard of the first variable by name
    names = [v.name for v in tf.trainable_variables()]
    assert powers_shard(layout, names) == layout["conv2d/bias"].home_shard


def test_mlp_template_placement_and_dtypes():
    tf.reset_default_graph()
    cluster = ClusterSpec({"ps": ["h:1", "h:2"], "worker": ["w:1", "w:2"]})
    with tf.device(tf.train.replica_device_setter(worker_device="/job:worker/task:1/cpu:0",
                                                  ps_device="/job:ps/cpu:0", cluster=cluster)):
        x = tf.placeholder(tf.float32, [None, 784])
        y_ = tf.placeholder(tf.float32, [None, 10])
        gs, xent, y = mnist_mlp.build_model(x, y_)
    vs = {v.name: v for v in tf.global_variables()}
    assert list(vs) == ["global_step", "hid_w", "hid_b", "sm_w", "sm_b"]
    assert vs["global_step"].dtype == "int32" and vs["global_step"].ps_task == 0  # created first
    assert [vs[n].ps_task for n in ("hid_w", "hid_b", "sm_w", "sm_b")] == [1, 0, 1, 0]
    assert vs["hid_w"].device == "/job:ps/task:1/device:CPU:0"
    assert x.device == "/job:worker/task:1/device:CPU:0"
    assert sum(v.num_elements for v in tf.trainable_variables()) == 79510


def test_device_setter_edge_cases():
    assert tf.train.replica_device_setter(ps_tasks=0) is None
    assert tf.train.replica_device_setter(cluster={"worker": ["w:1"]}) is None
    with pytest.raises(TypeError):
        tf.train.replica_device_setter(ps_tasks=2, ps_strategy=3)
    fn = tf.train.replica_device_setter(ps_tasks=3)
    ops = [P.OpInfo("v%d" % i) for i in range(4)]
    assert [fn(o) for o in ops] == ["/job:ps/task:0", "/job:ps/task:1", "/job:ps/task:2", "/job:ps/task:0"]
    assert fn(P.OpInfo("m", type="MatMul")) == "/job:worker"
    # merge_devices: an outer, more specific device wins over the setter's fields
    assert fn(P.OpInfo("v", device="/job:ps/task:2")) == "/job:ps/task:2"


def test_arena_layout_with_async_mailboxes_is_consistent():
    """Device arena of the B200 data plane (parallel/nvl.py): var / slots / serving copy regions,
    the per-shard mailbox flag block and `depth` mailbox buffers per worker must never overlap,
    every region must keep the alignment its kernels assume, and every task must derive the same
    layout (between-graph replication: nobody ships offsets around)."""
    from distributedtensorflow_b200.ops import _lib as L
    from distributedtensorflow_b200.parallel.nvl import build_arena_layout
    from distributedtensorflow_b200.train.optimizers import OptimizerSpec
    vs = _build_cnn(8, P.FlatSlicedStrategy(8))
    layout = build_layout(vs, 8)
    names = [v.name for v in tf.trainable_variables()]
    spec = OptimizerSpec("adam", 5e-4)
    W, depth = 8, 4
    al = build_arena_layout(layout, 8, spec, names, sync_workers=W, mailbox_depth=depth, mirrors=False)
    again = build_arena_layout(layout, 8, spec, names, sync_workers=W, mailbox_depth=depth, mirrors=False)
    assert al.shard_bytes == again.shard_bytes and al.mailbox_off == again.mailbox_off
    for sh in range(8):
        spans = [(0, L.CTRL_BYTES)]
        for name, regs in al.regions.items():
            for r in regs:
                if r.shard != sh:
                    continue
                assert r.var_off % 256 == 0
                spans.append((r.var_off, r.var_off + r.var_bytes))
                for so in r.slot_off:
                    spans.append((so, so + r.var_bytes))
                if r.shadow_off is not None:
                    assert r.shadow_off % 256 == 0 and r.shadow_ld % 8 == 0
                    spans.append((r.shadow_off, r.shadow_off + r.shadow_bytes))
        f = al.mbox_flags_off[sh]
        assert f % 256 == 0 and al.mbox_slot_base[sh] == f + L.MBOX_FLAGS_BYTES
        spans.append((f, f + L.MBOX_FLAGS_BYTES))
        stride = al.mailbox_stride[sh]
        assert stride % 256 == 0
        slices_here = [(n_, r) for n_ in names for r in al.regions[n_] if r.shard == sh]
        assert stride >= sum(r.var_bytes for _, r in slices_here)
        for w in range(W):
            for b in range(depth):
                lo = al.mbox_slot_base[sh] + (w * depth + b) * stride
                spans.append((lo, lo + stride))
        for n_, r in slices_here:        # a slice's place inside a slot lies within the slot
            off = al.mailbox_off[(n_, sh)] - al.mbox_slot_base[sh]
            assert 0 <= off and off + r.var_bytes <= stride and off % 256 == 0
        spans.sort()
        for (a0, a1), (b0, b1) in zip(spans, spans[1:]):
            assert a1 <= b0, f"shard {sh}: regions [{a0},{a1}) and [{b0},{b1}) overlap"
        assert spans[-1][1] <= al.shard_bytes[sh]
    # synchronous layout (depth 1 + mirrors at the same offset in every arena) still lines up
    sync = build_arena_layout(layout, 8, spec, names, sync_workers=8)
    assert sync.mailbox_depth == 1 and set(sync.mirror_f32) == set(n for n in sync.regions if sync.regions[n])
    assert len(set(sync.shard_bytes)) == 1
