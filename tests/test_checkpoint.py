"""TF1 tensor-bundle checkpoints: golden bytes assembled by hand from the format spec, writer <->
reader round trip, sharded save + merge, re-sharding on restore, checkpoint-state file
(SURVEY.md §4 item 1, §5.4)."""
import os
import struct

import numpy as np
import pytest

import distributedtensorflow_b200 as tf
from distributedtensorflow_b200.ckpt import BundleReader, BundleWriter, merge_bundles
from distributedtensorflow_b200.train import saver as S
from distributedtensorflow_b200.utils import events, hostlib
from distributedtensorflow_b200.utils import protowire as pw


def _masked(b: bytes) -> int:
    return events.mask_crc(events.crc32c_py(b))


def _golden_index(entries):
    """Independent python restatement of the LevelDB table layout for a single data block."""
    def block(kvs, interval):
        out, restarts, last, cnt = b"", [0], b"", 0
        for k, v in kvs:
            shared = 0
            if cnt < interval:
                while shared < min(len(last), len(k)) and last[shared] == k[shared]:
                    shared += 1
            else:
                restarts.append(len(out))
                cnt = 0
            out += pw.varint(shared) + pw.varint(len(k) - shared) + pw.varint(len(v)) + k[shared:] + v
            last, cnt = k, cnt + 1
        for r in restarts:
            out += struct.pack("<I", r)
        return out + struct.pack("<I", len(restarts))

    def with_trailer(b):
        return b + b"\x00" + struct.pack("<I", _masked(b + b"\x00"))

    data = block(entries, 16)
    meta = block([], 1)
    off = 0
    f = with_trailer(data)
    data_handle = pw.varint(0) + pw.varint(len(data))
    off += len(data) + 5
    f += with_trailer(meta)
    meta_handle = pw.varint(off) + pw.varint(len(meta))
    off += len(meta) + 5
    index = block([(entries[-1][0], data_handle)], 1)
    f += with_trailer(index)
    index_handle = pw.varint(off) + pw.varint(len(index))
    footer = meta_handle + index_handle
    footer += b"\x00" * (40 - len(footer)) + struct.pack("<Q", 0xdb4775248b80fb57)
    return f + footer


def test_golden_index_bytes(tmp_path):
    prefix = str(tmp_path / "model.ckpt-7")
    a = np.arange(6, dtype=np.float32).reshape(2, 3)
    gs = np.asarray(7, np.int64)
    with BundleWriter(prefix) as w:
        w.add("hid_w", a)
        w.add("global_step", gs)
    data = open(prefix + ".data-00000-of-00001", "rb").read()
    assert data == a.tobytes() + gs.tobytes()            # raw little-endian bytes in add order
    header = pw.field_varint(1, 1) + pw.field_bytes(3, pw.field_varint(1, 1))
    shape_a = pw.field_bytes(2, pw.field_varint(1, 2)) + pw.field_bytes(2, pw.field_varint(1, 3))
    e_a = (pw.field_varint(1, 1) + pw.field_bytes(2, shape_a) + pw.field_varint(5, 24)
           + pw.field_fixed32(6, _masked(a.tobytes())))
    e_gs = (pw.field_varint(1, 9) + pw.field_bytes(2, b"") + pw.field_varint(4, 24) + pw.field_varint(5, 8)
            + pw.field_fixed32(6, _masked(gs.tobytes())))
    golden = _golden_index([(b"", header), (b"global_step", e_gs), (b"hid_w", e_a)])
    assert open(prefix + ".index", "rb").read() == golden


def test_bundle_round_trip_many_tensors_and_blocks(tmp_path):
    prefix = str(tmp_path / "big")
    rng = np.random.default_rng(0)
    tensors = {"var_%04d/Adam_1" % i: rng.standard_normal((i % 7 + 1, 3)).astype(np.float32) for i in range(300)}
    tensors["step"] = np.asarray(12345678901, np.int64)
    tensors["i32"] = np.arange(10, dtype=np.int32)
    tensors["empty"] = np.zeros((0, 4), np.float32)
    lib = hostlib.lib()
    with BundleWriter(prefix) as w:
        for k in rng.permutation(list(tensors)):       # insertion order need not be sorted
            w.add(k, tensors[k])
    with BundleReader(prefix) as r:
        assert r.num_shards == 1 and sorted(r.keys()) == sorted(tensors)
        assert r.keys() == sorted(tensors)             # table keys come back in bytewise order
        for k, v in tensors.items():
            got = r.get_tensor(k)
            assert got.dtype == v.dtype and got.shape == v.shape and np.array_equal(got, v)
        with pytest.raises(KeyError):
            r.get_tensor("missing")
    # a small block size forces many data blocks + a multi-entry index block
    import ctypes as C
    path = str(tmp_path / "t.ldb")
    tb = lib.dtf_table_builder_open(path.encode(), 64)
    keys = [b"key%05d" % i for i in range(500)]
    for k in keys:
        v = k * 3
        assert lib.dtf_table_builder_add(tb, k, len(k), v, len(v)) == 0
    assert lib.dtf_table_builder_add(tb, b"a", 1, b"", 0) != 0   # out-of-order key is refused
    assert lib.dtf_table_builder_finish(tb) == 0


def test_checksum_mismatch_is_detected(tmp_path):
    prefix = str(tmp_path / "c")
    with BundleWriter(prefix) as w:
        w.add("x", np.ones(100, np.float32))
    p = prefix + ".data-00000-of-00001"
    raw = bytearray(open(p, "rb").read())
    raw[17] ^= 1
    open(p, "wb").write(raw)
    with BundleReader(prefix) as r:
        with pytest.raises(Exception, match="checksum"):
            r.get_tensor("x")
        assert r.get_tensor("x", verify_crc=False).shape == (100,)


def test_merge_shards(tmp_path):
    parts = [str(tmp_path / "tmp" / ("part-%05d-of-00003" % i)) for i in range(3)]
    expect = {}
    for i, p in enumerate(parts):
        with BundleWriter(p) as w:
            for j in range(3):
                name = "v%d_%d" % (i, j)
                expect[name] = (np.full((4, j + 1), i * 10 + j, np.float32), i)
                w.add(name, expect[name][0])
    merged = str(tmp_path / "model.ckpt-5")
    merge_bundles(parts, merged)
    assert sorted(os.listdir(tmp_path)) == ["model.ckpt-5.data-00000-of-00003", "model.ckpt-5.data-00001-of-00003",
                                            "model.ckpt-5.data-00002-of-00003", "model.ckpt-5.index"]
    with BundleReader(merged) as r:
        assert r.num_shards == 3
        for name, (val, shard) in expect.items():
            assert r.entries[name]["shard_id"] == shard
            assert np.array_equal(r.get_tensor(name), val)


def test_checkpoint_state_file(tmp_path):
    d = str(tmp_path)
    assert S.latest_checkpoint(d) is None
    for step in (10, 20):
        prefix = os.path.join(d, "model.ckpt-%d" % step)
        with BundleWriter(prefix) as w:
            w.add("x", np.zeros(1, np.float32))
        S.update_checkpoint_state(d, prefix, [os.path.join(d, "model.ckpt-10")] if step == 20 else None)
    text = open(os.path.join(d, "checkpoint")).read()
    assert text == ('model_checkpoint_path: "model.ckpt-20"\n'
                    'all_model_checkpoint_paths: "model.ckpt-10"\n'
                    'all_model_checkpoint_paths: "model.ckpt-20"\n')
    st = S.get_checkpoint_state(d)
    assert st.model_checkpoint_path == os.path.join(d, "model.ckpt-20")
    assert S.latest_checkpoint(d) == os.path.join(d, "model.ckpt-20")


def _train_graph(num_ps, opt="adam"):
    from distributedtensorflow_b200.models import mnist_mlp
    tf.reset_default_graph()
    tf.set_random_seed(5)
    cluster = tf.train.ClusterSpec({"ps": ["h:%d" % i for i in range(num_ps)], "worker": ["w:1"]})
    with tf.device(tf.train.replica_device_setter(worker_device="/job:worker/task:0", cluster=cluster)):
        x = tf.placeholder(tf.float32, [None, 784])
        y_ = tf.placeholder(tf.float32, [None, 10])
        gs, xent, _ = mnist_mlp.build_model(x, y_, hidden_units=16)
        o = tf.train.AdamOptimizer(0.01) if opt == "adam" else tf.train.AdagradOptimizer(0.01)
        train_op = o.minimize(xent, global_step=gs)
    return x, y_, gs, xent, train_op


def test_saver_sharded_save_restore_and_reshard(tmp_path):
    """Save on 3 PS shards (one data file per shard, TF tensor names), restore on 1."""
    rng = np.random.default_rng(1)
    bx = rng.random((20, 784), dtype=np.float32)
    by = np.eye(10, dtype=np.float32)[rng.integers(0, 10, 20)]
    x, y_, gs, xent, train_op = _train_graph(3)
    server = tf.train.Server.create_local_server(num_ps=3)
    ckdir = str(tmp_path / "ck")
    with tf.train.MonitoredTrainingSession(master=server.target, checkpoint_dir=ckdir, save_checkpoint_steps=2,
                                           save_summaries_steps=None, log_step_count_steps=None,
                                           config=tf.ConfigProto(backend="cpu")) as sess:
        for _ in range(5):
            sess.run(train_op, feed_dict={x: bx, y_: by})
        raw = sess._tf_sess()
        before = raw.ps_client.pull_all(raw.layout, ["hid_w", "hid_b", "sm_w", "sm_b"])
    ck = tf.train.latest_checkpoint(ckdir)
    assert ck.endswith("model.ckpt-5")
    with BundleReader(ck) as r:
        assert r.num_shards == 3
        keys = set(r.keys())
        assert {"global_step", "hid_w", "hid_w/Adam", "hid_w/Adam_1", "sm_b/Adam_1", "beta1_power",
                "beta2_power"} <= keys
        assert int(r.get_tensor("global_step")) == 5 and r.get_tensor("global_step").dtype == np.int32
        assert abs(float(r.get_tensor("beta1_power")) - 0.9 ** 6) < 1e-6
        # round-robin: global_step, hid_b, sm_b... land in different data shards
        assert r.entries["hid_w"]["shard_id"] == 1 and r.entries["hid_b"]["shard_id"] == 2
        assert np.array_equal(r.get_tensor("hid_w"), before["hid_w"])
    # restore into a 1-PS cluster and continue: the chief restores instead of re-initialising
    x, y_, gs, xent, train_op = _train_graph(1)
    server1 = tf.train.Server.create_local_server(num_ps=1)
    with tf.train.MonitoredTrainingSession(master=server1.target, checkpoint_dir=ckdir, save_checkpoint_steps=100,
                                           save_summaries_steps=None, log_step_count_steps=None,
                                           config=tf.ConfigProto(backend="cpu")) as sess:
        assert int(sess.run(gs)) == 5
        raw = sess._tf_sess()
        after = raw.ps_client.pull_all(raw.layout, ["hid_w", "sm_b"])
        assert np.array_equal(after["hid_w"], before["hid_w"]) and np.array_equal(after["sm_b"], before["sm_b"])
        sess.run(train_op, feed_dict={x: bx, y_: by})
        assert int(sess.run(gs)) == 6
    assert tf.train.latest_checkpoint(ckdir).endswith("model.ckpt-6")   # saved again at end()


def test_saver_max_to_keep(tmp_path):
    x, y_, gs, xent, train_op = _train_graph(1, opt="adagrad")
    server = tf.train.Server.create_local_server()
    sv = tf.train.Saver(max_to_keep=2)
    with tf.Session(server.target, config=tf.ConfigProto(backend="cpu")) as sess:
        sess.run(tf.global_variables_initializer())
        for i in range(4):
            sv.save(sess, str(tmp_path / "model.ckpt"), global_step=i)
    names = sorted(f for f in os.listdir(tmp_path) if f.endswith(".index"))
    assert names == ["model.ckpt-2.index", "model.ckpt-3.index"]
    with BundleReader(str(tmp_path / "model.ckpt-3")) as r:
        assert "hid_w/Adagrad" in r.keys() and np.allclose(r.get_tensor("hid_w/Adagrad"), 0.1)


def test_crc32c_combine_equals_checksum_of_the_concatenation():
    """Every PS task checksums only the slice it wrote; the chief derives the tensor's crc32c for
    the index from the pieces (dtf_crc32c_combine) without ever seeing the bytes."""
    from distributedtensorflow_b200.ckpt import bundle as B
    rng = np.random.RandomState(3)
    for lens in [(1, 1), (0, 5), (5, 0), (1000, 77777, 3), (4096, 4096, 4096, 1)]:
        parts = [rng.bytes(n) for n in lens]
        whole = B.crc32c(np.frombuffer(b"".join(parts), np.uint8))
        crc = B.crc32c(np.frombuffer(parts[0], np.uint8))
        for p in parts[1:]:
            crc = B.crc32c_combine(crc, B.crc32c(np.frombuffer(p, np.uint8)), len(p))
        assert crc == whole, lens


def test_index_only_writer_describes_data_files_written_elsewhere(tmp_path):
    """The parallel sharded save: data files are filled with pwrite by their owners, the index is
    written from (shard, offset, size, crc) alone and the result is an ordinary V2 checkpoint."""
    from distributedtensorflow_b200.ckpt import bundle as B
    prefix = str(tmp_path / "model.ckpt-7")
    w = B.BundleIndexWriter(prefix, 2)
    kernel = np.arange(24, dtype=np.float32).reshape(6, 4)
    step = np.asarray(7, np.int64)
    paths = [w.data_path(i) for i in range(2)]
    assert paths[1].endswith(".data-00001-of-00002")
    for p in paths:
        open(p, "wb").close()
    # two "PS tasks" each store their row slice of the kernel into shard 0, out of order
    top, bottom = kernel[:2], kernel[2:]
    fd = os.open(paths[0], os.O_WRONLY)
    os.pwrite(fd, bottom.tobytes(), top.nbytes)
    os.pwrite(fd, top.tobytes(), 0)
    os.close(fd)
    with open(paths[1], "r+b") as f:
        f.write(b"\0" * 16 + step.tobytes())
    crc = B.crc32c_combine(B.crc32c(top), B.crc32c(bottom), bottom.nbytes)
    w.add("dense/kernel", B.DT_FLOAT, kernel.shape, 0, 0, kernel.nbytes, crc)
    w.add("global_step", B.DT_INT64, (), 1, 16, 8, B.crc32c(step))
    w.finish()
    with B.BundleReader(prefix) as r:
        assert r.num_shards == 2 and sorted(r.keys()) == ["dense/kernel", "global_step"]
        np.testing.assert_array_equal(r.get_tensor("dense/kernel"), kernel)      # crc verified
        assert int(r.get_tensor("global_step")) == 7
    w2 = B.BundleIndexWriter(str(tmp_path / "bad"), 1)
    with pytest.raises(Exception):
        w2.add("x", B.DT_FLOAT, (1,), 3, 0, 4, 0)        # shard id outside the declared shards


def test_saver_writes_a_meta_graph_with_the_between_graph_placement(tmp_path):
    """Saver.save exports <prefix>.meta like TF1: one VariableV2 node per variable carrying the
    device replica_device_setter chose, the variables / trainable_variables collections, a V2 SaverDef."""
    from distributedtensorflow_b200.train.meta_graph import read_meta_graph
    from distributedtensorflow_b200.models import mnist_mlp
    tf.reset_default_graph()
    cluster = tf.train.ClusterSpec({"ps": ["h:1", "h:2"], "worker": ["w:1"]})
    with tf.device(tf.train.replica_device_setter(worker_device="/job:worker/task:0", cluster=cluster)):
        x = tf.placeholder(tf.float32, [None, 784])
        y_ = tf.placeholder(tf.float32, [None, 10])
        gs, xent, _ = mnist_mlp.build_model(x, y_, hidden_units=100)
        train_op = tf.train.AdamOptimizer(0.01).minimize(xent, global_step=gs)
    server = tf.train.Server.create_local_server(num_ps=2)
    with tf.train.MonitoredTrainingSession(master=server.target, config=tf.ConfigProto(backend="cpu")) as sess:
        raw = sess._tf_sess()
        path = tf.train.Saver(max_to_keep=3).save(raw, str(tmp_path / "model.ckpt"), global_step=0)
    server.stop()
    assert os.path.exists(path + ".meta") and os.path.exists(path + ".index")
    mg = read_meta_graph(path + ".meta")
    by_name = {n["name"]: n for n in mg["nodes"]}
    assert all(n["op"] == "VariableV2" for n in mg["nodes"])
    assert by_name["global_step"]["dtype"] == "int32" and by_name["global_step"]["shape"] == ()
    assert by_name["hid_w"]["shape"] == (784, 100) and by_name["hid_w"]["dtype"] == "float32"
    # round-robin in creation order over the two PS tasks
    devices = [by_name[n]["device"] for n in ("global_step", "hid_w", "hid_b", "sm_w", "sm_b")]
    assert devices == ["/job:ps/task:0", "/job:ps/task:1", "/job:ps/task:0", "/job:ps/task:1", "/job:ps/task:0"]
    assert mg["collections"]["trainable_variables"] == ["hid_w:0", "hid_b:0", "sm_w:0", "sm_b:0"]
    assert "global_step:0" in mg["collections"]["variables"]
    assert mg["saver"] == dict(max_to_keep=3, sharded=True, version=2, filename_tensor_name="save/Const:0")
    assert mg["producer"] == 27 and mg["tags"] == ["train"]
    S.remove_checkpoint(path)
    assert not os.path.exists(path + ".meta")
