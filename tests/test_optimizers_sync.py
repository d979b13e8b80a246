"""TF1 optimizer rules vs a float64 numpy restatement, shared beta powers, and
SyncReplicasOptimizer semantics on the in-process cluster."""
import math
import threading
import time

import numpy as np
import pytest
import torch

import distributedtensorflow_b200 as tf
from distributedtensorflow_b200.parallel.ps_core import ConditionalAccumulator, FIFOQueue, OutOfRangeError, PsShard
from distributedtensorflow_b200.train.optimizers import Optimizer, OptimizerSpec


def _ref_step(kind, var, s0, s1, g, h, t):
    if kind == "sgd":
        return var - h["lr"] * g, s0, s1
    if kind == "momentum":
        s0 = s0 * h["momentum"] + g
        return var - h["lr"] * s0, s0, s1
    if kind == "adagrad":
        s0 = s0 + g * g
        return var - h["lr"] * g / np.sqrt(s0), s0, s1
    b1, b2 = h["beta1"], h["beta2"]
    lr_t = h["lr"] * math.sqrt(1 - b2 ** t) / (1 - b1 ** t)
    s0 = b1 * s0 + (1 - b1) * g
    s1 = b2 * s1 + (1 - b2) * g * g
    return var - lr_t * s0 / (np.sqrt(s1) + h["eps"]), s0, s1


@pytest.mark.parametrize("kind", ["sgd", "momentum", "adagrad", "adam"])
def test_apply_dense_matches_fp64_oracle(kind):
    rng = np.random.default_rng(0)
    h = dict(lr=0.01, beta1=0.9, beta2=0.999, eps=1e-8, momentum=0.9)
    spec = OptimizerSpec(kind, h["lr"], h["beta1"], h["beta2"], h["eps"], h["momentum"])
    var = rng.standard_normal(1000)
    s0 = np.full(1000, spec.slot_init(0))
    s1 = np.zeros(1000)
    tv, t0, t1 = (torch.tensor(a, dtype=torch.float32) for a in (var, s0, s1))
    powers = {"beta1_power": torch.tensor(h["beta1"]), "beta2_power": torch.tensor(h["beta2"])}
    for t in range(1, 6):
        g = rng.standard_normal(1000)
        var, s0, s1 = _ref_step(kind, var, s0, s1, g, h, t)
        Optimizer.apply_dense(spec, tv, torch.tensor(g, dtype=torch.float32), [t0, t1][:spec.num_slots], powers)
        Optimizer.finish(spec, powers)
        assert np.abs(tv.numpy() - var).max() < 2e-5
    if kind == "adam":
        assert abs(float(powers["beta1_power"]) - 0.9 ** 6) < 1e-6


def test_slot_names_follow_tf_checkpoint_convention():
    assert tf.train.AdamOptimizer().spec().slot_names() == ["Adam", "Adam_1"]
    assert tf.train.AdagradOptimizer(0.1).spec().slot_names() == ["Adagrad"]
    assert tf.train.MomentumOptimizer(0.1, 0.9).spec().slot_names() == ["Momentum"]
    assert tf.train.GradientDescentOptimizer(0.1).spec().slot_names() == []
    assert tf.train.AdamOptimizer().spec().nonslot_names() == ["beta1_power", "beta2_power"]
    with pytest.raises(ValueError):
        tf.train.AdagradOptimizer(0.1, initial_accumulator_value=0.0)


def test_conditional_accumulator_mean_stale_drop_and_blocking():
    acc = ConditionalAccumulator((3,))
    assert acc.apply_grad(torch.tensor([1.0, 2.0, 3.0]), 0)
    assert acc.apply_grad(torch.tensor([3.0, 2.0, 1.0]), 5)      # newer than the global step is fine
    assert acc.num_accumulated() == 2
    out = acc.take_grad(2)
    assert torch.allclose(out, torch.tensor([2.0, 2.0, 2.0])) and acc.global_step == 1
    assert not acc.apply_grad(torch.ones(3), 0) and acc.dropped == 1   # stale: silently dropped
    acc.set_global_step(4)
    assert not acc.apply_grad(torch.ones(3), 3) and acc.apply_grad(torch.ones(3), 4)
    got = {}
    t = threading.Thread(target=lambda: got.setdefault("v", acc.take_grad(2)), daemon=True)
    t.start()
    time.sleep(0.05)
    assert t.is_alive()                                         # blocks until 2 fresh gradients
    acc.apply_grad(torch.full((3,), 3.0), 4)
    t.join(2)
    assert torch.allclose(got["v"], torch.full((3,), 2.0))
    with pytest.raises(TimeoutError):
        acc.take_grad(1, timeout=0.05)


def test_fifo_queue():
    q = FIFOQueue()
    q.enqueue_many([1, 2])
    assert q.dequeue() == 1 and q.size() == 1 and q.dequeue() == 2
    with pytest.raises(TimeoutError):
        q.dequeue(timeout=0.02)
    q.close()
    with pytest.raises(OutOfRangeError):
        q.dequeue()


def _mlp_graph(num_ps, sync=None, lr=0.05, opt="sgd"):
    from distributedtensorflow_b200.models import mnist_mlp
    tf.reset_default_graph()
    tf.set_random_seed(11)
    cluster = tf.train.ClusterSpec({"ps": ["h:%d" % i for i in range(num_ps)], "worker": ["w:1", "w:2", "w:3"]})
    with tf.device(tf.train.replica_device_setter(worker_device="/job:worker/task:0", cluster=cluster)):
        x = tf.placeholder(tf.float32, [None, 784])
        y_ = tf.placeholder(tf.float32, [None, 10])
        gs, xent, _ = mnist_mlp.build_model(x, y_, hidden_units=8)
        o = tf.train.GradientDescentOptimizer(lr) if opt == "sgd" else tf.train.AdamOptimizer(lr)
        if sync is not None:
            o = tf.train.SyncReplicasOptimizer(o, replicas_to_aggregate=sync[0], total_num_replicas=sync[1],
                                               name="mnist_sync_replicas")
        train_op = o.minimize(xent, global_step=gs)
    return x, y_, gs, xent, train_op, o


def _data(n, seed):
    rng = np.random.default_rng(seed)
    return rng.random((n, 784), dtype=np.float32), np.eye(10, dtype=np.float32)[rng.integers(0, 10, n)]


def test_shared_beta_powers_count_applies_from_all_workers():
    x, y_, gs, xent, train_op, _ = _mlp_graph(2, opt="adam", lr=0.001)
    server = tf.train.Server.create_local_server(num_ps=2)
    cfg = tf.ConfigProto(backend="cpu")
    bx, by = _data(10, 0)
    chief = tf.train.MonitoredTrainingSession(master=server.target, is_chief=True, config=cfg)
    other = tf.train.MonitoredTrainingSession(master=server.target, is_chief=False, config=cfg,
                                              recovery_wait_secs=0.01)
    for s in (chief, other, chief, other, other):
        s.run(train_op, feed_dict={x: bx, y_: by})
    assert int(chief.run(gs)) == 5                      # global_step counts applies from ALL workers
    shard = server.local_cluster.shards
    powers = [sh.read_powers() for sh in shard if sh.powers]
    assert len(powers) == 1                             # one shared copy, colocated with the first variable
    assert abs(powers[0]["beta1_power"] - 0.9 ** 6) < 1e-6
    chief.close()
    other.close()


def test_sync_replicas_equals_large_batch_sgd():
    """N replicas each pushing the gradient of their shard of the batch, aggregated (mean) and
    applied once per round == single-process SGD on the mean gradient (bitwise on CPU up to the
    order of fp32 additions: compared with a tight tolerance)."""
    N, rounds = 3, 4
    data = [_data(8, s) for s in range(N)]
    # reference: one process, mean of the three per-replica SUM losses' gradients
    x, y_, gs, xent, train_op, _ = _mlp_graph(1, lr=0.05)
    server = tf.train.Server.create_local_server()
    with tf.Session(server.target, config=tf.ConfigProto(backend="cpu")) as sess:
        sess.run(tf.global_variables_initializer())
        names = [v.name for v in tf.trainable_variables()]
        from distributedtensorflow_b200.framework.executor_cpu import evaluate
        for _ in range(rounds):
            vals = sess.ps_client.pull_all(sess.layout, names)
            grads = {n: np.zeros_like(v) for n, v in vals.items()}
            for bx, by in data:
                leaves = {n: torch.tensor(v, requires_grad=True) for n, v in vals.items()}
                loss = evaluate([xent], {x: bx, y_: by}, leaves)[0]
                for n, g in zip(names, torch.autograd.grad(loss, [leaves[n] for n in names])):
                    grads[n] += g.numpy() / N
            sess.ps_client.push_apply(sess.layout, tf.train.GradientDescentOptimizer(0.05).spec(), grads, "global_step")
        ref = sess.ps_client.pull_all(sess.layout, names)
    # sync replicas on a 2-PS in-process cluster, one thread per worker
    x, y_, gs, xent, train_op, opt = _mlp_graph(2, sync=(N, N), lr=0.05)
    server = tf.train.Server.create_local_server(num_ps=2)
    cfg = tf.ConfigProto(backend="cpu")
    results, errors = {}, []

    def worker(i):
        try:
            # no initial tokens: nobody runs ahead, every round holds exactly one gradient per
            # replica (with TF's default priming each replica's 2nd gradient arrives stale)
            hook = opt.make_session_run_hook(is_chief=(i == 0), num_tokens=0)
            with tf.train.MonitoredTrainingSession(master=server.target, is_chief=(i == 0), hooks=[hook],
                                                   config=cfg, recovery_wait_secs=0.01) as sess:
                for _ in range(rounds):
                    _, step = sess.run([train_op, gs], feed_dict={x: data[i][0], y_: data[i][1]})
                results[i] = int(step)
                if i == 0:
                    raw = sess._tf_sess()
                    # the last round's apply runs on the chief's queue-runner thread: wait for it
                    # (bounded) instead of sleeping a fixed time - the box may be busy
                    t_end = time.time() + 20
                    while time.time() < t_end and raw.ps_client.read_global_step(raw.layout, "global_step") < rounds:
                        time.sleep(0.01)
                    time.sleep(0.05)
                    results["vars"] = raw.ps_client.pull_all(raw.layout, names)
        except Exception as e:  # pragma: no cover
            errors.append(e)

    threads = [threading.Thread(target=worker, args=(i,), daemon=True) for i in range(N)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(60)
    assert not errors, errors
    assert results[0] == rounds and all(results[i] <= rounds for i in range(N))
    for n in names:
        assert np.abs(results["vars"][n] - ref[n]).max() < 1e-5, n


def test_sync_backup_workers_and_token_accounting():
    """replicas_to_aggregate < total_num_replicas: a round closes after 2 of 3 gradients; the
    straggler's gradient arrives stale and is dropped."""
    x, y_, gs, xent, train_op, opt = _mlp_graph(1, sync=(2, 3))
    with pytest.raises(ValueError):
        opt.get_init_tokens_op(num_tokens=-2)
    server = tf.train.Server.create_local_server()
    sess = tf.Session(server.target, config=tf.ConfigProto(backend="cpu"))
    sess.run(tf.global_variables_initializer())
    sess.run(opt.chief_init_op)
    sess.run(opt.get_init_tokens_op())             # primes replicas_to_aggregate = 2 tokens
    client, layout = sess.ps_client, sess.layout
    shard = server.local_cluster.shards[0]
    assert shard.queue_size("sync_token_q") == 2
    names = [v.name for v in tf.trainable_variables()]
    g = {n: np.ones(layout[n].shape, np.float32) for n in names}
    before = client.pull_all(layout, names)
    # two fast replicas push (local_step 0), the chief aggregates once
    for _ in range(2):
        for n in names:
            client.call(0, "accum_apply", {n: g[n].reshape(-1)}, 0)
    assert opt.run_sync_op(client, layout, timeout=1.0)
    assert int(sess.run(gs)) == 1
    assert shard.queue_size("sync_token_q") == 2 + 3   # tokens_per_step = max(total, to_aggregate)
    after = client.pull_all(layout, names)
    lr = 0.05
    for n in names:
        assert np.allclose(after[n], before[n] - lr * 1.0, atol=1e-6)   # mean of two identical grads
    # the straggler still carries local_step 0 < accumulator step 1 -> dropped, nothing pending
    assert client.call(0, "accum_apply", {names[0]: g[names[0]].reshape(-1)}, 0) == 0
    assert shard.accum_stats()[names[0]][0] == 0 and shard.accum_stats()[names[0]][2] == 1
    assert not opt.run_sync_op(client, layout, timeout=0.05)           # no fresh gradients: times out
    sess.close()


def test_shard_take_is_all_or_nothing_on_timeout():
    """accum_take_apply first waits (consuming nothing) until EVERY accumulator of the shard holds
    enough gradients: a timeout on the second variable must not discard the first one's mean nor
    advance its step (which would make its next gradients look stale)."""
    from distributedtensorflow_b200.parallel.ps_core import PsShard
    from distributedtensorflow_b200.parallel.transport import spec_to_dict
    sh = PsShard(0)
    spec = spec_to_dict(tf.train.GradientDescentOptimizer(0.5).spec())
    sh.create([dict(name="a", first=0, n=2, dtype="float32"), dict(name="b", first=0, n=2, dtype="float32")], spec)
    sh.accum_create([("a", 2), ("b", 2)])
    sh.accum_apply({"a": np.asarray([1.0, 2.0], np.float32)}, 0)        # only `a` has a gradient
    with pytest.raises(TimeoutError):
        sh.accum_take_apply(["a", "b"], 1, spec, None, timeout=0.05)
    stats = sh.accum_stats()
    assert stats["a"][0] == 1 and stats["b"][0] == 0                    # a's gradient is still there ...
    assert stats["a"][1] == 0                                            # ... and its step did not move
    sh.accum_apply({"b": np.asarray([3.0, 4.0], np.float32)}, 0)
    assert sh.accum_take_apply(["a", "b"], 1, spec, None, timeout=1.0) == 2
    np.testing.assert_allclose(sh.pull(["a", "b"]), [[-0.5, -1.0], [-1.5, -2.0]])
