"""End-to-end training through the public session API on the B200 backend, checked against the
CPU executor (PyTorch fp32 autograd) started from the same initial values."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run(backend, build, steps, batches, num_ps=1, strategy=None, options=None, probe=None):
    import distributedtensorflow_b200 as tf
    tf.reset_default_graph()
    tf.set_random_seed(1234)
    server = tf.train.Server.create_local_server(num_ps=num_ps)
    cluster = tf.train.ClusterSpec({"ps": ["localhost:%d" % (1000 + i) for i in range(num_ps)],
                                    "worker": ["localhost:2000"]})
    ps_strategy = None
    if strategy == "flat":
        ps_strategy = tf.train.FlatSlicedStrategy(num_ps)
    with tf.device(tf.train.replica_device_setter(worker_device="/job:worker/task:0", cluster=cluster,
                                                  ps_strategy=ps_strategy)):
        x_ph, y_ph, loss, gs, train_op = build(tf)
    cfg = tf.ConfigProto(backend=backend, device="cuda:0" if backend == "b200" else None, **(options or {}))
    losses, gsteps = [], []
    with tf.train.MonitoredTrainingSession(master=server.target, is_chief=True, config=cfg) as sess:
        for i in range(steps):
            bx, by = batches[i % len(batches)]
            _, l, g = sess.run([train_op, loss, gs], feed_dict={x_ph: bx, y_ph: by})
            losses.append(float(l))
            gsteps.append(int(g))
        raw = sess._tf_sess()
        names = [v.name for v in tf.trainable_variables()]
        if backend == "b200":
            raw._runner.flush_pushes()
            finals = {n: raw._runner.store.read_variable(n) for n in names}
            shadows = {n: raw._runner.store.read_shadow(n) for n in names if len(finals[n].shape) >= 2}
            if probe is not None:
                probe(raw._runner, finals)
        else:
            finals = raw.ps_client.pull_all(raw.layout, names)
            shadows = {}
    server.stop()
    return losses, gsteps, finals, shadows


def _build_cnn(tf):
    from distributedtensorflow_b200.models import mnist_cnn
    x = tf.placeholder(tf.float32, [None, 784])
    y = tf.placeholder(tf.int32, [None])
    _, loss, gs, train_op = mnist_cnn.build_train_graph(x, y, learning_rate=0.0005)
    return x, y, loss, gs, train_op


def _build_mlp(opt_name):
    def build(tf):
        from distributedtensorflow_b200.models import mnist_mlp
        x = tf.placeholder(tf.float32, [None, 784])
        y_ = tf.placeholder(tf.float32, [None, 10])
        gs, xent, _ = mnist_mlp.build_model(x, y_, hidden_units=100)
        opt = {"adam": lambda: tf.train.AdamOptimizer(0.01),
               "sgd": lambda: tf.train.GradientDescentOptimizer(0.001),
               "momentum": lambda: tf.train.MomentumOptimizer(0.001, 0.9),
               "adagrad": lambda: tf.train.AdagradOptimizer(0.01)}[opt_name]()
        train_op = opt.minimize(xent, global_step=gs)
        return x, y_, xent, gs, train_op
    return build


def _batches(n, B, one_hot):
    from distributedtensorflow_b200.data import mnist
    imgs, labels = mnist.synthetic_arrays(n * B, seed=3)
    x = imgs.astype(np.float32) / 255.0
    out = []
    for i in range(n):
        lab = labels[i * B:(i + 1) * B]
        y = np.eye(10, dtype=np.float32)[lab] if one_hot else lab.astype(np.int32)
        out.append((x[i * B:(i + 1) * B], y))
    return out


@pytest.mark.parametrize("opt", ["adam", "sgd", "momentum", "adagrad"])
def test_mlp_matches_cpu_oracle(opt):
    batches = _batches(4, 100, one_hot=True)
    ref_l, ref_g, ref_v, _ = _run("cpu", _build_mlp(opt), 12, batches)
    got_l, got_g, got_v, shadows = _run("b200", _build_mlp(opt), 12, batches)
    assert got_g == ref_g == list(range(1, 13))
    for a, b in zip(got_l, ref_l):
        assert abs(a - b) <= 0.05 * abs(b) + 0.5, (got_l, ref_l)
    assert got_l[-1] < got_l[0]
    # Adam's update is ~lr*sign(g) where |g| is tiny, so bf16 rounding flips individual elements:
    # compare the bulk (mean abs error) tightly and the worst element loosely
    steps_lr = 12 * {"adam": 0.01, "sgd": 0.001, "momentum": 0.001, "adagrad": 0.01}[opt]
    for n in ref_v:
        diff = np.abs(got_v[n] - ref_v[n])
        scale = max(1e-3, float(np.abs(ref_v[n]).max()))
        assert float(diff.mean()) < 0.02 * scale + 0.05 * steps_lr, n
        assert float(diff.max()) < 0.1 * scale + 1.5 * steps_lr, n
    for n, sh in shadows.items():  # the bf16 serving copy tracks the fp32 master
        assert float(np.abs(sh - got_v[n]).max()) <= 0.01 * float(np.abs(got_v[n]).max()) + 1e-6


def test_cnn_matches_cpu_oracle():
    batches = _batches(3, 128, one_hot=False)
    ref_l, ref_g, ref_v, _ = _run("cpu", _build_cnn, 8, batches)
    got_l, got_g, got_v, _ = _run("b200", _build_cnn, 8, batches)
    assert got_g == ref_g == list(range(1, 9))
    for a, b in zip(got_l, ref_l):
        assert abs(a - b) <= 0.06 * abs(b) + 0.03, (got_l, ref_l)
    assert got_l[-1] < got_l[0]


def test_cnn_sharded_round_robin_and_flat_sliced_agree():
    """8 shards on one GPU: TF round-robin placement and the flat-sliced layout train alike."""
    batches = _batches(2, 128, one_hot=False)
    base_l, _, base_v, _ = _run("b200", _build_cnn, 6, batches, num_ps=1)
    rr_l, rr_g, rr_v, _ = _run("b200", _build_cnn, 6, batches, num_ps=8)
    fl_l, fl_g, fl_v, _ = _run("b200", _build_cnn, 6, batches, num_ps=8, strategy="flat")
    assert rr_g == fl_g == list(range(1, 7))
    for a, b, c in zip(base_l, rr_l, fl_l):
        assert abs(a - b) <= 0.03 * abs(a) + 0.02 and abs(a - c) <= 0.03 * abs(a) + 0.02


def test_eager_mode_equals_cuda_graph_mode():
    batches = _batches(2, 128, one_hot=False)
    a, _, va, _ = _run("b200", _build_cnn, 5, batches, options={"cuda_graph": True})
    b, _, vb, _ = _run("b200", _build_cnn, 5, batches, options={"cuda_graph": False})
    for x, y in zip(a, b):
        assert abs(x - y) <= 0.02 * abs(x) + 0.01


def test_sync_round_kernel_single_replica_equals_async_adam():
    """With one replica a synchronous round (mailbox -> mean -> apply -> all-gather into the
    mirror, csrc/kernels/sync_step.cu) must train exactly like the asynchronous push+apply."""
    def build_sync(tf):
        from distributedtensorflow_b200.models import mnist_mlp
        x = tf.placeholder(tf.float32, [None, 784])
        y_ = tf.placeholder(tf.float32, [None, 10])
        gs, xent, _ = mnist_mlp.build_model(x, y_, hidden_units=100)
        opt = tf.train.SyncReplicasOptimizer(tf.train.AdamOptimizer(0.01), replicas_to_aggregate=1,
                                             total_num_replicas=1)
        return x, y_, xent, gs, opt.minimize(xent, global_step=gs)
    batches = _batches(4, 100, one_hot=True)
    a_l, a_g, a_v, _ = _run("b200", _build_mlp("adam"), 10, batches)
    s_l, s_g, s_v, s_sh = _run("b200", build_sync, 10, batches)
    assert s_g == a_g == list(range(1, 11))
    for x, y in zip(a_l, s_l):
        assert abs(x - y) <= 0.02 * abs(x) + 0.05, (a_l, s_l)
    for n in a_v:
        assert float(np.abs(a_v[n] - s_v[n]).mean()) < 2e-3, n


def test_eval_fetch_and_validation_batch():
    """Fetching only the loss runs a forward pass at any batch size (the MLP template's
    validation cross-entropy, templates/00_mnist_replica.py:267-271)."""
    import distributedtensorflow_b200 as tf
    from distributedtensorflow_b200.models import mnist_mlp
    tf.reset_default_graph()
    tf.set_random_seed(7)
    server = tf.train.Server.create_local_server()
    x = tf.placeholder(tf.float32, [None, 784])
    y_ = tf.placeholder(tf.float32, [None, 10])
    gs, xent, _ = mnist_mlp.build_model(x, y_)
    train_op = tf.train.AdamOptimizer(0.01).minimize(xent, global_step=gs)
    (bx, by), = _batches(1, 100, one_hot=True)
    (vx, vy), = _batches(1, 1000, one_hot=True)
    with tf.train.MonitoredTrainingSession(master=server.target, config=tf.ConfigProto(backend="b200")) as sess:
        v0 = float(sess.run(xent, feed_dict={x: vx, y_: vy}))
        for _ in range(30):
            sess.run(train_op, feed_dict={x: bx, y_: by})
        v1 = float(sess.run(xent, feed_dict={x: vx, y_: vy}))
        assert int(sess.run(gs)) == 30
    assert v1 < v0


@pytest.mark.parametrize("num_ps,depth,fused_push", [(1, 2, False), (2, 4, False), (2, 4, True)])
def test_mailbox_push_mode_owner_applies_every_gradient(num_ps, depth, fused_push):
    """push_mode='mailbox': the worker only stores gradients into its mailboxes; the PS shards'
    native apply daemons apply them. Every push must be applied exactly once (global_step and the
    daemons' counters), the loss must fall, and staleness is bounded by the mailbox depth."""
    import distributedtensorflow_b200 as tf
    tf.reset_default_graph()
    tf.set_random_seed(7)
    server = tf.train.Server.create_local_server(num_ps=num_ps)
    cluster = tf.train.ClusterSpec({"ps": ["localhost:%d" % (1000 + i) for i in range(num_ps)],
                                    "worker": ["localhost:2000"]})
    strategy = tf.train.FlatSlicedStrategy(num_ps) if num_ps > 1 else None
    with tf.device(tf.train.replica_device_setter(worker_device="/job:worker/task:0", cluster=cluster,
                                                  ps_strategy=strategy)):
        x_ph, y_ph, loss, gs, train_op = _build_mlp("adam")(tf)
    batches = _batches(4, 100, one_hot=True)
    cfg = tf.ConfigProto(backend="b200", device="cuda:0", push_mode="mailbox", mailbox_depth=depth,
                         fused_push=fused_push)
    steps = 40
    losses = []
    with tf.train.MonitoredTrainingSession(master=server.target, is_chief=True, config=cfg) as sess:
        raw = sess._tf_sess()
        for i in range(steps):
            bx, by = batches[i % len(batches)]
            _, l, g = sess.run([train_op, loss, gs], feed_dict={x_ph: bx, y_ph: by})
            losses.append(float(l))
            assert i + 1 - depth <= int(g) <= i + 1        # bounded staleness: at most `depth` pushes in flight
        runner = raw._runner
        assert runner.use_mailbox and runner.store.arena.mailbox_depth == depth
        assert int(sess.run(gs)) == steps                  # flushes my pushes first
        runner.replay(25)                                  # pipelined device steps go through the same admission
        runner.flush_pushes()
        assert runner.store.read_global_step() == steps + 25
        import time as _time
        for sh in range(num_ps):
            for _ in range(200):           # the daemon books retired gradients at its next poll
                st = raw.ps_client.call(sh, "mbox_stats")
                if st["applied"] == steps + 25:
                    break
                _time.sleep(0.005)
            assert st["running"] and st["applied"] == steps + 25
        w = runner.store.read_variable("hid_w")
        assert np.isfinite(w).all()
        np.testing.assert_array_equal(runner.store.read_shadow("hid_w"),
                                      _bf16_round(w))
        # checkpoint: the daemons are parked while their shards are read, then resume
        import tempfile, os
        from distributedtensorflow_b200.ckpt import BundleReader
        with tempfile.TemporaryDirectory() as d:
            path = tf.train.Saver().save(raw, os.path.join(d, "model.ckpt"), global_step=steps + 25)
            with BundleReader(path) as r:
                np.testing.assert_array_equal(r.get_tensor("hid_w"), w)
                assert int(r.get_tensor("global_step")) == steps + 25
        for i in range(3):
            sess.run([train_op, loss, gs], feed_dict={x_ph: batches[0][0], y_ph: batches[0][1]})
        assert int(sess.run(gs)) == steps + 28
    server.stop()
    assert np.mean(losses[-5:]) < 0.7 * np.mean(losses[:3])


def _check_mirror(runner, finals):
    """The worker-visible parameters (its mirror, written by the owners' drains) equal the owners'
    fp32 masters: exactly for the fp32 form, after bf16 rounding for the tensor-core form."""
    from distributedtensorflow_b200.ops import _lib as L
    assert runner.use_mirror and runner.my_shards == list(range(runner.store.num_shards))
    import time
    st = runner.store
    for _ in range(200):      # the header words trail the flag the flush waited for by a few stores
        ver = st.mirror_hdr_u32(L.MIRROR_OFF_VERSION, st.num_shards)
        con = st.mirror_hdr_u32(L.MIRROR_OFF_CONSUMED, st.num_shards)
        if list(ver) == [runner._pushes] * st.num_shards == list(con):
            break
        time.sleep(0.001)
    assert list(ver) == [runner._pushes] * st.num_shards and list(con) == list(ver)
    for n, master in finals.items():
        forms = st.read_mirror(n)
        assert forms, n
        if "f32" in forms:
            np.testing.assert_array_equal(forms["f32"], master, err_msg=n)
        if "bf16" in forms:
            np.testing.assert_array_equal(forms["bf16"], _bf16_round(master), err_msg=n)


@pytest.mark.parametrize("model", ["mlp", "cnn"])
def test_owner_pushes_mirror_with_in_graph_drain_matches_cpu_oracle(model):
    """Owner-applies push with worker mirrors and the drain as a node of the step graph
    (drain_at='end': every gradient is applied before the next step reads its mirror, so the run
    is the same sequence of applies as the CPU executor's)."""
    opts = {"push_mode": "mailbox", "drain_at": "end"}
    if model == "mlp":
        batches, build, steps = _batches(4, 100, one_hot=True), _build_mlp("adam"), 12
    else:
        batches, build, steps = _batches(3, 128, one_hot=False), _build_cnn, 8
    ref_l, ref_g, ref_v, _ = _run("cpu", build, steps, batches)
    got_l, got_g, got_v, _ = _run("b200", build, steps, batches, num_ps=2, strategy="flat", options=opts,
                                  probe=_check_mirror)
    # global_step as the step reports it comes from the mirror header: the count before this step's apply
    assert got_g == list(range(0, steps)), got_g
    for a, b in zip(got_l, ref_l):
        assert abs(a - b) <= 0.06 * abs(b) + (0.5 if model == "mlp" else 0.03), (got_l, ref_l)
    assert got_l[-1] < got_l[0]
    lr_steps = steps * (0.01 if model == "mlp" else 0.0005)
    for n in ref_v:
        diff = np.abs(got_v[n] - ref_v[n])
        scale = max(1e-3, float(np.abs(ref_v[n]).max()))
        assert float(diff.mean()) < 0.02 * scale + 0.05 * lr_steps, n


def test_owner_pushes_mirror_drain_at_start_counts_every_apply():
    """Default placement of the drain (start of the step, beside the forward pass): Hogwild reads,
    so only the accounting is exact - every push applied once, mirror == master after a flush."""
    batches = _batches(3, 128, one_hot=False)
    got_l, got_g, got_v, _ = _run("b200", _build_cnn, 10, batches, num_ps=2, strategy="flat",
                                  options={"push_mode": "mailbox"}, probe=_check_mirror)
    assert all(np.isfinite(got_l)) and got_l[-1] < got_l[0]


def _bf16_round(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).to(torch.bfloat16).float().numpy()


# ---------------------------------------------------------------------------------------------
# synchronous rounds on the mailboxes: PS-only tasks, backup workers (csrc/kernels/mailbox.cu)
# ---------------------------------------------------------------------------------------------
def _sync_run(backend, n_workers, replicas, steps, num_ps=1, slow_worker=None, options=None, stop_at=None):
    """``n_workers`` worker threads of one process against ``num_ps`` PS shards: the reference
    template's topology (templates/00_mnist_replica.py:80-83: one PS, two workers) on one GPU."""
    import threading
    import time
    import distributedtensorflow_b200 as tf
    from distributedtensorflow_b200.models import mnist_mlp
    tf.reset_default_graph()
    tf.set_random_seed(0)
    cluster = tf.train.ClusterSpec({"ps": ["h:%d" % i for i in range(num_ps)],
                                    "worker": ["w:%d" % i for i in range(n_workers)]})
    strategy = tf.train.FlatSlicedStrategy(num_ps) if num_ps > 1 else None
    with tf.device(tf.train.replica_device_setter(worker_device="/job:worker/task:0", cluster=cluster,
                                                  ps_strategy=strategy)):
        X = tf.placeholder(tf.float32, [None, 784])
        Y = tf.placeholder(tf.float32, [None, 10])
        gs, loss, _ = mnist_mlp.build_model(X, Y, 100)
        opt = tf.train.SyncReplicasOptimizer(tf.train.AdamOptimizer(0.001), replicas_to_aggregate=replicas,
                                             total_num_replicas=n_workers)
        train_op = opt.minimize(loss, global_step=gs)
    server = tf.train.Server.create_local_server(num_ps=num_ps)
    names = [v.name for v in tf.trainable_variables()]
    out, errs, seen = {}, [], {}
    t_start = time.time()
    # The worker threads of this test share ONE GPU. Pinned-memory and device allocations wait for every
    # kernel on the device, including another worker's step parked in its token wait, so every worker
    # finishes its first step (which allocates its state and cannot block: it needs no token) before
    # any worker starts a second one. Real deployments run one worker process per GPU.
    first_step_done = threading.Barrier(n_workers)
    opts = dict(cuda_graph=False, mailbox_timeout=25.0, sync_token_timeout=5.0)
    opts.update(options or {})

    def worker(i):
        try:
            from distributedtensorflow_b200.data import mnist
            imgs, labels = mnist.synthetic_arrays(4 * 100, seed=50 + i)
            x = imgs.astype(np.float32) / 255.0
            y = np.eye(10, dtype=np.float32)[labels]
            hooks = [opt.make_session_run_hook(i == 0, num_tokens=0)]
            if stop_at is not None:
                hooks.append(tf.train.StopAtStepHook(last_step=stop_at))
            cfg = tf.ConfigProto(backend=backend, device="cuda:0" if backend == "b200" else None,
                                 task_index=i, **(opts if backend == "b200" else {}))
            steps_done, last_g = 0, 0
            with tf.train.MonitoredTrainingSession(master=server.target, is_chief=(i == 0), hooks=hooks,
                                                   config=cfg, recovery_wait_secs=0.01) as sess:
                while (steps_done < steps) if stop_at is None else (not sess.should_stop()):
                    k = steps_done % 4
                    if slow_worker == i:
                        time.sleep(0.03)
                    _, l, g = sess.run([train_op, loss, gs], feed_dict={X: x[k * 100:(k + 1) * 100],
                                                                        Y: y[k * 100:(k + 1) * 100]})
                    steps_done += 1
                    last_g = int(g)
                    if steps_done == 1 and backend == "b200":
                        first_step_done.wait(60)
                raw = sess._tf_sess()
                if i == 0 and raw is not None:
                    if backend == "b200":
                        raw._runner.stream.synchronize()
                        time.sleep(0.05)          # the last round's apply runs on the PS daemon
                        out.update({n: raw._runner.store.read_variable(n) for n in names})
                        out["__global_step__"] = raw._runner.store.read_global_step()
                        out["__generic__"] = raw._runner.sync_generic
                    else:
                        out.update(raw.ps_client.pull_all(raw.layout, names))
            seen[i] = (steps_done, last_g)
        except Exception as e:  # pragma: no cover
            import traceback
            traceback.print_exc()
            errs.append(e)
    ts = [threading.Thread(target=worker, args=(i,), daemon=True) for i in range(n_workers)]
    [t.start() for t in ts]
    [t.join(180) for t in ts]
    server.stop()
    assert not errs, errs
    assert not any(t.is_alive() for t in ts), "a worker thread is stuck"
    return out, seen


def _isolated(fn_name, *args, timeout=150):
    """Run ``fn_name(*args)`` of this module in a process of its own. The synchronous-round tests put
    several worker threads on ONE GPU, where a worker's token wait is a kernel that spins until another
    worker's kernels have run: streams of one process normally run side by side, but nothing guarantees
    it (B200_PROFILING.md). If they ever serialise, the in-kernel watchdog traps - in the child, not in
    the pytest process whose CUDA context every later test needs. More hardware queues than streams
    (CUDA_DEVICE_MAX_CONNECTIONS) keep unrelated streams from sharing one."""
    import os
    import subprocess
    import sys
    here = os.path.abspath(__file__)
    code = ("import importlib.util, sys; sys.path.insert(0, %r); "
            "spec = importlib.util.spec_from_file_location('t_exec_gpu', %r); m = importlib.util.module_from_spec(spec); "
            "spec.loader.exec_module(m); getattr(m, %r)(*%r); print('ISOLATED-OK')"
            % (os.path.dirname(os.path.dirname(here)), here, fn_name, tuple(args)))
    env = dict(os.environ, CUDA_DEVICE_MAX_CONNECTIONS="32")
    p = subprocess.run([sys.executable, "-c", code], env=env, timeout=timeout, capture_output=True, text=True)
    assert p.returncode == 0 and "ISOLATED-OK" in p.stdout, (p.stdout[-2000:] + p.stderr[-3000:])


@pytest.mark.parametrize("num_ps", [1, 2])
def test_sync_on_mailboxes_ps_only_topology_matches_cpu_sync_replicas(num_ps):
    _isolated("_impl_sync_on_mailboxes_ps_only_topology", num_ps)


def test_sync_on_mailboxes_backup_worker_is_dropped_not_waited_for():
    _isolated("_impl_sync_on_mailboxes_backup_worker")


def _impl_sync_on_mailboxes_ps_only_topology(num_ps):
    """The MLP template's own topology - PS task(s) without a worker, two workers, all replicas
    aggregated - on the device path: tagged pushes, the PS daemon's mean+apply, epoch tokens."""
    steps = 8
    got, seen = _sync_run("b200", 2, 2, steps, num_ps=num_ps)
    ref, _ = _sync_run("cpu", 2, 2, steps, num_ps=num_ps)
    assert got.pop("__generic__") is True
    assert got.pop("__global_step__") == steps            # ONE global step per aggregated round
    for name, r in ref.items():
        diff = np.abs(got[name] - r)
        assert float(diff.mean()) < 2e-3 and float(diff.max()) < 0.03, (name, float(diff.mean()), float(diff.max()))


def _impl_sync_on_mailboxes_backup_worker():
    """replicas_to_aggregate=2 of 3 replicas, one of them slow: rounds complete with the two fast
    gradients, the straggler's arrive stale and are retired unapplied, and when the others have
    stopped the straggler's token wait ends like a closed token queue (clean stop)."""
    import time
    t0 = time.time()
    got, seen = _sync_run("b200", 3, 2, steps=None, slow_worker=2, stop_at=12,
                          options={"sync_token_timeout": 1.0})
    assert got.pop("__generic__") is True
    gstep = got.pop("__global_step__")
    assert 12 <= gstep <= 14, gstep
    fast = seen[0][0] + seen[1][0]
    assert seen[0][0] >= 10 and seen[1][0] >= 10           # the fast replicas took part in (almost) every round
    assert seen[2][0] < seen[0][0]                         # the straggler contributed to fewer
    assert all(np.isfinite(v).all() for v in got.values())
    assert time.time() - t0 < 120


def test_parallel_sharded_save_every_shard_writes_its_own_bytes_and_restores_into_one_shard(tmp_path):
    """Checkpoint of a flat-sliced model on 4 shards: every shard stores its own slices into the data
    files (PsShard.save_arena_ops), the chief writes only the index; each variable is one contiguous,
    checksummed tensor of the result. Restoring into a 1-shard cluster gives the same parameters,
    optimizer slots and step (save on N shards -> restore on M)."""
    import distributedtensorflow_b200 as tf
    from distributedtensorflow_b200.ckpt import BundleReader
    batches = _batches(2, 128, one_hot=False)

    def session(num_ps):
        tf.reset_default_graph()
        tf.set_random_seed(5)
        server = tf.train.Server.create_local_server(num_ps=num_ps)
        cluster = tf.train.ClusterSpec({"ps": ["localhost:%d" % (1000 + i) for i in range(num_ps)],
                                        "worker": ["localhost:2000"]})
        strategy = tf.train.FlatSlicedStrategy(num_ps) if num_ps > 1 else None
        with tf.device(tf.train.replica_device_setter(worker_device="/job:worker/task:0", cluster=cluster,
                                                      ps_strategy=strategy)):
            x, y, loss, gs, train_op = _build_cnn(tf)
        sess = tf.train.MonitoredTrainingSession(master=server.target, is_chief=True,
                                                 config=tf.ConfigProto(backend="b200", device="cuda:0"))
        return server, sess, (x, y, loss, gs, train_op)

    server, sess, (x, y, loss, gs, train_op) = session(4)
    with sess:
        for i in range(5):
            sess.run([train_op, loss, gs], feed_dict={x: batches[i % 2][0], y: batches[i % 2][1]})
        raw = sess._tf_sess()
        store = raw._runner.store
        names = [v.name for v in tf.trainable_variables()]
        want = {n: store.read_variable(n) for n in names}
        want_m = {n: store.read_slot(n, 0) for n in names}
        saver = tf.train.Saver()
        path = saver.save(raw, str(tmp_path / "model.ckpt"), global_step=5)
        with BundleReader(path) as r:
            assert r.num_shards == 4
            assert len(r.keys()) == 3 * len(names) + 3           # vars + m + v, global_step, beta powers
            for n in names:
                np.testing.assert_array_equal(r.get_tensor(n), want[n])             # crc verified on read
                np.testing.assert_array_equal(r.get_tensor(n + "/Adam"), want_m[n])
            assert int(r.get_tensor("global_step")) == 5
            # the big flat-sliced kernel is ONE tensor although four shards wrote it
            assert r.entries["dense/kernel"]["shape"] == (3136, 1024)
        assert saver.last_save_bytes == sum(v.nbytes for v in want.values()) * 3 + 8 + 8
    server.stop()
    server, sess, (x, y, loss, gs, train_op) = session(1)
    with sess:
        raw = sess._tf_sess()
        tf.train.Saver().restore(raw, path)
        store = raw._runner.store
        for n in names:
            np.testing.assert_array_equal(store.read_variable(n), want[n])
            np.testing.assert_array_equal(store.read_slot(n, 0), want_m[n])
        assert store.read_global_step() == 5
        _, l, g = sess.run([train_op, loss, gs], feed_dict={x: batches[0][0], y: batches[0][1]})
        assert int(g) == 6 and np.isfinite(l)
    server.stop()
