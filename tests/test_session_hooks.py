"""Session layers, hooks, Supervisor and fault recovery on the in-process fake cluster
(SURVEY.md §4 items 2, 3, 7)."""
import os
import threading
import time

import numpy as np
import pytest

import distributedtensorflow_b200 as tf
from distributedtensorflow_b200.models import mnist_mlp
from distributedtensorflow_b200.parallel.transport import LocalCluster

CPU = dict(backend="cpu")


def _graph(num_ps=1, lr=0.01):
    tf.reset_default_graph()
    tf.set_random_seed(3)
    cluster = tf.train.ClusterSpec({"ps": ["h:%d" % i for i in range(num_ps)], "worker": ["w:0", "w:1"]})
    with tf.device(tf.train.replica_device_setter(worker_device="/job:worker/task:0", cluster=cluster)):
        x = tf.placeholder(tf.float32, [None, 784])
        y_ = tf.placeholder(tf.float32, [None, 10])
        gs, xent, y = mnist_mlp.build_model(x, y_, hidden_units=8)
        train_op = tf.train.AdamOptimizer(lr).minimize(xent, global_step=gs)
    rng = np.random.default_rng(0)
    feed = {x: rng.random((16, 784), dtype=np.float32), y_: np.eye(10, dtype=np.float32)[rng.integers(0, 10, 16)]}
    return x, y_, gs, xent, y, train_op, feed


def test_stop_at_step_hook_and_fetch_structures():
    x, y_, gs, xent, y, train_op, feed = _graph()
    server = tf.train.Server.create_local_server()
    calls = []

    class Spy(tf.train.SessionRunHook):
        def begin(self): calls.append("begin")
        def after_create_session(self, session, coord): calls.append("created")
        def before_run(self, ctx):
            calls.append("before")
            return tf.train.SessionRunArgs({"loss": xent})
        def after_run(self, ctx, values):
            assert set(values.results) == {"loss"}
            calls.append("after")
        def end(self, session): calls.append("end")

    with tf.train.MonitoredTrainingSession(master=server.target, hooks=[tf.train.StopAtStepHook(last_step=3), Spy()],
                                           config=tf.ConfigProto(**CPU)) as sess:
        n = 0
        while not sess.should_stop():
            out = sess.run({"t": train_op, "pair": [xent, gs], "probs": y}, feed_dict=feed)
            n += 1
            assert out["t"] is None and out["probs"].shape == (16, 10)
            assert int(out["pair"][1]) == n
        assert n == 3
        with pytest.raises(RuntimeError):
            sess.run(train_op, feed_dict=feed)         # run after should_stop is an error
    assert calls[:2] == ["begin", "created"] and calls[-1] == "end" and calls.count("before") == 3
    with pytest.raises(ValueError):
        tf.train.StopAtStepHook()
    with pytest.raises(ValueError):
        tf.train.StopAtStepHook(num_steps=1, last_step=2)


def test_num_steps_counts_from_the_step_at_session_creation():
    x, y_, gs, xent, y, train_op, feed = _graph()
    server = tf.train.Server.create_local_server()
    cfg = tf.ConfigProto(**CPU)
    with tf.train.MonitoredTrainingSession(master=server.target, hooks=[tf.train.StopAtStepHook(last_step=4)], config=cfg) as s:
        while not s.should_stop():
            s.run(train_op, feed_dict=feed)
    # a non-chief joins later: variables are already initialised, num_steps=2 -> stops at 6
    with tf.train.MonitoredTrainingSession(master=server.target, is_chief=False, config=cfg, recovery_wait_secs=0.01,
                                           hooks=[tf.train.StopAtStepHook(num_steps=2)]) as s:
        while not s.should_stop():
            _, g = s.run([train_op, gs], feed_dict=feed)
        assert int(g) == 6


def test_chief_inits_others_wait():
    x, y_, gs, xent, y, train_op, feed = _graph(num_ps=2)
    server = tf.train.Server.create_local_server(num_ps=2)
    cfg = tf.ConfigProto(**CPU)
    order, box = [], {}

    def non_chief():
        s = tf.train.MonitoredTrainingSession(master=server.target, is_chief=False, config=cfg, recovery_wait_secs=0.02)
        order.append("worker_ready")
        box["s"] = s

    t = threading.Thread(target=non_chief, daemon=True)
    t.start()
    time.sleep(0.3)
    assert order == [] and t.is_alive()                 # blocked: nothing is initialised yet
    with pytest.raises(TimeoutError):
        tf.train.WorkerSessionCreator(master=server.target, config=cfg, max_wait_secs=0.1,
                                      recovery_wait_secs=0.02).create_session()
    order.append("chief_start")
    chief = tf.train.MonitoredTrainingSession(master=server.target, is_chief=True, config=cfg)
    t.join(10)
    assert order == ["chief_start", "worker_ready"]
    box["s"].run(train_op, feed_dict=feed)
    assert int(chief.run(gs)) == 1
    box["s"].close()
    chief.close()


def test_uninitialized_report_and_plain_session():
    x, y_, gs, xent, y, train_op, feed = _graph()
    server = tf.train.Server.create_local_server()
    with tf.Session(server.target, config=tf.ConfigProto(**CPU)) as sess:
        missing = list(sess.run(tf.report_uninitialized_variables()))
        assert "hid_w" in missing and "global_step" in missing
        with pytest.raises(ValueError):                 # placeholders must be fed
            sess.run(tf.global_variables_initializer())
            sess.run(xent)
        assert len(sess.run(tf.report_uninitialized_variables())) == 0
        l0 = float(sess.run(xent, feed_dict=feed))
        for _ in range(5):
            sess.run(train_op, feed_dict=feed)
        assert float(sess.run(xent, feed_dict=feed)) < l0
    with pytest.raises(RuntimeError):
        sess.run(gs)                                    # closed


def test_recoverable_session_survives_ps_preemption(tmp_path):
    """Kill the PS mid-run: the step raises AbortedError inside, the monitored session recreates
    itself, the chief restores the last checkpoint and training continues (TF _RecoverableSession)."""
    x, y_, gs, xent, y, train_op, feed = _graph()
    server = tf.train.Server.create_local_server()
    lc = LocalCluster.lookup(server.target)
    ckdir = str(tmp_path / "ck")
    with tf.train.MonitoredTrainingSession(master=server.target, checkpoint_dir=ckdir, save_checkpoint_steps=2,
                                           save_summaries_steps=None, log_step_count_steps=None,
                                           config=tf.ConfigProto(**CPU)) as sess:
        for _ in range(5):
            sess.run(train_op, feed_dict=feed)
        assert tf.train.latest_checkpoint(ckdir).endswith("model.ckpt-4")
        lc.kill_ps(0)                                   # state is gone
        threading.Timer(0.3, lambda: lc.restart_ps(0)).start()
        _, g = sess.run([train_op, gs], feed_dict=feed)  # blocks through the outage, then retries
        assert sess.recoveries >= 1
        assert int(g) == 5                              # restored step 4, +1
    assert tf.train.latest_checkpoint(ckdir).endswith("model.ckpt-5")


def test_step_counter_logging_nan_and_summary_hooks(tmp_path):
    x, y_, gs, xent, y, train_op, feed = _graph()
    server = tf.train.Server.create_local_server()
    counter = tf.train.StepCounterHook(every_n_steps=2, output_dir=str(tmp_path / "sc"))
    logging_hook = tf.train.LoggingTensorHook({"loss": xent}, every_n_iter=2)
    summary = tf.train.SummarySaverHook(save_steps=1, output_dir=str(tmp_path / "sum"), summary_op={"xent": xent})
    with tf.train.MonitoredTrainingSession(master=server.target, hooks=[counter, logging_hook, summary],
                                           config=tf.ConfigProto(**CPU)) as sess:
        for _ in range(6):
            sess.run(train_op, feed_dict=feed)
    assert counter.last_rate is not None and counter.last_rate > 0
    assert len(logging_hook.logged) == 3
    from distributedtensorflow_b200 import logger
    df = logger.read_tb(str(tmp_path / "sum"))
    assert "xent" in df.columns and df["xent"].notna().sum() >= 5
    assert "global_step/sec" in logger.read_tb(str(tmp_path / "sc")).columns
    # NaN detection
    x, y_, gs, xent, y, train_op, feed = _graph(lr=1e30)
    server = tf.train.Server.create_local_server()
    with pytest.raises(tf.train.NanLossDuringTrainingError):
        with tf.train.MonitoredTrainingSession(master=server.target, hooks=[tf.train.NanTensorHook(xent)],
                                               config=tf.ConfigProto(**CPU)) as sess:
            for _ in range(20):
                sess.run(train_op, feed_dict=feed)


def test_supervisor_chief_services_and_worker_wait(tmp_path):
    x, y_, gs, xent, y, train_op, feed = _graph()
    server = tf.train.Server.create_local_server()
    cfg = tf.ConfigProto(allow_soft_placement=True, device_filters=["/job:ps", "/job:worker/task:0"], **CPU)
    sv = tf.train.Supervisor(is_chief=True, logdir=str(tmp_path / "sv"), init_op=tf.global_variables_initializer(),
                             recovery_wait_secs=1, global_step=gs, save_model_secs=0.2, save_summaries_secs=0.2)
    sess = sv.prepare_or_wait_for_session(server.target, config=cfg)
    sv2 = tf.train.Supervisor(is_chief=False, logdir=str(tmp_path / "ignored"), recovery_wait_secs=0.05, global_step=gs)
    sess2 = sv2.prepare_or_wait_for_session(server.target, config=cfg)
    with pytest.raises(RuntimeError):
        sv2.start_standard_services(sess2)
    for _ in range(4):
        sess.run(train_op, feed_dict=feed)
        sess2.run(train_op, feed_dict=feed)
        time.sleep(0.1)
    time.sleep(0.3)
    assert not sv.should_stop()
    sv.stop()
    sv2.stop()
    assert sv.should_stop()
    ck = tf.train.latest_checkpoint(str(tmp_path / "sv"))
    assert ck is not None and os.path.basename(ck).startswith("model.ckpt-")
    assert int(sess.run(gs)) == 8
    sess.close()
    sess2.close()


def test_coordinator_propagates_thread_exception():
    coord = tf.train.Coordinator()

    def bad():
        with coord.stop_on_exception():
            raise ValueError("boom")
    t = threading.Thread(target=bad, daemon=True)
    coord.register_thread(t)
    t.start()
    with pytest.raises(ValueError, match="boom"):
        coord.join()
    assert coord.should_stop()


def test_tracing_helpers_are_noops_without_nvtx():
    from distributedtensorflow_b200.utils import tracing
    assert not tracing.nvtx_enabled() or True
    tracing.enable_nvtx(False)
    with tracing.nvtx_range("dtf/test"):
        pass
    assert tracing.nvtx_range("a") is tracing.nvtx_range("b")      # shared null context when off


def test_data_plane_stats_hook_logs_rates_and_roofline_fractions():
    import time as _time
    import distributedtensorflow_b200 as tf
    from distributedtensorflow_b200 import logger

    class FakeRunner:
        def traffic_per_step(self):
            return dict(pull_remote_bytes=5_000_000, pull_total_bytes=6_000_000,
                        push_remote_bytes=11_000_000, push_total_bytes=13_000_000, push_mode="mailbox")

    class FakeSess:
        _runner = FakeRunner()

        def _tf_sess(self):
            return self

    class Ctx:
        session = FakeSess()

    hook = tf.train.DataPlaneStatsHook(every_n_steps=5)
    hook.begin()
    for _ in range(6):
        _time.sleep(0.002)
        hook.after_run(Ctx(), None)
    st = hook.last
    assert st["push_mode"] == "mailbox" and 50 < st["local_step/sec"] < 1000
    assert abs(st["push_GB/s"] - 11e6 * st["local_step/sec"] / 1e9) < 1e-9
    assert abs(st["pull_pct_nvlink"] - 100 * st["pull_GB/s"] / 770.0) < 1e-9
    assert logger.getkvs()["push_GB/s"] == st["push_GB/s"]
