"""Two-GPU checks (one process per GPU, CUDA IPC + NVLink peer memory): asynchronous PS training
with exact global_step accounting, and the synchronous round kernel (reduce-scatter + mean + apply
+ all-gather) against the CPU SyncReplicasOptimizer path."""
import json
import os
import subprocess
import sys
import threading

import numpy as np
import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _torchrun(nproc, script, args, port, timeout=150):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc),
           "--master-addr", "127.0.0.1", "--master-port", str(port), script, *args]
    p = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=timeout)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    line = [l for l in p.stdout.splitlines() if l.startswith("{")][-1]
    return json.loads(line)


def test_async_two_gpus_counts_every_apply():
    out = _torchrun(2, "bench.py", ["--gpus", "2", "--steps", "20", "--warmup", "3"], 29611)
    assert out["n_gpus"] == 2 and out["value"] > 0
    assert out["config"]["push"].startswith("mailbox")     # the default with more than one worker
    # 2 workers x (3 warm-up + 20 timed + 20 back-to-back + 20 e2e) applies, one atomic each
    assert out["final_global_step"] == 2 * (3 + 20 + 20 + 20)


def test_async_two_gpus_remote_rmw_push():
    """The worker-side push (remote read-modify-write over NVLink) stays available and exact."""
    out = _torchrun(2, "bench.py", ["--gpus", "2", "--steps", "10", "--warmup", "3", "--push-mode", "rmw"], 29631)
    assert out["config"]["push"].startswith("rmw")
    assert out["final_global_step"] == 2 * (3 + 10 + 10 + 10)


def test_async_mailbox_mode_with_a_ps_only_task():
    """1 PS task (GPU 0) + 1 worker (GPU 1), push_mode=mailbox: the PS process has no training loop,
    its apply daemon alone turns the worker's posted gradients into updates."""
    out = _torchrun(2, "bench/ps_bench.py", ["--model", "mlp", "--topology", "ps1", "--steps", "30", "--warmup", "3",
                                             "--push-mode", "mailbox"], 29641)
    assert out["config"]["workers"] == 1 and out["config"]["ps_shards"] == 1 and out.get("push") == "mailbox"
    assert out["final_step"] == 33 and np.isfinite(out["final_loss"])


def test_async_one_ps_shard_plus_worker():
    out = _torchrun(2, "bench/ps_bench.py", ["--model", "mlp", "--topology", "ps1", "--steps", "30", "--warmup", "3"], 29621)
    assert out["config"]["workers"] == 1 and out["config"]["ps_shards"] == 1
    assert out["final_step"] == 33 and np.isfinite(out["final_loss"])


def _cpu_sync_oracle(steps, batch, hidden, lr, world):
    """The same sync-replicas run on the CPU transport (threads as workers, no initial tokens so
    every round holds exactly one gradient per replica)."""
    import distributedtensorflow_b200 as tf
    from distributedtensorflow_b200.models import mnist_mlp
    sys.path.insert(0, os.path.join(ROOT, "bench"))
    import ps_bench

    class A:
        pass
    args = A()
    args.batch, args.model = batch, "mlp"
    tf.reset_default_graph()
    tf.set_random_seed(0)
    cluster = tf.train.ClusterSpec({"ps": ["h:%d" % i for i in range(world)], "worker": ["w:%d" % i for i in range(world)]})
    with tf.device(tf.train.replica_device_setter(worker_device="/job:worker/task:0", cluster=cluster,
                                                  ps_strategy=tf.train.FlatSlicedStrategy(world))):
        X = tf.placeholder(tf.float32, [None, 784])
        Y = tf.placeholder(tf.float32, [None, 10])
        gs, loss, _ = mnist_mlp.build_model(X, Y, hidden)
        opt = tf.train.SyncReplicasOptimizer(tf.train.AdamOptimizer(lr), replicas_to_aggregate=world,
                                             total_num_replicas=world)
        train_op = opt.minimize(loss, global_step=gs)
    server = tf.train.Server.create_local_server(num_ps=world)
    names = [v.name for v in tf.trainable_variables()]
    out, errs = {}, []

    def worker(i):
        try:
            batches = ps_bench.make_batches(args, i)
            hook = opt.make_session_run_hook(i == 0, num_tokens=0)
            with tf.train.MonitoredTrainingSession(master=server.target, is_chief=(i == 0), hooks=[hook],
                                                   config=tf.ConfigProto(backend="cpu"), recovery_wait_secs=0.01) as sess:
                for k in range(steps):
                    bx, by = batches[k % len(batches)]
                    sess.run(train_op, feed_dict={X: bx.astype(np.float32) / 255.0, Y: by})
                if i == 0:
                    raw = sess._tf_sess()
                    out.update(raw.ps_client.pull_all(raw.layout, names))
        except Exception as e:  # pragma: no cover
            errs.append(e)
    ts = [threading.Thread(target=worker, args=(i,), daemon=True) for i in range(world)]
    [t.start() for t in ts]
    [t.join(120) for t in ts]
    assert not errs, errs
    return out


def test_sync_round_kernel_matches_cpu_sync_replicas():
    steps, warm = 10, 2
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    res = _torchrun(2, "bench/ps_bench.py",
                    ["--model", "mlp", "--mode", "sync", "--placement", "flat", "--steps", str(steps), "--warmup",
                     str(warm), "--check", "--lr", "0.001", "--hidden", "100", "--batch", "100"], 29631)
    assert res["final_step"] == steps + warm          # ONE global step per aggregated round
    got0 = np.load(os.path.join(ROOT, "gpurun_out", "sync_params_rank0.npz"))
    got1 = np.load(os.path.join(ROOT, "gpurun_out", "sync_params_rank1.npz"))
    ref = _cpu_sync_oracle(steps + warm, 100, 100, 0.001, 2)
    for name, r in ref.items():
        a, b = got0[name.replace("/", "__")], got1[name.replace("/", "__")]
        assert np.array_equal(a, b), name             # both ranks see the same PS state
        diff = np.abs(a - r)
        assert float(diff.mean()) < 2e-3 and float(diff.max()) < 0.03, (name, float(diff.mean()), float(diff.max()))


def test_sync_checkpoint_sharded_save(tmp_path):
    res = _torchrun(2, "bench/ps_bench.py",
                    ["--model", "mlp", "--mode", "sync", "--placement", "round_robin", "--steps", "5", "--warmup", "2",
                     "--ckpt-dir", str(tmp_path)], 29641)
    assert res["ckpt_shards"] == 2 and res["ckpt_roundtrip_ok"] is True
    assert res["ckpt_tensors"] == 4 * 3 + 1 + 2        # vars + Adam slots, global_step, beta powers


def _gpus():
    import torch
    return torch.cuda.device_count()


def test_async_mirror_lockstep_matches_cpu_oracle(tmp_path):
    """Two workers + two shards on two GPUs, owner-applies push with worker mirrors and in-graph
    drains, run in lock step (workers take turns, every push flushed): the sequence of applies is then
    deterministic and must match a single-process CPU run that replays it; afterwards every worker's
    mirror equals the owners' masters (fp32 exactly, bf16 after rounding) and its header carries the
    owners' counts."""
    steps = 6
    pattern = str(tmp_path / "params_rank%d.npz")
    res = _torchrun(2, "bench/ps_bench.py",
                    ["--model", "mlp", "--placement", "flat", "--steps", str(steps), "--warmup", "0", "--lockstep",
                     "--lr", "0.001", "--hidden", "100", "--batch", "100", "--dump-params", pattern], 29651)
    assert res["mirror"] is True and res["graph_drain_shards"] == [0]
    assert res["final_step"] == 2 * steps
    got = [np.load(pattern % r) for r in range(2)]
    import torch
    for r in range(2):
        assert list(got[r]["mirror_consumed"]) == [steps, steps]            # each owner applied all of MY pushes
        assert list(got[r]["mirror_version"]) == [2 * steps, 2 * steps]     # ... and everybody's
    names = [k for k in got[0].files if not k.startswith("mirror_")]
    for n in names:
        np.testing.assert_array_equal(got[0][n], got[1][n])                 # both ranks read the same masters
        for r in range(2):
            if "mirror_f32__" + n in got[r].files:
                np.testing.assert_array_equal(got[r]["mirror_f32__" + n], got[r][n])
            if "mirror_bf16__" + n in got[r].files:
                want = torch.from_numpy(got[r][n]).to(torch.bfloat16).float().numpy()
                np.testing.assert_array_equal(got[r]["mirror_bf16__" + n], want)
    # CPU oracle: one process applies the same gradients in the same order (w0 step i, w1 step i, ...)
    import distributedtensorflow_b200 as tf
    from distributedtensorflow_b200.models import mnist_mlp
    sys.path.insert(0, os.path.join(ROOT, "bench"))
    import ps_bench

    class A:
        pass
    args = A()
    args.batch, args.model = 100, "mlp"
    tf.reset_default_graph()
    tf.set_random_seed(0)
    X = tf.placeholder(tf.float32, [None, 784])
    Y = tf.placeholder(tf.float32, [None, 10])
    gs, loss, _ = mnist_mlp.build_model(X, Y, 100)
    train_op = tf.train.AdamOptimizer(0.001).minimize(loss, global_step=gs)
    server = tf.train.Server.create_local_server()
    batches = [ps_bench.make_batches(args, w) for w in range(2)]
    with tf.train.MonitoredTrainingSession(master=server.target, config=tf.ConfigProto(backend="cpu")) as sess:
        for i in range(steps):
            for w in range(2):
                bx, by = batches[w][i % len(batches[w])]
                sess.run(train_op, feed_dict={X: bx.astype(np.float32) / 255.0, Y: by})
        raw = sess._tf_sess()
        ref = raw.ps_client.pull_all(raw.layout, [v.name for v in tf.trainable_variables()])
    server.stop()
    for name, r in ref.items():
        diff = np.abs(got[0][name.replace("/", "__")] - r)
        assert float(diff.mean()) < 2e-3 and float(diff.max()) < 0.03, (name, float(diff.mean()), float(diff.max()))


def test_sync_on_mailboxes_ps_only_task_two_gpus():
    """The MLP template's sync mode with a PS task that has no worker (GPU 0) and a worker (GPU 1):
    the round runs on the mailboxes - tagged push, the PS daemon's mean + apply, epoch token."""
    res = _torchrun(2, "bench/ps_bench.py",
                    ["--model", "mlp", "--mode", "sync", "--topology", "ps1", "--steps", "10", "--warmup", "2",
                     "--opt", "sync_path=mailbox"], 29661)
    assert res.get("sync_path") == "mailbox" and res["config"]["workers"] == 1
    assert res["final_step"] == 12 and np.isfinite(res["final_loss"])


def test_sync_backup_workers_four_gpus():
    """1 PS task + 3 workers on 4 GPUs, replicas_to_aggregate = 2: every round is the mean of the
    fresh gradients that arrived (>= 2), stale ones are dropped, nobody waits for a third."""
    if _gpus() < 4:
        pytest.skip("needs 4 GPUs")
    res = _torchrun(4, "bench/ps_bench.py",
                    ["--model", "mlp", "--mode", "sync", "--topology", "ps1", "--steps", "20", "--warmup", "2",
                     "--replicas", "2", "--opt", "sync_token_timeout=2"], 29671)
    assert res.get("sync_path") == "mailbox" and res["config"]["workers"] == 3
    # every round applies the mean of the 2 or 3 fresh gradients that arrived; training ends at the target
    assert 22 <= res["final_step"] <= 24 and np.isfinite(res["final_loss"])


def test_async_eight_gpus_counts_every_apply():
    if _gpus() < 8:
        pytest.skip("needs 8 GPUs")
    out = _torchrun(8, "bench.py", ["--gpus", "8", "--steps", "20", "--warmup", "3"], 29681, timeout=240)
    assert out["n_gpus"] == 8 and out["config"]["push"].startswith("mailbox")
    assert out["final_global_step"] == 8 * (3 + 20 + 20 + 20)
    assert out["final_loss"] < 2.0


def test_device_path_survives_a_worker_killed_mid_training():
    """Three OS processes per the reference's CLI on the device data plane - PS on its own GPU, two
    workers on theirs - and worker 1 is killed (os._exit) at its 30th step, mailboxes half-used and its
    mirror still registered with the PS. The chief must reach max_steps, the PS must keep applying and
    shut down cleanly (reference: a dead worker only stops contributing,
    run_mnist_distributed.py:143-150)."""
    if _gpus() < 3:
        pytest.skip("needs 3 GPUs (a PS task and two workers, one process per GPU)")
    cmd = [sys.executable, os.path.join(ROOT, "bench", "gpu_cli_baseline.py"), "--steps", "60000", "--workers", "2",
           "--ps", "1", "--gpus", "3", "--fault", "1:kill@30", "--timeout", "150"]
    p = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=200)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-3000:]
    res = json.loads([l for l in p.stdout.splitlines() if l.startswith("{")][-1])
    assert res["global_steps"] >= 60000               # (long enough for the second worker to join: it starts ~1 s late)
    assert res["workers"][1]["exit_code"] == 17
    assert 25 <= res["global_steps"] - res["workers"][0]["local_steps"] <= 40     # what the dead worker contributed
