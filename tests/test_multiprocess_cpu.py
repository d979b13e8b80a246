"""BASELINE config 1 as an integration test: 1 PS + 2 workers as three OS processes on localhost
over the TCP transport, driven through the same CLI as the reference."""
import json
import os
import re
import signal
import socket
import subprocess
import sys
import time

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_ports(n):
    socks, ports = [], []
    for _ in range(n):
        s = socket.socket()
        s.bind(("127.0.0.1", 0))
        socks.append(s)
        ports.append(s.getsockname()[1])
    for s in socks:
        s.close()
    return ports


def _cluster_file(tmp_path, n_ps, n_workers):
    ports = _free_ports(n_ps + n_workers)
    cfg = {"ps": {"ps:%d" % i: "127.0.0.1:%d" % ports[i] for i in range(n_ps)},
           "workers": {"worker:%d" % i: "127.0.0.1:%d" % ports[n_ps + i] for i in range(n_workers)}}
    path = tmp_path / "config.json"
    path.write_text(json.dumps(cfg))
    return str(path)


def _launch(job, idx, cfg_path, tmp_path, extra=()):
    env = dict(os.environ, PYTHONPATH=ROOT, PYTHONUNBUFFERED="1", OMP_NUM_THREADS="2")
    cmd = [sys.executable, os.path.join(ROOT, "run_mnist_distributed.py"), "--job_name", job, "--task_index",
           str(idx), "--config_file", cfg_path, "--backend", "cpu", "--data_dir", str(tmp_path / "data"),
           "--log_root", str(tmp_path / "logs"), *extra]
    return subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env, cwd=ROOT)


def _finish(p, timeout):
    try:
        out, _ = p.communicate(timeout=timeout)
    except subprocess.TimeoutExpired:
        p.kill()
        out, _ = p.communicate()
        out += "\n<<killed after timeout>>"
    return out


@pytest.mark.slow
def test_one_ps_two_workers_async(tmp_path):
    from distributedtensorflow_b200.data import mnist
    mnist.synthetic_mnist(str(tmp_path / "data"), num_train=2048, num_test=256)
    cfg = _cluster_file(tmp_path, 1, 2)
    steps = 40
    ps = _launch("ps", 0, cfg, tmp_path)
    w1 = _launch("worker", 1, cfg, tmp_path, ["--max_steps", str(steps)])     # non-chief first: must wait
    time.sleep(1.0)
    w0 = _launch("worker", 0, cfg, tmp_path, ["--max_steps", str(steps), "--shutdown_ps_when_done", "true"])
    out0 = _finish(w0, 240)
    out1 = _finish(w1, 180)
    outp = _finish(ps, 120)
    assert w0.returncode == 0, out0
    assert w1.returncode == 0, out1
    assert ps.returncode == 0 and "Close Parameter Server" in outp, outp   # clean PS shutdown
    pairs = [(float(a), int(b)) for a, b in re.findall(r"loss = ([0-9.]+) \(global step: (\d+)\)", out0)]
    steps_seen = [s for _, s in pairs]
    assert steps_seen == sorted(steps_seen) and len(set(steps_seen)) == len(steps_seen)   # monotone
    assert steps <= steps_seen[-1] <= steps + 1           # StopAtStepHook overshoot <= W - 1
    assert pairs[-1][0] < pairs[0][0]                     # loss decreases
    m0 = re.search(r"Worker 0: (\d+) local steps", out0)
    m1 = re.search(r"Worker 1: (\d+) local steps", out1)
    total_local = int(m0.group(1)) + int(m1.group(1))
    # global_step counts applies from ALL workers: sum of local steps == final global step
    final = max(steps_seen[-1], int(re.search(r"\(global step (\d+)\)", out1).group(1)) if "global step" in out1 else 0)
    assert steps <= total_local <= steps + 2, (total_local, out0[-400:], out1[-400:])
    assert int(m1.group(1)) > 0, "the second worker contributed steps"
    assert os.listdir(tmp_path / "logs"), "chief wrote TensorBoard events"


@pytest.mark.slow
def test_ps_survives_worker_exit_and_worker_reattaches(tmp_path):
    from distributedtensorflow_b200.data import mnist
    mnist.synthetic_mnist(str(tmp_path / "data"), num_train=1024, num_test=128)
    cfg = _cluster_file(tmp_path, 1, 1)
    ps = _launch("ps", 0, cfg, tmp_path)
    try:
        w = _launch("worker", 0, cfg, tmp_path, ["--max_steps", "6"])
        out_a = _finish(w, 180)
        assert w.returncode == 0 and "global step: 6" in out_a, out_a
        # a new chief re-attaches: the PS still holds the variables but this chief re-runs init_op
        # (no checkpoint_dir), exactly like the reference; training proceeds again to the target
        w = _launch("worker", 0, cfg, tmp_path, ["--max_steps", "4", "--shutdown_ps_when_done", "true"])
        out_b = _finish(w, 180)
        assert w.returncode == 0 and "global step: 4" in out_b, out_b
        outp = _finish(ps, 120)
        assert ps.returncode == 0, outp
    finally:
        if ps.poll() is None:
            ps.kill()


def test_ps_clean_shutdown_on_sigint(tmp_path):
    """The reference's README TODO: Ctrl-C must end a parameter server cleanly."""
    cfg = _cluster_file(tmp_path, 1, 1)
    ps = _launch("ps", 0, cfg, tmp_path)
    deadline = time.time() + 60
    line = ""
    while time.time() < deadline:
        line = ps.stdout.readline()
        if "Started Parameter Server" in line or not line:
            break
    assert "Started Parameter Server" in line
    time.sleep(0.3)
    ps.send_signal(signal.SIGINT)
    out = _finish(ps, 120)
    assert ps.returncode == 0 and "Close Parameter Server" in out, out


def test_collective_ps_comparator_matches_local_adam_on_gloo():
    """parallel/nccl_baseline.py (the NCCL/gloo comparator of the PS API): two gloo processes, sync
    round == Adam on the mean gradient, async-emulated push == W sequential Adam applies."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", "29877", "-m",
           "distributedtensorflow_b200.parallel.nccl_baseline"]
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    p = subprocess.run(cmd, cwd=root, capture_output=True, text=True, timeout=240, env=env)
    assert p.returncode == 0 and "SELFCHECK_OK" in p.stdout, p.stdout[-2000:] + p.stderr[-2000:]


def _launch_template(script, args, tmp_path):
    env = dict(os.environ, PYTHONPATH=ROOT, PYTHONUNBUFFERED="1", OMP_NUM_THREADS="2", CUDA_VISIBLE_DEVICES="")
    cmd = [sys.executable, os.path.join(ROOT, "templates", script), *args]
    return subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env, cwd=str(tmp_path))


@pytest.mark.slow
@pytest.mark.parametrize("sync", [False, True])
def test_mlp_template_one_ps_two_workers(tmp_path, sync):
    """templates/00_mnist_replica.py (the reference's MLP template: Supervisor, feed-dict loop,
    optional SyncReplicasOptimizer) as 1 PS + 2 workers on the CPU backend."""
    from distributedtensorflow_b200.data import mnist
    data = tmp_path / "data"
    mnist.synthetic_mnist(str(data), num_train=8192, num_test=256)     # read_data_sets keeps 5000 for validation
    ports = _free_ports(3)
    common = ["--ps_hosts", "127.0.0.1:%d" % ports[0],
              "--worker_hosts", "127.0.0.1:%d,127.0.0.1:%d" % (ports[1], ports[2]),
              "--data_dir", str(data), "--backend", "cpu", "--train_steps", "30", "--batch_size", "50"]
    if sync:
        common.append("--sync_replicas")
    ps = _launch_template("00_mnist_replica.py", common + ["--job_name", "ps", "--task_index", "0"], tmp_path)
    w1 = _launch_template("00_mnist_replica.py", common + ["--job_name", "worker", "--task_index", "1"], tmp_path)
    time.sleep(1.0)
    w0 = _launch_template("00_mnist_replica.py", common + ["--job_name", "worker", "--task_index", "0",
                                                           "--shutdown_ps"], tmp_path)
    out0 = _finish(w0, 240)
    out1 = _finish(w1, 180)
    outp = _finish(ps, 120)
    assert w0.returncode == 0, out0[-3000:]
    assert w1.returncode == 0, out1[-3000:]
    assert ps.returncode == 0, outp[-2000:]
    steps = [int(s) for s in re.findall(r"training step \d+ done \(global step: (\d+)\)", out0)]
    assert steps and steps == sorted(steps) and steps[-1] >= 30
    m = re.search(r"After \d+ training step\(s\), validation cross entropy = ([0-9.eE+-]+)", out0)
    assert m is not None and float(m.group(1)) > 0.0, out0[-1500:]
    assert "Training elapsed time" in out0


@pytest.mark.slow
def test_async_template_checkpoints_and_restores_after_a_full_restart(tmp_path):
    """templates/00_between_graph_replication_async_mnist.py: Adagrad, MonitoredTrainingSession with
    checkpoint_dir. Phase 1 trains to step 30 and leaves a TF1-format checkpoint; then EVERY task is
    gone (the PS holds no state any more). Phase 2 starts a fresh PS + chief, which must restore the
    checkpoint and continue from >= 30 to 45. The PS never exits by itself, like the
    reference's (README TODO): SIGINT ends it cleanly."""
    from distributedtensorflow_b200.ckpt import BundleReader
    ckpt = tmp_path / "train_logs"
    script = "00_between_graph_replication_async_mnist.py"

    def run_phase(last_step, n_workers):
        ports = _free_ports(1 + n_workers)
        common = ["--ps_hosts", "127.0.0.1:%d" % ports[0],
                  "--worker_hosts", ",".join("127.0.0.1:%d" % p for p in ports[1:]),
                  "--backend", "cpu", "--last_step", str(last_step), "--checkpoint_dir", str(ckpt)]
        ps = _launch_template(script, common + ["--job_name", "ps", "--task_index", "0"], tmp_path)
        ws = [_launch_template(script, common + ["--job_name", "worker", "--task_index", str(i)], tmp_path)
              for i in reversed(range(n_workers))]
        outs = [_finish(w, 240) for w in ws]
        for w, o in zip(ws, outs):
            assert w.returncode == 0, o[-3000:]
        ps.send_signal(signal.SIGINT)
        outp = _finish(ps, 120)
        assert ps.returncode == 0, outp[-2000:]
        return "\n".join(outs)

    def latest_step():
        state = (ckpt / "checkpoint").read_text()
        m = re.search(r'model_checkpoint_path: "([^"]+)"', state)
        path = m.group(1)
        path = path if os.path.isabs(path) else str(ckpt / path)
        with BundleReader(path) as r:
            keys = set(r.keys())
            assert {"hid_w", "hid_b", "sm_w", "sm_b", "global_step", "hid_w/Adagrad"} <= keys, keys
            return int(r.get_tensor("global_step")), r.get_tensor("hid_w").copy()

    out1 = run_phase(30, 2)
    assert "Restoring parameters from" not in out1
    step1, w1 = latest_step()
    assert 30 <= step1 <= 31
    out2 = run_phase(45, 1)
    assert "Restoring parameters from" in out2 and "model.ckpt-%d" % step1 in out2
    step2, w2 = latest_step()
    assert 45 <= step2 <= 46 and step2 > step1          # continued from the checkpoint, did not restart at 0
    assert not (w1 == w2).all()


@pytest.mark.slow
def test_fault_injection_worker_killed_mid_run_async_continues(tmp_path):
    """kill -9 a (non-chief) worker mid-run. Async PS training is elastic on the
    worker side: the surviving worker keeps stepping and reaches the target alone; the PS survives."""
    from distributedtensorflow_b200.data import mnist
    mnist.synthetic_mnist(str(tmp_path / "data"), num_train=2048, num_test=256)
    cfg = _cluster_file(tmp_path, 1, 2)
    ps = _launch("ps", 0, cfg, tmp_path)
    try:
        w1 = _launch("worker", 1, cfg, tmp_path, ["--max_steps", "40", "--fault_inject", "kill@5"])
        time.sleep(1.0)
        w0 = _launch("worker", 0, cfg, tmp_path, ["--max_steps", "40", "--shutdown_ps_when_done", "true"])
        out1 = _finish(w1, 120)
        out0 = _finish(w0, 240)
        outp = _finish(ps, 120)
        assert w1.returncode not in (0, None), out1[-1500:]            # died by SIGKILL (negative) / non-zero
        assert w0.returncode == 0, out0[-3000:]
        steps_seen = [int(s) for s in re.findall(r"\(global step: (\d+)\)", out0)]
        assert steps_seen and steps_seen[-1] >= 40
        assert ps.returncode == 0, outp[-2000:]
    finally:
        for p in (ps,):
            if p.poll() is None:
                p.kill()


@pytest.mark.slow
def test_fault_injection_ps_killed_then_restarted_chief_restores_checkpoint(tmp_path):
    """kill -9 the PS mid-run, start a new PS on the same address. The worker's
    mon_sess.run surfaces AbortedError internally, the recoverable session rebuilds the session, the
    chief restores the last checkpoint into the fresh PS and training continues to the target
    (run_mnist_distributed.py:146 "mon_sess.run handles AbortedError in case of preempted PS")."""
    from distributedtensorflow_b200.data import mnist
    mnist.synthetic_mnist(str(tmp_path / "data"), num_train=2048, num_test=256)
    cfg = _cluster_file(tmp_path, 1, 1)
    ckpt = tmp_path / "ckpt"
    ps = _launch("ps", 0, cfg, tmp_path)
    w0 = None
    try:
        w0 = _launch("worker", 0, cfg, tmp_path, ["--max_steps", "150", "--checkpoint_dir", str(ckpt),
                                                  "--shutdown_ps_when_done", "true"])
        # wait until a checkpoint exists and training is under way, then kill the PS hard
        deadline = time.time() + 180
        while time.time() < deadline and not (ckpt / "checkpoint").exists():
            time.sleep(0.2)
        assert (ckpt / "checkpoint").exists(), "no checkpoint was written"
        time.sleep(1.0)
        ps.kill()
        ps.wait(timeout=20)
        time.sleep(0.5)
        ps = _launch("ps", 0, cfg, tmp_path)            # new incarnation, same address, empty state
        out0 = _finish(w0, 300)
        outp = _finish(ps, 120)
        assert w0.returncode == 0, out0[-4000:]
        assert "Restoring parameters from" in out0, out0[-3000:]
        steps_seen = [int(s) for s in re.findall(r"\(global step: (\d+)\)", out0)]
        assert steps_seen and steps_seen[-1] >= 150
        assert ps.returncode == 0, outp[-2000:]
    finally:
        for p in (ps, w0):
            if p is not None and p.poll() is None:
                p.kill()
