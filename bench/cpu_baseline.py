#!/usr/bin/env python
"""BASELINE config 1: the reference loop (CNN, batch 128, Adam 5e-4, 1000 global steps, async) as
1 PS + 2 workers = three OS processes on localhost over the TCP transport, CPU backend.
Prints wall-clock seconds of the chief's training loop and global steps/s (README.md:21 = 162 s)."""
import json, os, re, socket, subprocess, sys, tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def main():
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
    workers = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    tmp = tempfile.mkdtemp(prefix="dtf_cpu_baseline_")
    cfg = {"ps": {"ps:0": "127.0.0.1:%d" % free_port()},
           "workers": {"worker:%d" % i: "127.0.0.1:%d" % free_port() for i in range(workers)}}
    path = os.path.join(tmp, "config.json")
    json.dump(cfg, open(path, "w"))
    env = dict(os.environ, PYTHONPATH=ROOT, PYTHONUNBUFFERED="1")
    base = [sys.executable, os.path.join(ROOT, "run_mnist_distributed.py"), "--config_file", path, "--backend", "cpu",
            "--data_dir", os.path.join(tmp, "data"), "--log_root", os.path.join(tmp, "logs"), "--max_steps", str(steps),
            "--quiet", "true"]
    sys.path.insert(0, ROOT)
    from distributedtensorflow_b200.data import mnist
    mnist.synthetic_mnist(os.path.join(tmp, "data"))
    ps = subprocess.Popen(base + ["--job_name", "ps", "--task_index", "0"], env=env, stdout=subprocess.PIPE,
                          stderr=subprocess.STDOUT, text=True)
    ws = [subprocess.Popen(base + ["--job_name", "worker", "--task_index", str(i)] +
                           (["--shutdown_ps_when_done", "true"] if i == 0 else []), env=env,
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for i in range(workers)]
    outs = [w.communicate(timeout=3600)[0] for w in ws]
    ps.communicate(timeout=60)
    m = re.search(r"Worker 0: (\d+) local steps, ([0-9.]+) s wall clock \(global step (\d+)\)", outs[0])
    secs, gstep = float(m.group(2)), int(m.group(3))
    print(json.dumps({"config": "1 PS + %d workers, CPU, TCP localhost, CNN B=128 Adam 5e-4" % workers,
                      "global_steps": gstep, "seconds": secs, "global_steps_per_sec": round(gstep / secs, 2),
                      "cpus": os.cpu_count(), "reference_readme_seconds": 162}))


if __name__ == "__main__":
    main()
