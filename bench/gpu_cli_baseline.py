#!/usr/bin/env python
"""The reference's own benchmark, run the reference's own way, on GPUs: `run_mnist_distributed.py`
started once per task from a config.json (README.md:17-21, run_mnist_distributed.py:119-161) - CNN,
batch 128, Adam 5e-4, StopAtStepHook(last_step=1000), asynchronous - with `host:port/gpu:N`
addresses and --backend b200. Every task is its own OS process: the PS tasks own their arenas and
apply daemons, the workers map them over CUDA IPC. The number reported is the wall clock of the
chief's training loop (session creation to the last step), as in the README's "2 min 42 s".

    python bench/gpu_cli_baseline.py [--steps 1000] [--workers 2] [--ps 1] [--gpus 2] [--out FILE]

--workers W / --ps P are spread round-robin over --gpus G devices (README topology: --ps 1
--workers 2; one of each per GPU: --ps 8 --workers 8 --gpus 8)."""
import argparse
import json
import os
import re
import socket
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--workers", type=int, default=2)
    ap.add_argument("--ps", type=int, default=1)
    ap.add_argument("--gpus", type=int, default=2)
    ap.add_argument("--placement", default="auto")
    ap.add_argument("--timeout", type=float, default=240.0)
    ap.add_argument("--fault", default="", metavar="W:ACTION@N",
                    help="fault injection for one worker, e.g. 1:kill@30 (that worker exits with code 17 at its "
                         "30th local step; the others must finish the run)")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    tmp = tempfile.mkdtemp(prefix="dtf_gpu_cli_")
    # Every task is its own process. A PS task that shared a GPU with a worker PROCESS would have its
    # drain kernels time-sliced against that worker's kernels (two contexts on one device), and a worker
    # kernel parked in its flow control would then wait for a kernel that cannot run beside it: give the
    # PS tasks GPUs of their own whenever there are enough devices, workers first.
    if args.gpus >= args.workers + args.ps:
        ps_gpu = [args.workers + i for i in range(args.ps)]
    else:
        ps_gpu = [i % args.gpus for i in range(args.ps)]
    cfg = {"ps": {"ps:%d" % i: "127.0.0.1:%d/gpu:%d" % (free_port(), ps_gpu[i]) for i in range(args.ps)},
           "workers": {"worker:%d" % i: "127.0.0.1:%d/gpu:%d" % (free_port(), i % args.gpus)
                       for i in range(args.workers)}}
    path = os.path.join(tmp, "config.json")
    json.dump(cfg, open(path, "w"))
    placement = args.placement
    if placement == "auto":
        placement = "flat" if args.ps > 1 else "round_robin"
    env = dict(os.environ, PYTHONPATH=ROOT, PYTHONUNBUFFERED="1")
    base = [sys.executable, os.path.join(ROOT, "run_mnist_distributed.py"), "--config_file", path,
            "--backend", "b200", "--placement", placement, "--data_dir", os.path.join(tmp, "data"),
            "--log_root", os.path.join(tmp, "logs"), "--max_steps", str(args.steps), "--quiet", "true"]
    sys.path.insert(0, ROOT)
    from distributedtensorflow_b200.data import mnist
    mnist.synthetic_mnist(os.path.join(tmp, "data"))
    t0 = time.time()
    logs = [open(os.path.join(tmp, "ps%d.log" % i), "w") for i in range(args.ps)]
    pss = [subprocess.Popen(base + ["--job_name", "ps", "--task_index", str(i)], env=env, stdout=logs[i],
                            stderr=subprocess.STDOUT, text=True) for i in range(args.ps)]
    fault_w, _, fault_spec = args.fault.partition(":")
    ws = [subprocess.Popen(base + ["--job_name", "worker", "--task_index", str(i)] +
                           (["--shutdown_ps_when_done", "true"] if i == 0 else []) +
                           (["--fault_inject", fault_spec] if (args.fault and int(fault_w) == i) else []), env=env,
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
          for i in range(args.workers)]
    outs = []
    try:
        for w in ws:
            outs.append(w.communicate(timeout=max(5.0, args.timeout - (time.time() - t0)))[0])
        for p in pss:
            p.wait(timeout=30)
    except subprocess.TimeoutExpired:
        for p in ws + pss:
            if p.poll() is None:
                p.kill()
        for i, w in enumerate(ws):
            try:
                out = w.communicate(timeout=5)[0]
            except Exception:
                out = ""
            sys.stderr.write("---- worker %d ----\n%s\n" % (i, (out or "")[-3000:]))
        for i in range(args.ps):
            sys.stderr.write("---- ps %d ----\n%s\n" % (i, open(os.path.join(tmp, "ps%d.log" % i)).read()[-2000:]))
        print(json.dumps({"error": "timeout after %.0f s" % args.timeout}))
        return 1
    total = time.time() - t0
    per_worker = []
    for i, out in enumerate(outs):
        if args.fault and int(fault_w) == i:
            per_worker.append(dict(local_steps=None, seconds=None, last_global_step=None, exit_code=ws[i].returncode))
            continue
        m = re.search(r"Worker %d: (\d+) local steps, ([0-9.]+) s wall clock \(global step (\d+)\)" % i, out)
        if m is None:
            sys.stderr.write("---- worker %d ----\n%s\n" % (i, out[-3000:]))
            print(json.dumps({"error": "worker %d did not finish" % i}))
            return 1
        per_worker.append(dict(local_steps=int(m.group(1)), seconds=float(m.group(2)), last_global_step=int(m.group(3))))
    chief = per_worker[0]
    steady = re.search(r"steady state: (\d+) global steps in ([0-9.]+) s = ([0-9.]+) global steps/s", outs[0])
    result = {"config": "%d PS + %d workers as %d OS processes on %d GPU(s), run_mnist_distributed.py --backend b200, "
                        "CNN B=128 Adam 5e-4, StopAtStepHook(last_step=%d), %s placement"
                        % (args.ps, args.workers, args.ps + args.workers, args.gpus, args.steps, placement),
              "global_steps": chief["last_global_step"], "seconds_chief_loop": chief["seconds"],
              "global_steps_per_sec": round(chief["last_global_step"] / chief["seconds"], 2),
              "seconds_including_process_start": round(total, 2), "workers": per_worker,
              "steady_state": (dict(global_steps=int(steady.group(1)), seconds=float(steady.group(2)),
                                    global_steps_per_sec=float(steady.group(3)),
                                    seconds_per_1000_global_steps=round(1000.0 * float(steady.group(2)) / max(1, int(steady.group(1))), 4))
                               if steady else None),
              "reference_readme_seconds": 162, "speedup_vs_readme": round(162.0 / chief["seconds"], 1),
              "data": "synthetic MNIST-shaped idx files"}
    print(json.dumps(result))
    if args.out:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        json.dump(result, open(args.out, "w"), indent=1)
    return 0


if __name__ == "__main__":
    sys.exit(main())
