#!/usr/bin/env python
"""Headline benchmark: global steps/sec of asynchronous parameter-server training of the
reference MNIST CNN (run_mnist_distributed.py: batch 128 per worker, Adam 5e-4, 1000 global
steps; README.md:16-23) on N B200s.

    python bench.py --gpus 1 --steps 200 --warmup 10
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

One process per GPU; rank r hosts PS shard r (its slice of every variable, optimizer state, bf16
serving copy) AND worker r, exactly the reference's ps/worker roles colocated per GPU. Weak
scaling: every worker keeps batch 128, a global step is one worker's apply, so N workers running
K steps each perform N*K global steps. ``value`` = N*K / max-over-ranks(device time of K steps).

Timing: CUDA events around EACH of the K timed step launches with an L2 flush (256 MiB write)
between them, on the step's stream, max over ranks; clocks are sampled with nvidia-smi during
the timed region. ``e2e`` is the same metric through the public API
(``MonitoredTrainingSession.run([train_op, loss, global_step])``) including the per-step H2D copy
of the batch from pinned host memory and the D2H read of loss + global_step.
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

README_BASELINE_STEPS_PER_SEC = 1000.0 / 162.0   # README.md:21 — 1 PS + 2 workers, 2 min 42 s
BATCH = 128


def reference_arm(args) -> int:
    """The unmodified reference needs TensorFlow 1.x (tf.train.Server, tf.layers, tf.app): not
    installable on this image (Python 3.12, no TF wheel in /opt/wheelhouse, no network) and the
    reference tree has no setup.py/pyproject to pip-install. See DESIGN.md "Reference arm"."""
    why = None
    ref = os.path.join(ROOT, "baseline", "_ref")
    try:
        sys.path.insert(0, ref)
        import tensorflow as tf  # noqa: F401
        if not hasattr(tf, "train") or not hasattr(tf.train, "Server"):
            why = "tensorflow is importable but is not TF1 (no tf.train.Server)"
    except Exception as e:  # ModuleNotFoundError on this image
        why = "reference needs TensorFlow 1.x: %s: %s" % (type(e).__name__, e)
    if why is None:
        why = "TF1 found but running the reference's gRPC cluster is not wired up on this box"
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps({"impl": "reference", "unavailable": why[:300]}))
    return 0


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""

    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_indices):
        self.gpus = set(gpu_indices)
        self.lines = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "--query-gpu=" + self.FIELDS, "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 9:
                continue
            try:
                idx = int(parts[0])
                if idx not in self.gpus:
                    continue
                sm.append(float(parts[1]))
                smax.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(names, parts[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(smax) if smax else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="cnn", choices=["cnn", "mlp"])
    ap.add_argument("--placement", default="auto", choices=["auto", "round_robin", "flat"])
    ap.add_argument("--no-l2-flush", action="store_true")
    ap.add_argument("--no-cuda-graph", action="store_true")
    ap.add_argument("--no-pdl", action="store_true", help="(default) no programmatic dependent launch")
    ap.add_argument("--pdl", action="store_true", help="enable programmatic dependent launch (measured slower)")
    ap.add_argument("--no-split-k", action="store_true", help="disable split-K for skinny GEMMs")
    ap.add_argument("--no-overlap", action="store_true", help="single-stream backward (no side stream)")
    ap.add_argument("--fused-apply", action="store_true", help="dW GEMM epilogue does the push+apply")
    ap.add_argument("--opt", action="append", default=[], metavar="KEY=0|1",
                    help="extra executor option, e.g. --opt zero_copy_input=0 --opt host_sink=0")
    ap.add_argument("--num-ps", type=int, default=0,
                    help="PS shards (default: one per GPU). --gpus 2 --num-ps 1 is the README's '1 PS, 2 workers'")
    ap.add_argument("--timeline", default="", metavar="PATH",
                    help="after the measurements: CUPTI kernel timeline of 6 more replayed steps on rank 0 (json)")
    ap.add_argument("--push-mode", default="auto", choices=["auto", "rmw", "mailbox"],
                    help="gradient push: worker-side remote read-modify-write, or posted stores + the "
                         "owner's apply daemon (auto = mailbox when there is more than one worker)")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    if args.warmup < 3:
        args.warmup = 3

    import numpy as np
    import torch
    import torch.distributed as dist
    import __graft_entry__ as entry
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    N = args.gpus
    if world != N and not (N == 1 and world == 1):
        if rank == 0:
            print(json.dumps({"error": "launch with torchrun --nproc-per-node %d (WORLD_SIZE=%d)" % (N, world)}))
        return 2
    if not torch.cuda.is_available():
        print(json.dumps({"error": "no CUDA device"}))
        return 2
    torch.cuda.set_device(local_rank)
    if local_rank == 0:
        entry.build()
    if os.environ.get("DTF_B200_HANG_DUMP"):
        import faulthandler
        faulthandler.dump_traceback_later(float(os.environ["DTF_B200_HANG_DUMP"]), exit=True)
    cpu_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        dist.barrier()
        # Barriers between phases run on the CPU (gloo): a rank parked inside an NCCL barrier keeps a
        # spinning kernel on its GPU, and while that kernel runs the same process cannot map a late
        # peer's mirror (cudaIpcOpenMemHandle waits for the device) - the peer it is waiting for.
        cpu_group = dist.new_group(backend="gloo")

    def barrier():
        if world > 1:
            dist.barrier(group=cpu_group)

    import distributedtensorflow_b200 as tf
    from distributedtensorflow_b200.data import mnist
    from distributedtensorflow_b200.models import mnist_cnn, mnist_mlp
    from distributedtensorflow_b200.ops import kernels as K

    # -- cluster: PS shard r and worker r on GPU r --------------------------------------------------
    base_port = int(os.environ.get("MASTER_PORT", "29500"))
    NPS = args.num_ps if 0 < args.num_ps <= N else N
    ps_addrs = ["127.0.0.1:%d/gpu:%d" % (base_port + 101 + i, i) for i in range(NPS)]
    wk_addrs = ["127.0.0.1:%d/gpu:%d" % (base_port + 201 + i, i) for i in range(N)]
    cluster = tf.train.ClusterSpec({"ps": ps_addrs, "worker": wk_addrs})
    ps_server = tf.train.Server(cluster, job_name="ps", task_index=rank) if rank < NPS else None
    wk_server = tf.train.Server(cluster, job_name="worker", task_index=rank)
    barrier()

    placement = args.placement
    if placement == "auto":
        placement = "flat" if NPS > 1 else "round_robin"
    strategy = tf.train.FlatSlicedStrategy(NPS) if placement == "flat" else None

    # -- input pipeline: synthetic MNIST-shaped uint8 images in pinned host memory ------------------
    # low-contrast blobs with 10% label noise: the achievable loss has a floor (0.50 nats), so the
    # final loss of a run is informative - a wrong gradient scale or a dropped shard does not reach it
    LABEL_NOISE = 0.10
    imgs, labels = mnist.synthetic_arrays(BATCH * 64, seed=rank, contrast=90.0, label_noise=LABEL_NOISE)
    imgs_t = torch.from_numpy(imgs).pin_memory()
    labels_t = torch.from_numpy(labels.astype(np.int32)).pin_memory()
    ds = tf.data.Dataset.from_tensor_slices((imgs_t.numpy(), labels_t.numpy()))
    ds = ds.repeat().batch(BATCH).prefetch(BATCH)
    iterator = tf.data.Iterator.from_structure(ds.output_types, ds.output_shapes)
    iter_init = iterator.make_initializer(ds)

    tf.set_random_seed(0)
    with tf.device(tf.train.replica_device_setter(worker_device="/job:worker/task:%d" % rank,
                                                  cluster=cluster, ps_strategy=strategy)):
        X, Y = iterator.get_next()
        if args.model == "cnn":
            logits, loss, global_step, train_op = mnist_cnn.build_train_graph(X, Y, 0.0005)
            model_name = "mnist_cnn (run_mnist_distributed.py:46-70)"
        else:
            y1h = None
            raise SystemExit("--model mlp is benchmarked by bench/ps_bench.py --model mlp")

    cfg = tf.ConfigProto(allow_soft_placement=True, backend="b200", device="cuda:%d" % local_rank,
                         cuda_graph=not args.no_cuda_graph, pdl=bool(args.pdl and not args.no_pdl),
                         split_k=not args.no_split_k, overlap=not args.no_overlap, fused_apply=args.fused_apply,
                         push_mode=args.push_mode,
                         **{k: (v == "1") if v in ("0", "1") else (int(v) if v.lstrip("-").isdigit() else v)
                            for k, v in (o.split("=", 1) for o in args.opt)})
    result = {}
    with tf.train.MonitoredTrainingSession(master=wk_server.target, is_chief=(rank == 0), config=cfg,
                                           recovery_wait_secs=0.05) as sess:
        sess.run(iter_init)
        raw = sess._tf_sess()
        # -- warm-up through the public API (also captures the CUDA graph) ---------------------
        for _ in range(args.warmup):
            _, l, g = sess.run([train_op, loss, global_step])
        runner = raw._runner
        state = runner.train_state()
        launches_per_step = state.launches_per_step
        barrier()
        torch.cuda.synchronize()

        # -- device-timed region: K steps, L2 flushed between them ------------------------------
        flush = None if args.no_l2_flush else torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
        starts = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
        ends = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
        sampler = ClockSampler(range(N)) if rank == 0 else None
        if sampler:
            sampler.start()
        barrier()
        torch.cuda.synchronize()
        t_wall0 = time.perf_counter()
        with torch.cuda.stream(runner.stream):
            for i in range(args.steps):
                if flush is not None:
                    flush.fill_(i & 0xFF)
                starts[i].record(runner.stream)
                runner.replay(1)
                ends[i].record(runner.stream)
        runner.stream.synchronize()
        torch.cuda.synchronize()
        barrier()
        t_wall = time.perf_counter() - t_wall0
        dev_ms = sum(s.elapsed_time(e) for s, e in zip(starts, ends))

        # back-to-back (warm L2) for context
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(runner.stream):
            e0.record(runner.stream)
            runner.replay(args.steps)
            e1.record(runner.stream)
        runner.stream.synchronize()
        b2b_ms = e0.elapsed_time(e1)
        clocks = sampler.stop() if sampler else None

        # -- end-to-end through the public API ---------------------------------------------------
        barrier()
        torch.cuda.synchronize()
        runner.host_prof = dict(stage=0.0, admit=0.0, launch=0.0, wait=0.0, n=0)
        tail_losses = []
        t0 = time.perf_counter()
        for k in range(args.steps):
            _, l, g = sess.run([train_op, loss, global_step])
            if k >= args.steps - 32:
                tail_losses.append(float(l))
        torch.cuda.synchronize()
        e2e_s = time.perf_counter() - t0
        host_prof, runner.host_prof = runner.host_prof, None
        final_loss = float(sum(tail_losses) / max(1, len(tail_losses)))   # mean of the last 32 training batches
        runner.flush_pushes()        # "owner applies" mode: my last gradients have been applied
        barrier()                    # every worker has finished: the shared counter is final
        final_step = int(sess.run(global_step))
        # diagnostics of the data plane: every rank's loss, and whether every owner's last broadcast
        # reached this rank's mirror (version[s] = gradients shard s had applied when it last wrote here)
        extra_diag = {}
        if world > 1:
            ver = None
            try:    # local part first: nothing in here may make one rank skip the collective below
                if getattr(runner, "use_mirror", False):
                    from distributedtensorflow_b200.ops import _lib as L_
                    ver = [int(v) for v in runner.store.mirror_hdr_u32(L_.MIRROR_OFF_VERSION, runner.store.num_shards)]
            except Exception:
                ver = None
            rows = [None] * world
            dist.all_gather_object(rows, (round(final_loss, 5), ver), group=cpu_group)
            extra_diag["final_loss_per_rank"] = [r[0] for r in rows]
            if all(r[1] is not None for r in rows):
                # every shard applies every worker's every push: all entries equal the final global step
                extra_diag["mirror_versions_current"] = bool(all(x == final_step for r in rows for x in r[1]))
                if not extra_diag["mirror_versions_current"]:
                    extra_diag["mirror_versions"] = [r[1] for r in rows]

        if args.timeline:
            from torch.profiler import profile, ProfilerActivity
            barrier()
            if rank == 0:
                with profile(activities=[ProfilerActivity.CUDA]) as prof:
                    runner.replay(6)
                    runner.stream.synchronize()
                evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
                evs.sort(key=lambda e: e.time_range.start)
                t00 = evs[0].time_range.start if evs else 0
                rows = [dict(t_us=round(e.time_range.start - t00, 2),
                             dur_us=round(e.time_range.end - e.time_range.start, 2), name=e.name[:70]) for e in evs]
                os.makedirs(os.path.dirname(os.path.abspath(args.timeline)), exist_ok=True)
                json.dump(rows, open(args.timeline, "w"), indent=0)
            else:
                runner.replay(6)
                runner.stream.synchronize()
            runner.flush_pushes()
            barrier()

        times = torch.tensor([dev_ms, b2b_ms, e2e_s * 1e3], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(times, op=dist.ReduceOp.MAX)
        dev_ms, b2b_ms, e2e_ms = [float(t) for t in times.tolist()]
        h2d = int(state.io_pin.numel())      # images + labels, one pinned staging buffer, one copy
        d2h = int(state.res_pin.numel() * 4) if runner.host_sink else 4 + 32   # loss + seq + global_step
        value = N * args.steps / (dev_ms / 1e3)
        result = {
            "metric": "global_steps_per_sec", "value": round(value, 2), "unit": "global steps/s",
            "n_gpus": N, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round(dev_ms / args.steps, 5), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": round(value / README_BASELINE_STEPS_PER_SEC, 1), "dtype": "bf16",
            "data": "synthetic MNIST-shaped uint8 28x28, 10 classes (blob contrast 90/255, 10% label noise); "
                    "random-init weights",
            "impl": "ours",
            "config": {"model": model_name, "global_batch": BATCH * N, "batch_per_worker": BATCH,
                       "seq_len": None, "optimizer": "Adam lr=5e-4 (TF1 rule), fp32 master + bf16 serving copy",
                       "parallelism": "async PS: %d PS shard(s) + %d worker(s), %s, %s placement"
                                      % (NPS, N, "one of each per GPU" if NPS == N else
                                         "PS task(s) on the first GPU(s), one worker per GPU", placement),
                       "l2": "not flushed" if args.no_l2_flush else "flushed between timed steps (256 MiB write)",
                       "cuda_graph": not args.no_cuda_graph, "pdl": bool(args.pdl and not args.no_pdl),
                       "split_k": not args.no_split_k, "overlap": not args.no_overlap,
                       "fused_apply": args.fused_apply,
                       "step_io": ("inputs: the step's first kernel reads the batch in place from pinned host memory "
                                   "over PCIe (%d B/step); results: the last kernel stores loss + global_step into "
                                   "pinned host memory (%d B/step)" % (h2d, d2h)) if runner.zero_copy_input and runner.host_sink
                                  else "cudaMemcpyAsync H2D of the pinned batch, D2H of loss + global_step",
                       "push": ("mailbox: posted gradient stores + owner-side apply daemon, depth %d"
                                % runner.store.arena.mailbox_depth) if runner.use_mailbox
                               else "rmw: worker-side fused push+apply"},
            "time_to_1000_global_steps_s": round(1000.0 / value, 5),
            "back_to_back_ms_per_step": round(b2b_ms / args.steps, 5),
            "e2e": {"value": round(N * args.steps / (e2e_ms / 1e3), 2), "unit": "global steps/s",
                    "ms_per_step": round(e2e_ms / args.steps, 5),
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "host_us_per_step_rank0": {k: round(v / max(1, host_prof["n"]) * 1e6, 1)
                                               for k, v in host_prof.items() if k != "n"}},
            "gpu_launches": int(launches_per_step * args.steps),
            "launches_per_step": int(launches_per_step),
            "clocks": clocks, "final_global_step": final_step, "final_loss": round(final_loss, 5),
            "loss_floor_label_noise": round(mnist.label_noise_xent_floor(LABEL_NOISE), 4),
            **extra_diag,
        }
    barrier()
    if ps_server is not None:
        ps_server.stop()
    wk_server.stop()
    if rank == 0:
        print(json.dumps(result))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
