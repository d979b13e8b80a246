#!/usr/bin/env python
"""Configurable parameter-server benchmark / multi-GPU check for the BASELINE.json configs.

    torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench/ps_bench.py --model {cnn,mlp,wide} --mode {async,sync} --topology {colocated,ps1} ...

* config 2  ``--model mlp --topology ps1``        1 PS shard (rank 0) + N-1 workers, async
* config 3  ``--model mlp --mode sync --placement round_robin --ckpt-dir D``   sync replicas on N
            shards + TF1 checkpoint save / restore check
* config 4  ``--model wide --batch 4096``          comm-bound wide MLP, flat-sliced shards
* ``--check`` compares the final parameters of a sync run with the single-process large-batch
  oracle on the CPU executor.

Device-timed like bench.py: CUDA events around each step, max over ranks; one JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def build_model(tf, args, X, Y):
    from distributedtensorflow_b200.models import mnist_cnn, mnist_mlp, wide_mlp
    if args.model == "cnn":
        _, loss, gs, _ = None, None, None, None
        logits = mnist_cnn.create_model(X)
        loss = tf.losses.sparse_softmax_cross_entropy(labels=Y, logits=logits)
        gs = tf.train.get_or_create_global_step()
    elif args.model == "mlp":
        gs, loss, _ = mnist_mlp.build_model(X, Y, args.hidden)
    else:
        logits = wide_mlp.create_model(X, hidden=(args.hidden,) * 3)
        loss = tf.losses.sparse_softmax_cross_entropy(labels=Y, logits=logits)
        gs = tf.train.get_or_create_global_step()
    opt = {"adam": lambda: tf.train.AdamOptimizer(args.lr), "sgd": lambda: tf.train.GradientDescentOptimizer(args.lr),
           "momentum": lambda: tf.train.MomentumOptimizer(args.lr, 0.9),
           "adagrad": lambda: tf.train.AdagradOptimizer(args.lr)}[args.optimizer]()
    return loss, gs, opt


def make_batches(args, rank, n=8):
    import numpy as np
    from distributedtensorflow_b200.data import mnist
    imgs, labels = mnist.synthetic_arrays(n * args.batch, seed=100 + rank)
    out = []
    for i in range(n):
        x = imgs[i * args.batch:(i + 1) * args.batch]
        lab = labels[i * args.batch:(i + 1) * args.batch]
        y = np.eye(10, dtype=np.float32)[lab] if args.model == "mlp" else lab.astype(np.int32)
        out.append((x, y))
    return out


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="mlp", choices=["cnn", "mlp", "wide"])
    ap.add_argument("--mode", default="async", choices=["async", "sync"])
    ap.add_argument("--topology", default="colocated", choices=["colocated", "ps1"])
    ap.add_argument("--placement", default="flat", choices=["round_robin", "flat", "greedy"])
    ap.add_argument("--optimizer", default="adam")
    ap.add_argument("--lr", type=float, default=0.001)
    ap.add_argument("--batch", type=int, default=100)
    ap.add_argument("--hidden", type=int, default=100)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--ckpt-dir", default="")
    ap.add_argument("--no-l2-flush", action="store_true")
    ap.add_argument("--push-mode", default="auto", choices=["auto", "rmw", "mailbox"])
    ap.add_argument("--opt", action="append", default=[], metavar="KEY=VAL", help="extra executor option")
    ap.add_argument("--timeline", default="", help="rank of the first worker: CUPTI kernel timeline of 3 more steps (json)")
    ap.add_argument("--replicas", type=int, default=0,
                    help="sync mode: replicas_to_aggregate (default: every worker; fewer = backup workers)")
    ap.add_argument("--lockstep", action="store_true",
                    help="async correctness mode: workers take turns, one step each, and every push is "
                         "flushed before the next worker steps - the run is then a deterministic "
                         "sequence of applies that a single-process oracle can replay")
    ap.add_argument("--dump-params", default="", help="write the final parameters (+ this worker's mirror) as npz")
    args = ap.parse_args()

    import numpy as np
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    if world > 1:
        # gloo on purpose: a PS-only rank must wait on the CPU. Parked inside an NCCL collective it
        # would occupy its GPU with a spinning kernel while its RPC thread needs cudaMalloc for the
        # shard arena, which needs the device idle -> deadlock.
        dist.init_process_group("gloo")
        dist.barrier()
    import distributedtensorflow_b200 as tf

    base_port = int(os.environ.get("MASTER_PORT", "29500"))
    if args.topology == "ps1" and world > 1:
        ps_ranks, worker_ranks = [0], list(range(1, world))
    else:
        ps_ranks, worker_ranks = list(range(world)), list(range(world))
    ps_addrs = ["127.0.0.1:%d/gpu:%d" % (base_port + 101 + i, r) for i, r in enumerate(ps_ranks)]
    wk_addrs = ["127.0.0.1:%d/gpu:%d" % (base_port + 201 + i, r) for i, r in enumerate(worker_ranks)]
    cluster = tf.train.ClusterSpec({"ps": ps_addrs, "worker": wk_addrs})
    servers = []
    if rank in ps_ranks:
        servers.append(tf.train.Server(cluster, job_name="ps", task_index=ps_ranks.index(rank)))
    is_worker = rank in worker_ranks
    wk_index = worker_ranks.index(rank) if is_worker else -1
    if is_worker:
        servers.append(tf.train.Server(cluster, job_name="worker", task_index=wk_index))
    if world > 1:
        dist.barrier()

    nps = len(ps_ranks)
    strategy = {"flat": tf.train.FlatSlicedStrategy(nps), "greedy": tf.train.GreedyLoadBalancingStrategy(nps),
                "round_robin": None}[args.placement]
    result = None
    times = torch.zeros(2, dtype=torch.float64)
    if is_worker:
        tf.set_random_seed(0)
        with tf.device(tf.train.replica_device_setter(worker_device="/job:worker/task:%d" % wk_index,
                                                      cluster=cluster, ps_strategy=strategy)):
            X = tf.placeholder(tf.uint8, [None, 784])
            Y = tf.placeholder(tf.float32 if args.model == "mlp" else tf.int32,
                               [None, 10] if args.model == "mlp" else [None])
            loss, gs, opt = build_model(tf, args, X, Y)
            nw = len(worker_ranks)
            if args.mode == "sync":
                opt = tf.train.SyncReplicasOptimizer(opt, replicas_to_aggregate=(args.replicas or nw),
                                                     total_num_replicas=nw)
            train_op = opt.minimize(loss, global_step=gs)
        hooks = [opt.make_session_run_hook(wk_index == 0)] if args.mode == "sync" else []
        cfg = tf.ConfigProto(backend="b200", device="cuda:%d" % local_rank, push_mode=args.push_mode,
                             **{k: (v == "1") if v in ("0", "1") else (int(v) if v.lstrip("-").isdigit() else v)
                                for k, v in (o.split("=", 1) for o in args.opt)})
        batches = make_batches(args, wk_index)
        with tf.train.MonitoredTrainingSession(master=servers[-1].target, is_chief=(wk_index == 0), config=cfg,
                                               hooks=hooks, recovery_wait_secs=0.05,
                                               checkpoint_dir=None) as sess:
            raw = sess._tf_sess()
            if args.lockstep:
                # round-robin turns over the workers; PS-only ranks just keep the barriers company
                for i in range(args.warmup + args.steps):
                    for turn in range(nw):
                        if turn == wk_index:
                            bx, by = batches[i % len(batches)]
                            _, l, g = sess.run([train_op, loss, gs], feed_dict={X: bx, Y: by})
                            raw._runner.flush_pushes()
                            raw._runner.stream.synchronize()
                        if world > 1:
                            dist.barrier()
                args.steps_done, args.steps, args.warmup = args.warmup + args.steps, 0, 0
            if args.mode == "sync" and args.replicas and args.replicas < nw:
                # backup workers: train until the SHARED step reaches the target (a replica whose gradient
                # arrived stale simply takes the next token); whoever is left waiting for a token after the
                # others have finished gets the closed-queue signal
                target, i, l, g = args.warmup + args.steps, 0, float("nan"), 0
                try:
                    while int(g) < target:
                        bx, by = batches[i % len(batches)]
                        _, l, g = sess.run([train_op, loss, gs], feed_dict={X: bx, Y: by})
                        i += 1
                except tf.errors.OutOfRangeError:
                    pass
                raw._runner.mark_done()
                args.local_steps, args.steps, args.warmup = i, 0, 0
            for i in range(args.warmup):
                bx, by = batches[i % len(batches)]
                _, l, g = sess.run([train_op, loss, gs], feed_dict={X: bx, Y: by})
            runner = raw._runner
            torch.cuda.synchronize()
            flush = None if args.no_l2_flush else torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
            dev_ms = 0.0
            t0 = time.perf_counter()
            for i in range(args.steps):
                bx, by = batches[(args.warmup + i) % len(batches)]
                if flush is not None:
                    flush.fill_(i & 0xFF)
                    torch.cuda.synchronize()
                s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s.record(runner.stream)
                _, l, g = sess.run([train_op, loss, gs], feed_dict={X: bx, Y: by})
                e.record(runner.stream)
                e.synchronize()
                dev_ms += s.elapsed_time(e)
            wall = time.perf_counter() - t0
            times[0], times[1] = dev_ms, wall * 1e3
            if args.timeline:
                from torch.profiler import profile, ProfilerActivity
                bx, by = batches[0]
                if wk_index == 0:
                    with profile(activities=[ProfilerActivity.CUDA]) as prof:
                        for _ in range(3):
                            sess.run([train_op, loss, gs], feed_dict={X: bx, Y: by})
                        runner.stream.synchronize()
                    evs = sorted([e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA],
                                 key=lambda e: e.time_range.start)
                    t00 = evs[0].time_range.start if evs else 0
                    json.dump([dict(t_us=round(e.time_range.start - t00, 2), dur_us=round(e.time_range.end - e.time_range.start, 2),
                                    name=e.name[:70]) for e in evs], open(args.timeline, "w"), indent=0)
                else:
                    for _ in range(3):
                        sess.run([train_op, loss, gs], feed_dict={X: bx, Y: by})
            final_step, final_loss = int(g), float(l)
            extra = {}
            if runner.use_mailbox:
                runner.flush_pushes()        # my last gradients have been applied by their owners
                extra["push"] = "mailbox"
            if getattr(runner, "sync_generic", False):
                extra["sync_path"] = "mailbox"
                extra["local_steps"] = getattr(args, "local_steps", None)
            if args.dump_params:
                if world > 1:
                    dist.barrier()           # every worker's last gradient is applied and broadcast
                names = [v.name for v in tf.trainable_variables()]
                dump = {}
                for n in names:
                    dump[n.replace("/", "__")] = runner.store.read_variable(n)
                    if runner.use_mirror:
                        for form, arr in runner.store.read_mirror(n).items():
                            dump["mirror_%s__%s" % (form, n.replace("/", "__"))] = arr
                if runner.use_mirror:
                    dump["mirror_version"] = runner.store.mirror_hdr_u32(128, nps)
                    dump["mirror_consumed"] = runner.store.mirror_hdr_u32(256, nps)
                    extra["mirror"] = True
                    extra["graph_drain_shards"] = list(runner.my_shards)
                np.savez(args.dump_params % rank if "%" in args.dump_params else args.dump_params, **dump)
            if args.check and args.mode == "sync":
                names = [v.name for v in tf.trainable_variables()]
                got = {n: runner.store.read_variable(n) for n in names}
                extra["param_checksum"] = float(sum(np.abs(v).sum() for v in got.values()))
                np.savez(os.path.join(ROOT, "gpurun_out", "sync_params_rank%d.npz" % rank),
                         **{n.replace("/", "__"): v for n, v in got.items()})
            if args.ckpt_dir and wk_index == 0:
                sv = tf.train.Saver()
                t_s = time.perf_counter()
                path = sv.save(raw, os.path.join(args.ckpt_dir, "model.ckpt"), global_step=final_step)
                extra["ckpt_save_s"] = round(time.perf_counter() - t_s, 4)
                extra["ckpt_bytes"] = sv.last_save_bytes
                extra["ckpt_write_gbs"] = round(sv.last_save_bytes / max(extra["ckpt_save_s"], 1e-9) / 1e9, 3)
                t_r = time.perf_counter()
                from distributedtensorflow_b200.train import saver as S_
                _t = S_.load_checkpoint_tensors(path)
                extra["ckpt_read_s"] = round(time.perf_counter() - t_r, 4)
                extra["ckpt_read_gbs"] = round(sum(v.nbytes for v in _t.values()) / max(extra["ckpt_read_s"], 1e-9) / 1e9, 3)
                del _t
                from distributedtensorflow_b200.ckpt import BundleReader
                with BundleReader(path) as r:
                    extra["ckpt_shards"] = r.num_shards
                    extra["ckpt_tensors"] = len(r.keys())
                    name0 = [v.name for v in tf.trainable_variables()][0]
                    extra["ckpt_roundtrip_ok"] = bool(np.array_equal(r.get_tensor(name0),
                                                                     runner.store.read_variable(name0)))
            if world > 1:
                dist.barrier()       # nobody tears its arena down while a peer may still touch it
            if runner.use_mailbox:   # every worker has flushed: the shared counter is final
                final_step = int(runner.store.read_global_step())
            result = dict(final_step=final_step, final_loss=final_loss, launches_per_step=runner.train_state().launches_per_step,
                          **extra)
    else:
        if args.lockstep and world > 1:
            for _ in range((args.warmup + args.steps) * len(worker_ranks)):
                dist.barrier()
        if args.dump_params and world > 1:
            dist.barrier()
        if world > 1:
            dist.barrier()
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
        dist.barrier()
    for s in servers:
        s.stop()
    out_rank = worker_ranks[0]
    if rank == out_rank:
        nw = len(worker_ranks)
        dev_ms, wall_ms = [float(t) for t in times.tolist()]
        steps_total = args.steps * (nw if args.mode == "async" else 1)
        print(json.dumps(dict(
            metric="global_steps_per_sec", value=round(steps_total / max(dev_ms / 1e3, 1e-9), 2), n_gpus=world,
            ms_per_step=round(dev_ms / max(args.steps, 1), 5), e2e_value=round(steps_total / max(wall_ms / 1e3, 1e-9), 2),
            config=dict(model=args.model, mode=args.mode, topology=args.topology, placement=args.placement,
                        optimizer=args.optimizer, batch_per_worker=args.batch, hidden=args.hidden,
                        workers=nw, ps_shards=nps, push_mode=args.push_mode, l2="not flushed" if args.no_l2_flush else "flushed per step"),
            **(result or {}))))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
