#!/usr/bin/env python
"""Between-graph-replicated, asynchronous parameter-server training of the MNIST CNN.

Same contract as the reference's product script (run_mnist_distributed.py:164-182 there):

    python run_mnist_distributed.py --job_name ps     --task_index 0
    python run_mnist_distributed.py --job_name worker --task_index 0      # chief
    python run_mnist_distributed.py --job_name worker --task_index 1

one OS process per task, the cluster read from ``config.json`` (``{"ps": {...}, "workers":
{...}}``; position = task index). Parameter servers park in ``server.join()`` (and, unlike the
reference, leave it on Ctrl-C / SIGTERM / a shutdown message); every worker builds the same graph
under ``replica_device_setter``, trains with Adam(5e-4) on batches of 128 until the shared
``global_step`` reaches 1000, and the chief prints the loss and logs it to TensorBoard.

The literals of the reference (:77,:78,:116,:119,:138) are the defaults of ``TrainConfig`` and can
be overridden by the union of both templates' flags. On a GPU box, addresses of the form
``host:port/gpu:N`` put a task on GPU N and the step runs on the hand-written sm_100a kernels.
"""
from __future__ import annotations

import argparse
import dataclasses
import datetime
import os
import sys
import time

import numpy as np

import distributedtensorflow_b200 as tf
from distributedtensorflow_b200 import logger
from distributedtensorflow_b200.data import mnist
from distributedtensorflow_b200.models import mnist_cnn


@dataclasses.dataclass
class TrainConfig:
    config_file: str = "config.json"
    batch_size: int = 128                       # :77
    data_dir: str = "/tmp/data/"                # :78
    learning_rate: float = 0.0005               # :116
    max_steps: int = 1000                       # :119
    log_root: str = "/tmp/distributed_logs"     # :138
    checkpoint_dir: str = ""                    # the reference leaves checkpointing commented out (:130)
    backend: str = "auto"
    placement: str = "round_robin"
    shutdown_ps_when_done: bool = False
    quiet: bool = False
    fault_inject: str = ""                      # e.g. "kill@50" / "stall@50"


def build_input(cfg: TrainConfig):
    """Unshuffled, unsharded stream: every worker consumes the identical sequence (:77-85)."""
    raw = cfg.backend != "cpu"
    ds = mnist.train(cfg.data_dir, raw_uint8=raw)
    ds = ds.repeat().batch(cfg.batch_size).prefetch(cfg.batch_size)
    iterator = tf.data.Iterator.from_structure(ds.output_types, ds.output_shapes)
    return iterator, iterator.make_initializer(ds)


def serve_parameters(server) -> None:
    print("Started Parameter Server ...")
    server.join()
    print("Close Parameter Server ...")


def train_worker(cfg: TrainConfig, cluster, server, task_index: int) -> int:
    is_chief = task_index == 0
    iterator, iter_init = build_input(cfg)
    strategy = None
    if cfg.placement == "flat":
        strategy = tf.train.FlatSlicedStrategy(cluster.num_tasks("ps"))
    elif cfg.placement == "greedy":
        strategy = tf.train.GreedyLoadBalancingStrategy(cluster.num_tasks("ps"))
    setter = tf.train.replica_device_setter(worker_device="/job:worker/task:%d" % task_index,
                                            cluster=cluster, ps_strategy=strategy)
    with tf.device(setter):
        X, Y = iterator.get_next()
        logits, loss, global_step, train_op = mnist_cnn.build_train_graph(X, Y, cfg.learning_rate)

    hooks = [tf.train.StopAtStepHook(last_step=cfg.max_steps)]
    if cfg.fault_inject:
        action, _, at = cfg.fault_inject.partition("@")
        hooks.append(tf.train.FaultInjectionHook(int(at), action))
    session_config = tf.ConfigProto(allow_soft_placement=True,
                                    intra_op_parallelism_threads=os.cpu_count(),
                                    inter_op_parallelism_threads=os.cpu_count(),
                                    backend=cfg.backend)
    t_begin = time.time()
    last_step = 0
    with tf.train.MonitoredTrainingSession(master=server.target, is_chief=is_chief,
                                           checkpoint_dir=cfg.checkpoint_dir or None,
                                           config=session_config, hooks=hooks,
                                           recovery_wait_secs=1.0) as mon_sess:
        mon_sess.run(iter_init)
        logs = None
        if is_chief:
            stamp = datetime.datetime.now().strftime("%Y_%m_%d__%H_%M_%S")
            log_directory = os.path.join(cfg.log_root, stamp)
            # no flush per step on the hot path (the reference flushes every call, logger.py:165)
            logs = logger.TensorBoardOutputFormat(dir=log_directory, flush_every=100)
            print("Logging to", log_directory)
        local_step = 0
        warm_at = warm_t = warm_gstep = None      # steady-state clock: starts once set-up (first steps,
        while not mon_sess.should_stop():         # CUDA-graph capture) is behind us
            if local_step == 10:
                warm_at, warm_t, warm_gstep = local_step, time.time(), last_step
            local_step += 1
            if is_chief:
                _, train_loss, gstep = mon_sess.run([train_op, loss, global_step])
                last_step = int(gstep)
                if not cfg.quiet:
                    print("Worker ({}): loss = {:0.2f} (global step: {})".format(task_index, train_loss, gstep))
                logs.writekvs({"Global Step": gstep, "Loss": train_loss}, global_step=gstep)
            else:
                mon_sess.run(train_op)
        if logs is not None:
            logs.close()
        if cfg.shutdown_ps_when_done and is_chief:
            mon_sess._tf_sess().ps_client.request_shutdown()
    t_end = time.time()
    elapsed = t_end - t_begin
    print("Worker %d: %d local steps, %.3f s wall clock (global step %d)" % (task_index, local_step, elapsed, last_step))
    if warm_t is not None and is_chief and t_end > warm_t:
        print("Worker %d steady state: %d global steps in %.4f s = %.1f global steps/s (after the first %d local steps)"
              % (task_index, last_step - warm_gstep, t_end - warm_t, (last_step - warm_gstep) / (t_end - warm_t), warm_at))
    return 0


def main(argv=None) -> int:
    p = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    p.register("type", "bool", lambda v: v.lower() == "true")
    p.add_argument("--job_name", type=str, default="", required=True, help="One of 'ps', 'worker'")
    p.add_argument("--task_index", type=int, required=True, help="Index of task within the job")
    for f in dataclasses.fields(TrainConfig):
        kind = "bool" if f.type in (bool, "bool") else {"int": int, "float": float}.get(str(f.type), str)
        p.add_argument("--" + f.name, type=kind, default=f.default, help="default: %r" % (f.default,))
    flags, _ = p.parse_known_args(argv)
    print("run main with args =", flags)
    cfg = TrainConfig(**{f.name: getattr(flags, f.name) for f in dataclasses.fields(TrainConfig)})

    ps_hosts, worker_hosts = tf.Config(cfg.config_file).get_ps_and_worker_hosts()
    cluster = tf.train.ClusterSpec({"ps": ps_hosts, "worker": worker_hosts})
    server = tf.train.Server(cluster, job_name=flags.job_name, task_index=flags.task_index)
    if flags.job_name == "ps":
        serve_parameters(server)
        return 0
    if flags.job_name == "worker":
        return train_worker(cfg, cluster, server, flags.task_index)
    raise ValueError("--job_name must be 'ps' or 'worker', got %r" % flags.job_name)


if __name__ == "__main__":
    sys.exit(main())
