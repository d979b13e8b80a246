#!/usr/bin/env python
"""Distributed MNIST training and validation with model replicas — the MLP template.

Counterpart of the reference's templates/00_mnist_replica.py (TF's tools/dist_test
mnist_replica.py): one hidden layer (784 -> hidden_units -> 10) whose parameters live on the
parameter servers while the ops run on the workers; asynchronous by default, or synchronous with
``--sync_replicas`` (``SyncReplicasOptimizer`` + ``Supervisor`` + chief queue runner), optional
backup workers through ``--replicas_to_aggregate``. All 13 flags of the reference keep their names
and defaults, and the progress lines it prints ("job name = ...", "Worker N: training step K done
(global step: G)", "Training elapsed time ...", the final validation cross entropy) keep their
wording on purpose: scripts and tests that watch the reference's output keep working. Start one process per task:

    python templates/00_mnist_replica.py --job_name ps     --task_index 0
    python templates/00_mnist_replica.py --job_name worker --task_index 0
    python templates/00_mnist_replica.py --job_name worker --task_index 1 [--sync_replicas]

Extras: ``--backend cpu|b200|auto`` and ``--shutdown_ps`` (chief asks the PS tasks to exit when it
is done; the reference's PS never exits).
"""
from __future__ import annotations

import os
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import distributedtensorflow_b200 as tf  # noqa: E402
from distributedtensorflow_b200.data import mnist as input_data  # noqa: E402
from distributedtensorflow_b200.models import mnist_mlp  # noqa: E402

flags = tf.app.flags
# The 13 flags of the reference template keep their NAMES and DEFAULTS (that is the interface a user
# of templates/00_mnist_replica.py:49-88 relies on); the descriptions are this repository's.
for _kind, _name, _default, _doc in (
        ("string", "data_dir", "/tmp/mnist-data", "where the MNIST IDX files live (or get generated)"),
        ("boolean", "download_only", False, "fetch / generate the data set, then exit without training"),
        ("integer", "task_index", None, "index of this task inside its job; worker 0 is the chief and "
                                        "initialises the variables"),
        ("integer", "num_gpus", 0, "GPUs per machine; worker j uses GPU j mod num_gpus, 0 keeps it on the CPU"),
        ("integer", "replicas_to_aggregate", None, "sync mode: gradients to collect per update "
                                                   "(fewer than the worker count = backup workers; "
                                                   "default: all workers)"),
        ("integer", "hidden_units", 100, "width of the hidden layer"),
        ("integer", "train_steps", 200, "stop once the shared global step reaches this value"),
        ("integer", "batch_size", 100, "examples per worker step"),
        ("float", "learning_rate", 0.01, "Adam step size"),
        ("boolean", "sync_replicas", False, "aggregate the workers' gradients into one update per round "
                                            "(SyncReplicasOptimizer) instead of applying each on arrival"),
        ("boolean", "existing_servers", False, "connect to servers that are already running at the "
                                               "worker addresses instead of starting one in this process"),
        ("string", "ps_hosts", "localhost:2222", "parameter-server tasks, host:port[,host:port...]"),
        ("string", "worker_hosts", "localhost:2223,localhost:2224", "worker tasks, host:port[,host:port...]"),
        ("string", "job_name", None, "'ps' or 'worker'"),
        # extras of this framework
        ("string", "backend", "auto", "cpu | b200 | auto"),
        ("boolean", "shutdown_ps", False, "chief: ask the PS tasks to exit after training"),
        ("boolean", "quiet", False, "do not print every step")):
    getattr(flags, "DEFINE_" + _kind)(_name, _default, _doc)

FLAGS = flags.FLAGS


def main(unused_argv):
    mnist = input_data.read_data_sets(FLAGS.data_dir, one_hot=True)
    if FLAGS.download_only:
        return 0
    if FLAGS.job_name is None or FLAGS.job_name == "":
        raise ValueError("Must specify an explicit `job_name`")
    if FLAGS.task_index is None or FLAGS.task_index == "":
        raise ValueError("Must specify an explicit `task_index`")
    print("job name = %s" % FLAGS.job_name)
    print("task index = %d" % FLAGS.task_index)

    cluster = tf.cluster.cluster_from_flags(FLAGS.ps_hosts, FLAGS.worker_hosts)
    num_workers = cluster.num_tasks("worker")
    server = None
    if not FLAGS.existing_servers:
        server = tf.train.Server(cluster, job_name=FLAGS.job_name, task_index=FLAGS.task_index)
        if FLAGS.job_name == "ps":
            server.join()
            return 0

    is_chief = FLAGS.task_index == 0
    if FLAGS.num_gpus > 0:
        # one GPU per worker task on a machine: task j -> gpu (j mod num_gpus)
        worker_device = "/job:worker/task:%d/gpu:%d" % (FLAGS.task_index, FLAGS.task_index % FLAGS.num_gpus)
    else:
        worker_device = "/job:worker/task:%d/cpu:0" % FLAGS.task_index

    with tf.device(tf.train.replica_device_setter(worker_device=worker_device,
                                                  ps_device="/job:ps/cpu:0", cluster=cluster)):
        x = tf.placeholder(tf.float32, [None, mnist_mlp.IMAGE_PIXELS ** 2])
        y_ = tf.placeholder(tf.float32, [None, 10])
        global_step, cross_entropy, _ = mnist_mlp.build_model(x, y_, FLAGS.hidden_units)
        opt = tf.train.AdamOptimizer(FLAGS.learning_rate)
        if FLAGS.sync_replicas:
            to_aggregate = FLAGS.replicas_to_aggregate if FLAGS.replicas_to_aggregate is not None else num_workers
            opt = tf.train.SyncReplicasOptimizer(opt, replicas_to_aggregate=to_aggregate,
                                                 total_num_replicas=num_workers, name="mnist_sync_replicas")
        train_step = opt.minimize(cross_entropy, global_step=global_step)

    sv_kwargs = dict(is_chief=is_chief, logdir=tempfile.mkdtemp(), init_op=tf.global_variables_initializer(),
                     recovery_wait_secs=1, global_step=global_step)
    if FLAGS.sync_replicas:
        sv_kwargs.update(local_init_op=opt.chief_init_op if is_chief else opt.local_step_init_op,
                         ready_for_local_init_op=opt.ready_for_local_init_op)
        chief_queue_runner = opt.get_chief_queue_runner()
        sync_init_op = opt.get_init_tokens_op()
    sv = tf.train.Supervisor(**sv_kwargs)

    gpu = FLAGS.task_index % FLAGS.num_gpus if FLAGS.num_gpus > 0 else None
    sess_config = tf.ConfigProto(allow_soft_placement=True, log_device_placement=False,
                                 device_filters=["/job:ps", "/job:worker/task:%d" % FLAGS.task_index],
                                 cluster_def=cluster, backend=FLAGS.backend if gpu is None or FLAGS.backend != "auto" else "b200",
                                 device=None if gpu is None else "cuda:%d" % gpu)
    if is_chief:
        print("Worker %d: Initializing session..." % FLAGS.task_index)
    else:
        print("Worker %d: Waiting for session to be initialized..." % FLAGS.task_index)
    if FLAGS.existing_servers:
        target = "grpc://" + cluster.task_address("worker", FLAGS.task_index).split("/")[0]
        print("Using existing server at: %s" % target)
    else:
        target = server.target
    sess = sv.prepare_or_wait_for_session(target, config=sess_config)
    print("Worker %d: Session initialization complete." % FLAGS.task_index)

    if FLAGS.sync_replicas and is_chief:
        sess.run(sync_init_op)
        sv.start_queue_runners(sess, [chief_queue_runner])

    time_begin = time.time()
    print("Training begins @ %f" % time_begin)
    local_step = 0
    try:
        while True:
            batch_xs, batch_ys = mnist.train.next_batch(FLAGS.batch_size)
            _, step = sess.run([train_step, global_step], feed_dict={x: batch_xs, y_: batch_ys})
            local_step += 1
            if not FLAGS.quiet:
                print("%f: Worker %d: training step %d done (global step: %d)"
                      % (time.time(), FLAGS.task_index, local_step, step))
            if step >= FLAGS.train_steps:
                break
    except tf.errors.OutOfRangeError:
        print("Worker %d: the chief closed the token queue; stopping" % FLAGS.task_index)
    time_end = time.time()
    print("Training ends @ %f" % time_end)
    print("Training elapsed time: %f s" % (time_end - time_begin))

    val_xent = sess.run(cross_entropy, feed_dict={x: mnist.validation.images, y_: mnist.validation.labels})
    print("After %d training step(s), validation cross entropy = %g" % (FLAGS.train_steps, val_xent))
    sv.stop()
    if FLAGS.shutdown_ps and is_chief:
        sess.ps_client.request_shutdown()
    sess.close()
    return 0


if __name__ == "__main__":
    tf.app.run(main)
