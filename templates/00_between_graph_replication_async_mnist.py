#!/usr/bin/env python
"""Skeleton of between-graph replicated, asynchronous training — the "deploy guide" template.

Counterpart of the reference's templates/00_between_graph_replication_async_mnist.py: cluster
from ``--ps_hosts/--worker_hosts``, a ``Server`` per task, variables placed by
``replica_device_setter``, Adagrad(0.01), ``StopAtStepHook``, and a
``MonitoredTrainingSession`` with ``checkpoint_dir="/tmp/train_logs"`` (so the chief restores /
saves TF1-format checkpoints). The reference leaves the model as ``loss = ...``; here the MNIST
MLP on synthetic data fills that gap so the script runs as is.
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import distributedtensorflow_b200 as tf  # noqa: E402
from distributedtensorflow_b200.data import mnist  # noqa: E402
from distributedtensorflow_b200.models import mnist_mlp  # noqa: E402

FLAGS = None


def build_loss():
    imgs, labels = mnist.synthetic_arrays(6400, seed=0)
    ds = tf.data.Dataset.from_tensor_slices((imgs.astype(np.float32) / 255.0,
                                             np.eye(10, dtype=np.float32)[labels]))
    iterator = ds.repeat().batch(64).make_one_shot_iterator()
    x, y_ = iterator.get_next()
    w1 = tf.Variable(tf.truncated_normal([784, 100], stddev=1.0 / 28), name="hid_w")
    b1 = tf.Variable(tf.zeros([100]), name="hid_b")
    w2 = tf.Variable(tf.truncated_normal([100, 10], stddev=0.1), name="sm_w")
    b2 = tf.Variable(tf.zeros([10]), name="sm_b")
    y = tf.nn.softmax(tf.nn.xw_plus_b(tf.nn.relu(tf.nn.xw_plus_b(x, w1, b1)), w2, b2))
    return -tf.reduce_sum(y_ * tf.log(tf.clip_by_value(y, 1e-10, 1.0)))


def main(_):
    cluster = tf.cluster.cluster_from_flags(FLAGS.ps_hosts, FLAGS.worker_hosts)
    server = tf.train.Server(cluster, job_name=FLAGS.job_name, task_index=FLAGS.task_index)
    if FLAGS.job_name == "ps":
        server.join()
        return 0
    with tf.device(tf.train.replica_device_setter(worker_device="/job:worker/task:%d" % FLAGS.task_index,
                                                  cluster=cluster)):
        loss = build_loss()
        global_step = tf.train.get_or_create_global_step()
        train_op = tf.train.AdagradOptimizer(0.01).minimize(loss, global_step=global_step)
    hooks = [tf.train.StopAtStepHook(last_step=FLAGS.last_step)]
    with tf.train.MonitoredTrainingSession(master=server.target, is_chief=(FLAGS.task_index == 0),
                                           checkpoint_dir=FLAGS.checkpoint_dir, hooks=hooks,
                                           config=tf.ConfigProto(backend=FLAGS.backend)) as mon_sess:
        while not mon_sess.should_stop():
            # asynchronous step; mon_sess.run retries after a PS pre-emption (AbortedError)
            mon_sess.run(train_op)
    return 0


if __name__ == "__main__":
    parser = argparse.ArgumentParser()
    parser.register("type", "bool", lambda v: v.lower() == "true")
    parser.add_argument("--ps_hosts", type=str, default="", help="Comma-separated list of hostname:port pairs")
    parser.add_argument("--worker_hosts", type=str, default="", help="Comma-separated list of hostname:port pairs")
    parser.add_argument("--job_name", type=str, default="", help="One of 'ps', 'worker'")
    parser.add_argument("--task_index", type=int, default=0, help="Index of task within the job")
    parser.add_argument("--last_step", type=int, default=1000000)
    parser.add_argument("--checkpoint_dir", type=str, default="/tmp/train_logs")
    parser.add_argument("--backend", type=str, default="auto")
    FLAGS, unparsed = parser.parse_known_args()
    tf.app.run(main=main, argv=[sys.argv[0]] + unparsed)
