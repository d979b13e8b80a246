#!/usr/bin/env python
"""cProfile of the end-to-end hot loop (sess.run([train_op, loss, global_step]) with a fresh batch
every step) on one GPU: where the host microseconds between two device steps go."""
import cProfile
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402

import distributedtensorflow_b200 as tf  # noqa: E402
from distributedtensorflow_b200.models import mnist_cnn  # noqa: E402


def main():
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
    tf.reset_default_graph()
    server = tf.train.Server.create_local_server(num_ps=1)
    x = tf.placeholder(tf.uint8, [None, 784])
    y = tf.placeholder(tf.int32, [None])
    _, loss, gs, train_op = mnist_cnn.build_train_graph(x, y, learning_rate=0.0005)
    rng = np.random.RandomState(0)
    bx = [rng.randint(0, 256, (128, 784)).astype(np.uint8) for _ in range(4)]
    by = [rng.randint(0, 10, (128,)).astype(np.int32) for _ in range(4)]
    cfg = tf.ConfigProto(backend="b200", device="cuda:0")
    with tf.train.MonitoredTrainingSession(master=server.target, is_chief=True, config=cfg) as sess:
        for i in range(20):
            sess.run([train_op, loss, gs], feed_dict={x: bx[i % 4], y: by[i % 4]})
        t0 = time.perf_counter()
        for i in range(steps):
            sess.run([train_op, loss, gs], feed_dict={x: bx[i % 4], y: by[i % 4]})
        dt = (time.perf_counter() - t0) / steps
        print("e2e per step: %.1f us (%.0f steps/s)" % (dt * 1e6, 1.0 / dt))
        pr = cProfile.Profile()
        pr.enable()
        for i in range(steps):
            sess.run([train_op, loss, gs], feed_dict={x: bx[i % 4], y: by[i % 4]})
        pr.disable()
        pstats.Stats(pr).sort_stats("tottime").print_stats(28)
    server.stop()


if __name__ == "__main__":
    main()
