#!/usr/bin/env python
"""Kernel timeline of one captured training step (CUPTI through torch.profiler): start offset,
duration and stream of every kernel, so the critical path through the two-stream graph is visible.
(ncu serialises kernels and nsys is not installed; CUPTI activity records keep concurrency.)"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import distributedtensorflow_b200 as tf  # noqa: E402
from distributedtensorflow_b200.models import mnist_cnn  # noqa: E402


def main():
    tf.reset_default_graph()
    server = tf.train.Server.create_local_server(num_ps=1)
    x = tf.placeholder(tf.uint8, [None, 784])
    y = tf.placeholder(tf.int32, [None])
    _, loss, gs, train_op = mnist_cnn.build_train_graph(x, y, learning_rate=0.0005)
    rng = np.random.RandomState(0)
    bx = rng.randint(0, 256, (128, 784)).astype(np.uint8)
    by = rng.randint(0, 10, (128,)).astype(np.int32)
    opts = {}
    for a in sys.argv[1:]:
        k, v = a.split("=")
        opts[k] = (v == "1") if v in ("0", "1") else v
    cfg = tf.ConfigProto(backend="b200", device="cuda:0", **opts)
    with tf.train.MonitoredTrainingSession(master=server.target, is_chief=True, config=cfg) as sess:
        for i in range(10):
            sess.run([train_op, loss, gs], feed_dict={x: bx, y: by})
        runner = sess._tf_sess()._runner
        from torch.profiler import profile, ProfilerActivity
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            runner.replay(4)
            runner.stream.synchronize()
        evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
        evs.sort(key=lambda e: e.time_range.start)
        # split into steps at the H2D copies (first event of a step is a memcpy)
        rows = [(e.time_range.start, e.time_range.end - e.time_range.start, e.name[:60]) for e in evs]
        starts = [i for i, r in enumerate(rows) if "Memcpy HtoD" in r[2] or "memcpy" in r[2].lower() and "HtoD" in r[2]]
        steps = starts                     # one H2D copy node opens every step
        if len(steps) >= 3:
            a, b = steps[-2], steps[-1]
        else:
            a, b = 0, len(rows)
        t0 = rows[a][0]
        out = []
        for st, du, nm in rows[a:b]:
            out.append(dict(t_us=round(st - t0, 2), dur_us=round(du, 2), name=nm))
            print("%8.2f  +%7.2f  %s" % (st - t0, du, nm))
        print("step span: %.2f us" % (rows[b - 1][0] + rows[b - 1][1] - t0))
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        json.dump(out, open(os.path.join(ROOT, "gpurun_out", "timeline_step.json"), "w"), indent=0)
    server.stop()


if __name__ == "__main__":
    main()
