"""The MNIST MLP of templates/00_mnist_replica.py:138-164: ``global_step`` (int32, created FIRST so
it lands on ps:0), ``hid_w[784,H]`` ~ truncated_normal(sigma = 1/28), ``hid_b[H]`` = 0,
``sm_w[H,10]`` ~ truncated_normal(sigma = 1/sqrt(H)), ``sm_b[10]`` = 0;
``y = softmax(relu(x hid_w + hid_b) sm_w + sm_b)``;
loss = ``-sum(y_ * log(clip(y, 1e-10, 1)))`` — a SUM over the batch with one-hot float labels."""
from __future__ import annotations

import math

from ..framework import graph as G

IMAGE_PIXELS = 28


def build_model(x: G.Tensor, y_: G.Tensor, hidden_units: int = 100):
    """Returns (global_step, cross_entropy, y)."""
    global_step = G.Variable(0, name="global_step", trainable=False,
                             collections=[G.GraphKeys.GLOBAL_VARIABLES, G.GraphKeys.GLOBAL_STEP])
    hid_w = G.Variable(G.truncated_normal([IMAGE_PIXELS * IMAGE_PIXELS, hidden_units],
                                          stddev=1.0 / IMAGE_PIXELS), name="hid_w")
    hid_b = G.Variable(G.zeros([hidden_units]), name="hid_b")
    sm_w = G.Variable(G.truncated_normal([hidden_units, 10], stddev=1.0 / math.sqrt(hidden_units)),
                      name="sm_w")
    sm_b = G.Variable(G.zeros([10]), name="sm_b")
    hid_lin = G.xw_plus_b(x, hid_w, hid_b)
    hid = G.relu(hid_lin)
    y = G.softmax(G.xw_plus_b(hid, sm_w, sm_b))
    cross_entropy = -G.reduce_sum(y_ * G.log(G.clip_by_value(y, 1e-10, 1.0)))
    return global_step, cross_entropy, y
