"""Wide 4-layer MLP (784 -> 8192 -> 8192 -> 8192 -> 10, batch 4096) for the communication-bound
push/pull benchmark (BASELINE.json config 4): 140 746 762 parameters = 563 MB fp32 per pull and
per push, about 3.46 TFLOP per step. Same layer ops and loss as the reference CNN's dense head
(run_mnist_distributed.py:67-69,113)."""
from __future__ import annotations

from typing import Sequence

from ..framework import graph as G


def create_model(inputs: G.Tensor, hidden: Sequence[int] = (8192, 8192, 8192), classes: int = 10) -> G.Tensor:
    h = inputs
    for units in hidden:
        h = G.dense(h, units, activation=G.relu)
    return G.dense(h, classes)


def num_params(hidden: Sequence[int] = (8192, 8192, 8192), inputs: int = 784, classes: int = 10) -> int:
    dims = [inputs, *hidden, classes]
    return sum(a * b + b for a, b in zip(dims[:-1], dims[1:]))
