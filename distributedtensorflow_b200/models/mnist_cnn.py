"""The reference's MNIST CNN (``create_model`` in run_mnist_distributed.py:46-70):
reshape [-1,28,28,1] -> conv5x5x32 SAME + ReLU -> maxpool 2x2/2 -> conv5x5x64 SAME + ReLU ->
maxpool 2x2/2 -> flatten 3136 -> dense 1024 + ReLU -> dense 10 (logits). No dropout.
3 274 634 parameters; variables are created kernel, bias, kernel, bias ... which is the order
``replica_device_setter`` round-robins over the PS tasks."""
from __future__ import annotations

from ..framework import graph as G

NUM_PARAMS = 3274634


def create_model(inputs: G.Tensor) -> G.Tensor:
    """Model to recognise digits in the MNIST dataset; returns the logits."""
    input_layer = G.reshape(inputs, [-1, 28, 28, 1])
    conv1 = G.conv2d(inputs=input_layer, filters=32, kernel_size=[5, 5], padding="same",
                     activation=G.relu)
    pool1 = G.max_pooling2d(inputs=conv1, pool_size=[2, 2], strides=2)
    conv2 = G.conv2d(inputs=pool1, filters=64, kernel_size=[5, 5], padding="same",
                     activation=G.relu)
    pool2 = G.max_pooling2d(inputs=conv2, pool_size=[2, 2], strides=2)
    pool2_flat = G.reshape(pool2, [-1, 7 * 7 * 64])
    dense = G.dense(inputs=pool2_flat, units=1024, activation=G.relu)
    logits = G.dense(inputs=dense, units=10)
    return logits


def build_train_graph(X: G.Tensor, Y: G.Tensor, learning_rate: float = 0.0005):
    """logits, loss, global_step, train_op exactly as run_mnist_distributed.py:111-116."""
    from ..train.optimizers import AdamOptimizer
    logits = create_model(X)
    loss = G.sparse_softmax_cross_entropy(labels=Y, logits=logits)
    global_step = G.get_or_create_global_step()
    train_op = AdamOptimizer(learning_rate).minimize(loss, global_step=global_step)
    return logits, loss, global_step, train_op
