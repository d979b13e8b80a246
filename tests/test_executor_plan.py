"""The B200 executor's graph -> plan compiler and GEMM planner (pure Python: runs without a GPU)."""
import pytest

import distributedtensorflow_b200 as tf
from distributedtensorflow_b200.framework import executor_b200 as E
from distributedtensorflow_b200.models import mnist_cnn, mnist_mlp, wide_mlp


def test_reference_cnn_is_recognised_as_the_compiled_topology():
    tf.reset_default_graph()
    x = tf.placeholder(tf.float32, [None, 784])
    y = tf.placeholder(tf.int32, [None])
    logits, loss, gs, train_op = mnist_cnn.build_train_graph(x, y, 0.0005)
    p = E.parse_plan(loss)
    assert p.kind == "cnn" and p.loss_kind == "sparse_mean" and p.num_classes == 10 and p.in_features == 784
    assert [tuple(c.k.shape) for c in p.convs] == [(5, 5, 1, 32), (5, 5, 32, 64)]
    assert [tuple(d.w.shape) for d in p.dense] == [(3136, 1024), (1024, 10)]
    assert [d.relu for d in p.dense] == [True, False]
    assert p.input_node is x and p.labels_node is y and p.logits_node is logits


def test_mlp_template_graph_with_clipped_sum_loss():
    tf.reset_default_graph()
    x = tf.placeholder(tf.float32, [None, 784])
    y_ = tf.placeholder(tf.float32, [None, 10])
    gs, xent, _ = mnist_mlp.build_model(x, y_, hidden_units=100)
    p = E.parse_plan(xent)
    assert p.kind == "mlp" and p.loss_kind == "dense_clipped_sum" and not p.convs
    assert [tuple(d.w.shape) for d in p.dense] == [(784, 100), (100, 10)]


def test_wide_mlp_and_unsupported_graphs():
    tf.reset_default_graph()
    x = tf.placeholder(tf.float32, [None, 784])
    y = tf.placeholder(tf.int32, [None])
    logits = wide_mlp.create_model(x, hidden=(256, 256))
    loss = tf.losses.sparse_softmax_cross_entropy(labels=y, logits=logits)
    p = E.parse_plan(loss)
    assert p.kind == "mlp" and [tuple(d.w.shape) for d in p.dense] == [(784, 256), (256, 256), (256, 10)]
    # a loss the compiler does not know, and a topology it does not know
    with pytest.raises(E.UnsupportedGraphError):
        E.parse_loss(tf.nn.softmax(logits))
    tf.reset_default_graph()
    x = tf.placeholder(tf.float32, [None, 784])
    img = tf.reshape(x, [-1, 28, 28, 1])
    c = tf.layers.conv2d(img, 16, 3, padding="same", activation=tf.nn.relu)       # not the reference's 5x5x32
    flat = tf.reshape(c, [-1, 28 * 28 * 16])
    lg = tf.layers.dense(flat, 10)
    with pytest.raises(E.UnsupportedGraphError):
        E.parse_plan(tf.losses.sparse_softmax_cross_entropy(labels=tf.placeholder(tf.int32, [None]), logits=lg))


def test_gemm_planner_fills_the_sms_for_skinny_shapes_and_pairs_ctas_for_large_ones():
    # dense1 forward of the CNN: 16 tiles of 128 x 64 -> split K 9 ways (144 CTAs)
    assert E.plan_gemm(128, 1024, 3136) == (64, 9)
    # its dX: 49 tiles: not split by default (SPLITK_MAX_TILES = 40)
    bn, splits = E.plan_gemm(128, 3136, 1024)
    assert bn == 64 and splits == 1
    # the logits layer: one tile, K = 1024 -> 8 slices at most (kb // 2)
    assert E.plan_gemm(128, 10, 1024) == (64, 8)
    # wide MLP layers: persistent CTA-pair kernel, no split
    assert E.plan_gemm(4096, 8192, 8192) == (256, 1)
    assert E.plan_gemm(8192, 8192, 4096) == (256, 1)
