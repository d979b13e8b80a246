"""distributedtensorflow_b200 — a B200-native parameter-server training framework with the
capabilities of SvenGronauer/distributedTensorFlow (TF1 between-graph replication on MNIST).

The top-level namespace mirrors the slice of ``tensorflow`` the reference's scripts use, so
``import distributedtensorflow_b200 as tf`` runs them nearly verbatim::

    cluster = tf.train.ClusterSpec({"ps": ps_hosts, "worker": worker_hosts})
    server = tf.train.Server(cluster, job_name=..., task_index=...)
    with tf.device(tf.train.replica_device_setter(worker_device=..., cluster=cluster)):
        logits = create_model(X); loss = tf.losses.sparse_softmax_cross_entropy(Y, logits)
        train_op = tf.train.AdamOptimizer(5e-4).minimize(loss, tf.train.get_or_create_global_step())
    with tf.train.MonitoredTrainingSession(master=server.target, is_chief=..., hooks=[...]) as s:
        while not s.should_stop(): s.run(train_op)

See DESIGN.md for the architecture and SURVEY.md for the reference inventory.
"""
__version__ = "0.1.0"

from . import cluster, placement  # noqa: F401
from .cluster import ClusterSpec, Config  # noqa: F401
from .framework import graph as _G
from .framework.graph import (GraphKeys, Graph, Tensor, Variable, add, argmax, bias_add, cast,  # noqa: F401
                              clip_by_value, constant, equal, float32, get_default_graph,
                              get_variable, global_variables, int32, int64, log, matmul, multiply,
                              placeholder, random_normal, reduce_mean, reduce_sum,
                              reset_default_graph, reshape, set_random_seed, trainable_variables,
                              truncated_normal, uint8, zeros, ones)
from .placement import device  # noqa: F401
from . import train  # noqa: F401
from .train.session import (ConfigProto, Session, global_variables_initializer,  # noqa: F401
                            report_uninitialized_variables)
from . import data  # noqa: F401


class _Namespace:
    def __init__(self, **kw):
        self.__dict__.update(kw)


nn = _Namespace(relu=_G.relu, softmax=_G.softmax, xw_plus_b=_G.xw_plus_b, bias_add=_G.bias_add,
                conv2d=_G.conv2d_op, max_pool=_G.max_pool,
                sparse_softmax_cross_entropy_with_logits=_G.sparse_softmax_cross_entropy_with_logits)
layers = _Namespace(conv2d=_G.conv2d, max_pooling2d=_G.max_pooling2d, dense=_G.dense)
losses = _Namespace(sparse_softmax_cross_entropy=_G.sparse_softmax_cross_entropy)
initializers = _Namespace(glorot_uniform=_G.glorot_uniform_initializer, zeros=_G.zeros_initializer,
                          constant=_G.constant_initializer)
errors = _Namespace(AbortedError=train.AbortedError, UnavailableError=train.UnavailableError,
                    OutOfRangeError=train.OutOfRangeError)

from .utils import flags as _flags  # noqa: E402
app = _Namespace(run=_flags.app_run, flags=_flags)
flags = _flags
