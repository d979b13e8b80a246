"""Input pipeline (IDX validation, tf.data-style Dataset / Iterator, numpy feed) and the flag
system."""
import os
import struct

import numpy as np
import pytest

import distributedtensorflow_b200 as tf
from distributedtensorflow_b200.data import mnist
from distributedtensorflow_b200.data.dataset import Dataset, Iterator


def test_idx_header_validation(tmp_path):
    d = str(tmp_path)
    imgs, labels = mnist.synthetic_arrays(32, seed=1)
    mnist.write_idx(d, mnist.TRAIN_IMAGES, mnist.TRAIN_LABELS, imgs, labels)
    assert mnist.check_image_file_header(os.path.join(d, mnist.TRAIN_IMAGES)) == 32
    assert mnist.check_labels_file_header(os.path.join(d, mnist.TRAIN_LABELS)) == 32
    bad = os.path.join(d, "bad-images")
    with open(bad, "wb") as f:
        f.write(struct.pack(">IIII", 1234, 1, 28, 28))
    with pytest.raises(ValueError, match="Invalid magic number 1234 in MNIST file"):
        mnist.check_image_file_header(bad)
    with open(bad, "wb") as f:
        f.write(struct.pack(">IIII", 2051, 1, 27, 28))
    with pytest.raises(ValueError, match="Expected 28x28 images, found 27x28"):
        mnist.check_image_file_header(bad)
    with open(bad, "wb") as f:
        f.write(struct.pack(">II", 2050, 1))
    with pytest.raises(ValueError, match="Invalid magic number 2050"):
        mnist.check_labels_file_header(bad)


def test_mnist_dataset_semantics(tmp_path):
    d = str(tmp_path)
    mnist.synthetic_mnist(d, num_train=300, num_test=50)
    ds = mnist.train(d)
    first = next(iter(ds))
    assert first[0].dtype == np.float32 and first[0].shape == (784,) and 0.0 <= first[0].min() and first[0].max() <= 1.0
    assert first[1].dtype == np.int32 and first[1].shape == ()
    assert len(list(mnist.test(d))) == 50
    raw = next(iter(mnist.train(d, raw_uint8=True)))
    assert raw[0].dtype == np.uint8 and np.allclose(raw[0].astype(np.float32) / 255.0, first[0])
    # repeat().batch(128): no shuffle, batches run across the epoch boundary, identical on every worker
    a = iter(mnist.train(d).repeat().batch(128).prefetch(128))
    b = iter(mnist.train(d).repeat().batch(128).prefetch(128))
    xs = [next(a) for _ in range(4)]
    ys = [next(b) for _ in range(4)]
    for (x1, l1), (x2, l2) in zip(xs, ys):
        assert x1.shape == (128, 784) and l1.shape == (128,)
        assert np.array_equal(x1, x2) and np.array_equal(l1, l2)
    imgs = mnist.load_idx_images(os.path.join(d, mnist.TRAIN_IMAGES))
    assert np.allclose(xs[2][0][44:], imgs[:84].astype(np.float32) / 255.0)   # 2*128+44 = 300 wraps around


def test_dataset_transformations():
    ds = Dataset.from_tensor_slices(np.arange(10))
    assert [int(x) for x in ds.map(lambda v: v * 2)] == list(range(0, 20, 2))
    assert [b.tolist() for b in ds.batch(4)] == [[0, 1, 2, 3], [4, 5, 6, 7], [8, 9]]
    assert [b.tolist() for b in ds.batch(4, drop_remainder=True)] == [[0, 1, 2, 3], [4, 5, 6, 7]]
    assert [int(x) for x in ds.shard(3, 1)] == [1, 4, 7]
    assert [int(x) for x in ds.repeat(2).take(13)] == list(range(10)) + [0, 1, 2]
    assert sorted(int(x) for x in ds.shuffle(5, seed=0)) == list(range(10))
    z = Dataset.zip((ds, Dataset.from_tensor_slices(np.arange(10) * 10)))
    assert [tuple(int(v) for v in e) for e in z.take(2)] == [(0, 0), (1, 10)]
    assert [b.tolist() for b in ds.repeat(2).batch(8)] == [[0, 1, 2, 3, 4, 5, 6, 7], [8, 9, 0, 1, 2, 3, 4, 5],
                                                          [6, 7, 8, 9]]


def test_reinitialisable_iterator_through_session():
    tf.reset_default_graph()
    ds1 = Dataset.from_tensor_slices((np.ones((6, 2), np.float32), np.arange(6, dtype=np.int32))).batch(3)
    ds2 = Dataset.from_tensor_slices((np.zeros((4, 2), np.float32), np.arange(4, dtype=np.int32) + 10)).batch(2)
    it = Iterator.from_structure(ds1.output_types, ds1.output_shapes)
    x, y = it.get_next()
    total = tf.reduce_sum(x)
    with tf.Session(config=tf.ConfigProto(backend="cpu")) as sess:
        with pytest.raises(RuntimeError):
            sess.run(total)                       # not initialised yet
        sess.run(it.make_initializer(ds1))
        assert float(sess.run(total)) == 6.0
        xv, yv = sess.run([x, y])                 # ONE element per run, shared by both outputs
        assert yv.tolist() == [3, 4, 5] and xv.shape == (3, 2)
        with pytest.raises(tf.errors.OutOfRangeError):
            sess.run(total)
        sess.run(it.make_initializer(ds2))        # re-initialise on another dataset
        assert sess.run(y).tolist() == [10, 11]


def test_numpy_feed_next_batch_and_validation_split(tmp_path):
    d = str(tmp_path)
    mnist.synthetic_mnist(d, num_train=6000, num_test=100)
    data = mnist.read_data_sets(d, one_hot=True, validation_size=500, seed=0)
    assert data.train.num_examples == 5500 and data.validation.num_examples == 500 and data.test.num_examples == 100
    xs, ys = data.train.next_batch(100)
    assert xs.shape == (100, 784) and ys.shape == (100, 10) and np.allclose(ys.sum(1), 1.0)
    seen = 100
    while data.train.epochs_completed == 0:
        xs, ys = data.train.next_batch(100)
        seen += len(xs)
    assert seen == 5600 and xs.shape == (100, 784)     # the batch that crosses the epoch is full
    with pytest.raises(ValueError):
        mnist.read_data_sets(d, validation_size=10 ** 6)


def test_flags_parse_defaults_booleans_and_unknown():
    from distributedtensorflow_b200.utils import flags as F
    F.FLAGS._reset()
    F.DEFINE_string("job_name", None, "job")
    F.DEFINE_integer("task_index", None, "idx")
    F.DEFINE_float("learning_rate", 0.01, "lr")
    F.DEFINE_boolean("sync_replicas", False, "sync")
    with pytest.raises(ValueError):
        F.DEFINE_string("job_name", None, "dup")
    left = F.FLAGS._parse(["--job_name=worker", "--task_index", "3", "--sync_replicas", "--extra", "1"])
    assert F.FLAGS.job_name == "worker" and F.FLAGS.task_index == 3 and F.FLAGS.sync_replicas is True
    assert F.FLAGS.learning_rate == 0.01 and left == ["--extra", "1"]
    F.FLAGS._parse(["--sync_replicas=false"])
    assert F.FLAGS.sync_replicas is False
    F.FLAGS._parse(["--nosync_replicas"])
    assert F.FLAGS.sync_replicas is False
    with pytest.raises(AttributeError):
        F.FLAGS.nope
    called = {}

    def main(argv):
        called["argv"] = argv
        return 0
    with pytest.raises(SystemExit) as e:
        F.app_run(main, ["prog", "--task_index=1", "leftover"])
    assert e.value.code == 0 and called["argv"] == ["prog", "leftover"]
    F.FLAGS._reset()


def test_synthetic_data_label_noise_has_a_loss_floor():
    """bench.py trains on low-contrast blobs with 10% label noise so that the final loss means
    something: the best possible predictor cannot get below the entropy of the noisy label."""
    from distributedtensorflow_b200.data import mnist
    x0, y0 = mnist.synthetic_arrays(4000, seed=3)
    x1, y1 = mnist.synthetic_arrays(4000, seed=3, contrast=90.0, label_noise=0.10)
    assert x0.shape == x1.shape == (4000, 784) and x1.dtype == np.uint8
    assert x1.mean() < x0.mean()                               # dimmer blobs
    flipped = float((y0 != y1).mean())
    assert 0.06 < flipped < 0.12                               # 10% redrawn, a tenth of those land on the same class
    assert mnist.label_noise_xent_floor(0.0) == 0.0
    assert abs(mnist.label_noise_xent_floor(0.10) - 0.5003) < 1e-3
    assert mnist.label_noise_xent_floor(1.0) == pytest.approx(np.log(10.0), rel=1e-6)


def test_synthetic_stand_in_for_mnist_is_marked(tmp_path, monkeypatch):
    """Without network access download() falls back to generated data under the real MNIST file names;
    a marker file says so, and loading from such a directory warns instead of passing for the real set."""
    from distributedtensorflow_b200.data import mnist
    import urllib.request

    def no_network(*a, **k):
        raise OSError("no route to host")
    monkeypatch.setattr(urllib.request, "urlretrieve", no_network)
    monkeypatch.setattr(mnist, "synthetic_mnist",
                        lambda d, **k: mnist.write_idx(d, mnist.TRAIN_IMAGES, mnist.TRAIN_LABELS,
                                                       *mnist.synthetic_arrays(64)))
    d = str(tmp_path / "data")
    assert not mnist.is_synthetic(d)
    path = mnist.download(d, mnist.TRAIN_IMAGES)
    assert os.path.exists(path) and mnist.is_synthetic(d)
    with pytest.raises(OSError):
        mnist.download(str(tmp_path / "strict"), mnist.TRAIN_IMAGES, allow_synthetic=False)
