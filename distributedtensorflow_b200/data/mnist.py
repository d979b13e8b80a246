"""MNIST input: IDX file parsing with the reference's integrity checks, the ``tf.data`` style
``train()`` / ``test()`` datasets, the numpy ``read_data_sets(...).train.next_batch()`` feed, and a
deterministic synthetic generator for boxes without network access.

Parity targets: dataset.py:30-117 (``read32``, header checks with the same error messages,
``download``, ``dataset``, ``train``, ``test``: images float32[784] in [0,1], labels int32
scalars, 60 000 / 10 000 examples) and ``input_data.read_data_sets(data_dir, one_hot=True)`` +
``mnist.train.next_batch(batch_size)`` + ``mnist.validation`` (5 000 images) of
templates/00_mnist_replica.py:92,249,268.

The GPU box has no network, so ``download`` only succeeds if the files are already there or the
URL is reachable; ``synthetic_mnist`` writes genuine IDX files (same magic numbers and layout)
filled with class-dependent blob images so that every code path — header validation, decoding,
training to a falling loss — is exercised without the real data.
"""
from __future__ import annotations

import gzip
import os
import shutil
import struct
import tempfile
from typing import Optional, Tuple

import numpy as np

from .dataset import Dataset

MNIST_URL = "https://storage.googleapis.com/cvdf-datasets/mnist/"
TRAIN_IMAGES, TRAIN_LABELS = "train-images-idx3-ubyte", "train-labels-idx1-ubyte"
TEST_IMAGES, TEST_LABELS = "t10k-images-idx3-ubyte", "t10k-labels-idx1-ubyte"


def read32(bytestream) -> int:
    """Read 4 bytes from bytestream as an unsigned 32-bit big-endian integer."""
    return struct.unpack(">I", bytestream.read(4))[0]


def check_image_file_header(filename: str) -> int:
    """Validates an IDX3 image file; returns the image count."""
    with open(filename, "rb") as f:
        magic = read32(f)
        num = read32(f)
        rows = read32(f)
        cols = read32(f)
        if magic != 2051:
            raise ValueError("Invalid magic number %d in MNIST file %s" % (magic, f.name))
        if rows != 28 or cols != 28:
            raise ValueError("Invalid MNIST file %s: Expected 28x28 images, found %dx%d" % (f.name, rows, cols))
    return num


def check_labels_file_header(filename: str) -> int:
    with open(filename, "rb") as f:
        magic = read32(f)
        num = read32(f)
        if magic != 2049:
            raise ValueError("Invalid magic number %d in MNIST file %s" % (magic, f.name))
    return num


def download(directory: str, filename: str, allow_synthetic: bool = True) -> str:
    """Returns the path of ``filename`` under ``directory``, downloading + gunzipping it on first
    use. Without network access the synthetic generator fills in (unless disabled)."""
    filepath = os.path.join(directory, filename)
    if os.path.exists(filepath):
        return filepath
    os.makedirs(directory, exist_ok=True)
    url = MNIST_URL + filename + ".gz"
    try:
        import urllib.request
        fd, zipped = tempfile.mkstemp(suffix=".gz")
        os.close(fd)
        print("Downloading %s to %s" % (url, zipped))
        urllib.request.urlretrieve(url, zipped)
        with gzip.open(zipped, "rb") as f_in, open(filepath, "wb") as f_out:
            shutil.copyfileobj(f_in, f_out)
        os.remove(zipped)
        return filepath
    except Exception as e:
        if not allow_synthetic:
            raise
        print("MNIST download unavailable (%s); generating synthetic MNIST-shaped data in %s"
              % (type(e).__name__, directory))
        synthetic_mnist(directory)
        return filepath


def load_idx_images(filename: str) -> np.ndarray:
    n = check_image_file_header(filename)
    with open(filename, "rb") as f:
        f.seek(16)
        data = np.frombuffer(f.read(n * 784), dtype=np.uint8)
    return data.reshape(-1, 784)


def load_idx_labels(filename: str) -> np.ndarray:
    n = check_labels_file_header(filename)
    with open(filename, "rb") as f:
        f.seek(8)
        return np.frombuffer(f.read(n), dtype=np.uint8).copy()


def load_arrays(directory: str, images_file: str, labels_file: str) -> Tuple[np.ndarray, np.ndarray]:
    images_path = download(directory, images_file)
    labels_path = download(directory, labels_file)
    return load_idx_images(images_path), load_idx_labels(labels_path)


def dataset(directory: str, images_file: str, labels_file: str, raw_uint8: bool = False) -> Dataset:
    """(image float32[784] in [0,1], label int32 scalar) pairs, in file order.

    ``raw_uint8=True`` keeps the pixels as uint8 (the B200 step normalises on the device, so only
    one byte per pixel crosses PCIe)."""
    imgs, labels = load_arrays(directory, images_file, labels_file)
    labels = labels.astype(np.int32)
    if not raw_uint8:
        imgs = imgs.astype(np.float32) / 255.0
    return Dataset.from_tensor_slices((imgs, labels))


def train(directory: str, raw_uint8: bool = False) -> Dataset:
    return dataset(directory, TRAIN_IMAGES, TRAIN_LABELS, raw_uint8)


def test(directory: str, raw_uint8: bool = False) -> Dataset:
    return dataset(directory, TEST_IMAGES, TEST_LABELS, raw_uint8)


# ------------------------------------------------------------------------------------------------
# synthetic MNIST-shaped data
# ------------------------------------------------------------------------------------------------
def synthetic_arrays(num: int, seed: int = 0) -> Tuple[np.ndarray, np.ndarray]:
    """``num`` 28x28 uint8 images whose class (0..9) determines where a bright blob sits, plus
    noise: learnable by a linear model, so losses fall like on the real data."""
    rng = np.random.default_rng(seed)
    labels = rng.integers(0, 10, size=num).astype(np.uint8)
    yy, xx = np.mgrid[0:28, 0:28]
    centers = [(6 + 8 * (c // 4), 5 + 6 * (c % 4)) for c in range(10)]
    protos = np.stack([np.exp(-(((yy - cy) ** 2 + (xx - cx) ** 2) / 18.0)) for cy, cx in centers])
    imgs = protos[labels] * 200.0 + rng.uniform(0, 55, size=(num, 28, 28))
    return np.clip(imgs, 0, 255).astype(np.uint8).reshape(num, 784), labels


def write_idx(directory: str, images_file: str, labels_file: str, imgs: np.ndarray, labels: np.ndarray) -> None:
    os.makedirs(directory, exist_ok=True)
    with open(os.path.join(directory, images_file), "wb") as f:
        f.write(struct.pack(">IIII", 2051, len(imgs), 28, 28))
        f.write(np.ascontiguousarray(imgs, np.uint8).tobytes())
    with open(os.path.join(directory, labels_file), "wb") as f:
        f.write(struct.pack(">II", 2049, len(labels)))
        f.write(np.ascontiguousarray(labels, np.uint8).tobytes())


def synthetic_mnist(directory: str, num_train: int = 60000, num_test: int = 10000, seed: int = 0) -> None:
    if not os.path.exists(os.path.join(directory, TRAIN_IMAGES)):
        write_idx(directory, TRAIN_IMAGES, TRAIN_LABELS, *synthetic_arrays(num_train, seed))
    if not os.path.exists(os.path.join(directory, TEST_IMAGES)):
        write_idx(directory, TEST_IMAGES, TEST_LABELS, *synthetic_arrays(num_test, seed + 1))


# ------------------------------------------------------------------------------------------------
# input_data.read_data_sets (numpy feed of the MLP template)
# ------------------------------------------------------------------------------------------------
class DataSet:
    """``mnist.train`` / ``.validation`` / ``.test``: ``next_batch`` reshuffles every epoch."""

    def __init__(self, images: np.ndarray, labels: np.ndarray, one_hot: bool, seed: Optional[int] = None):
        self._images = images.astype(np.float32) / 255.0
        if one_hot:
            oh = np.zeros((len(labels), 10), np.float32)
            oh[np.arange(len(labels)), labels] = 1.0
            self._labels = oh
        else:
            self._labels = labels.astype(np.uint8)
        self._num_examples = len(images)
        self._epochs_completed = 0
        self._index_in_epoch = 0
        self._rng = np.random.default_rng(seed)

    images = property(lambda self: self._images)
    labels = property(lambda self: self._labels)
    num_examples = property(lambda self: self._num_examples)
    epochs_completed = property(lambda self: self._epochs_completed)

    def _shuffle(self) -> None:
        perm = self._rng.permutation(self._num_examples)
        self._images, self._labels = self._images[perm], self._labels[perm]

    def next_batch(self, batch_size: int, shuffle: bool = True):
        start = self._index_in_epoch
        if self._epochs_completed == 0 and start == 0 and shuffle:
            self._shuffle()
        if start + batch_size > self._num_examples:
            self._epochs_completed += 1
            rest = self._num_examples - start
            img_rest, lab_rest = self._images[start:], self._labels[start:]
            if shuffle:
                self._shuffle()
            self._index_in_epoch = batch_size - rest
            end = self._index_in_epoch
            return (np.concatenate((img_rest, self._images[:end])),
                    np.concatenate((lab_rest, self._labels[:end])))
        self._index_in_epoch += batch_size
        return self._images[start:self._index_in_epoch], self._labels[start:self._index_in_epoch]


class Datasets:
    def __init__(self, train, validation, test):
        self.train, self.validation, self.test = train, validation, test


def read_data_sets(train_dir: str, one_hot: bool = False, validation_size: int = 5000,
                   seed: Optional[int] = None) -> Datasets:
    """``input_data.read_data_sets``: 55 000 train / 5 000 validation / 10 000 test."""
    tr_i, tr_l = load_arrays(train_dir, TRAIN_IMAGES, TRAIN_LABELS)
    te_i, te_l = load_arrays(train_dir, TEST_IMAGES, TEST_LABELS)
    if not 0 <= validation_size <= len(tr_i):
        raise ValueError("Validation size should be between 0 and {}. Received: {}."
                         .format(len(tr_i), validation_size))
    return Datasets(DataSet(tr_i[validation_size:], tr_l[validation_size:], one_hot, seed),
                    DataSet(tr_i[:validation_size], tr_l[:validation_size], one_hot, seed),
                    DataSet(te_i, te_l, one_hot, seed))
