"""Input pipelines: tf.data-shaped Dataset / Iterator, MNIST IDX reader, synthetic MNIST."""
from .dataset import Dataset, Iterator  # noqa: F401
