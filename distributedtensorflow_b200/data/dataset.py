"""A ``tf.data``-shaped input pipeline: ``Dataset`` (map / zip / repeat / batch / prefetch /
shuffle / shard) and the reinitialisable ``Iterator`` the reference builds
(run_mnist_distributed.py:77-85,111: ``ds.repeat().batch(128).prefetch(128)``,
``Iterator.from_structure`` + ``make_initializer`` + ``get_next``).

Datasets here are array-backed (the whole of MNIST is 47 MB): transformations are index
arithmetic and ``batch`` yields zero-copy contiguous slices whenever the stream is unshuffled,
which it is in the reference — no shuffle and no per-worker shard, so every worker consumes the
identical sequence. ``prefetch`` is accepted for API parity; on the B200 path the step executor
stages the NEXT batch into pinned host memory while the current step runs, which is what prefetch
is for.
"""
from __future__ import annotations

from typing import Callable, Iterator as PyIterator, List, Optional, Sequence, Tuple

import numpy as np

from ..framework import graph as G


class Dataset:
    """Lazily-evaluated dataset; ``_gen()`` returns a fresh python iterator of elements, where an
    element is a tuple of numpy arrays (or a single array)."""

    def __init__(self, gen: Callable[[], PyIterator], output_types, output_shapes,
                 arrays: Optional[Tuple[np.ndarray, ...]] = None, plan: Optional[dict] = None):
        self._gen = gen
        self.output_types = output_types
        self.output_shapes = output_shapes
        # fast path bookkeeping: array-backed, unshuffled pipelines keep (arrays, plan)
        self._arrays = arrays
        self._plan = plan or {}

    # -- sources ------------------------------------------------------------------------------------
    @staticmethod
    def from_tensor_slices(tensors) -> "Dataset":
        single = not isinstance(tensors, (tuple, list))
        arrs = (np.asarray(tensors),) if single else tuple(np.asarray(t) for t in tensors)
        n = len(arrs[0])
        assert all(len(a) == n for a in arrs), "all components must have the same leading dimension"

        def gen():
            for i in range(n):
                yield arrs[0][i] if single else tuple(a[i] for a in arrs)
        types = _types(arrs, single)
        shapes = arrs[0].shape[1:] if single else tuple(a.shape[1:] for a in arrs)
        return Dataset(gen, types, shapes, arrays=arrs, plan={"single": single, "n": n, "repeat": 1,
                                                             "batch": None, "drop": False})

    @staticmethod
    def zip(datasets: Sequence["Dataset"]) -> "Dataset":
        datasets = tuple(datasets)

        def gen():
            for elems in zip(*[d._gen() for d in datasets]):
                yield tuple(elems)
        arrays, plan = None, None
        if all(d._arrays is not None and d._plan.get("single") and d._plan.get("batch") is None
               and d._plan.get("repeat") == 1 for d in datasets):
            n = min(d._plan["n"] for d in datasets)
            arrays = tuple(d._arrays[0][:n] for d in datasets)
            plan = {"single": False, "n": n, "repeat": 1, "batch": None, "drop": False}
        return Dataset(gen, tuple(d.output_types for d in datasets),
                       tuple(d.output_shapes for d in datasets), arrays=arrays, plan=plan)

    # -- transformations ----------------------------------------------------------------------------
    def map(self, fn: Callable) -> "Dataset":
        src = self

        def gen():
            for e in src._gen():
                yield fn(*e) if isinstance(e, tuple) else fn(e)
        first = next(iter(gen()), None)
        if first is None:
            return Dataset(gen, self.output_types, self.output_shapes)
        single = not isinstance(first, tuple)
        arrs = (np.asarray(first),) if single else tuple(np.asarray(x) for x in first)
        types = _types(arrs, single)
        shapes = arrs[0].shape if single else tuple(a.shape for a in arrs)
        arrays, plan = None, None
        if self._arrays is not None and self._plan.get("batch") is None and self._plan.get("repeat") == 1:
            # materialise once: element-wise maps over array-backed sources stay array-backed
            n = self._plan["n"]
            cols: List[List[np.ndarray]] = [[] for _ in arrs]
            for e in gen():
                e = (e,) if single else e
                for c, x in zip(cols, e):
                    c.append(np.asarray(x))
            arrays = tuple(np.stack(c) for c in cols)
            plan = {"single": single, "n": n, "repeat": 1, "batch": None, "drop": False}
        return Dataset(gen, types, shapes, arrays=arrays, plan=plan)

    def repeat(self, count: Optional[int] = None) -> "Dataset":
        src = self

        def gen():
            i = 0
            while count is None or i < count:
                empty = True
                for e in src._gen():
                    empty = False
                    yield e
                if empty:
                    return
                i += 1
        plan = None
        if self._arrays is not None and self._plan.get("batch") is None:
            plan = dict(self._plan, repeat=None if count is None else self._plan["repeat"] * count)
        return Dataset(gen, self.output_types, self.output_shapes, self._arrays if plan else None, plan)

    def shuffle(self, buffer_size: int, seed: Optional[int] = None) -> "Dataset":
        src = self

        def gen():
            rng = np.random.default_rng(seed)
            buf = []
            for e in src._gen():
                buf.append(e)
                if len(buf) >= buffer_size:
                    j = int(rng.integers(len(buf)))
                    buf[j], buf[-1] = buf[-1], buf[j]
                    yield buf.pop()
            while buf:
                j = int(rng.integers(len(buf)))
                buf[j], buf[-1] = buf[-1], buf[j]
                yield buf.pop()
        return Dataset(gen, self.output_types, self.output_shapes)

    def shard(self, num_shards: int, index: int) -> "Dataset":
        src = self

        def gen():
            for i, e in enumerate(src._gen()):
                if i % num_shards == index:
                    yield e
        return Dataset(gen, self.output_types, self.output_shapes)

    def take(self, count: int) -> "Dataset":
        src = self

        def gen():
            for i, e in enumerate(src._gen()):
                if i >= count:
                    return
                yield e
        return Dataset(gen, self.output_types, self.output_shapes)

    def batch(self, batch_size: int, drop_remainder: bool = False) -> "Dataset":
        src = self
        single = not isinstance(self.output_types, tuple)

        if self._arrays is not None and self._plan.get("batch") is None:
            arrays, plan = self._arrays, dict(self._plan, batch=batch_size, drop=drop_remainder)

            def fast_gen():
                # TF semantics of repeat().batch(): batches run across epoch boundaries
                n, rep = plan["n"], plan["repeat"]
                total = None if rep is None else n * rep
                pos = 0
                while total is None or pos < total:
                    want = batch_size if total is None else min(batch_size, total - pos)
                    if want < batch_size and drop_remainder:
                        return
                    s = pos % n
                    if s + want <= n:
                        out = tuple(a[s:s + want] for a in arrays)      # zero-copy view
                    else:
                        idx = (np.arange(want) + s) % n
                        out = tuple(a[idx] for a in arrays)
                    pos += want
                    yield out[0] if plan["single"] else out
            shapes = _batched_shapes(self.output_shapes, single)
            return Dataset(fast_gen, self.output_types, shapes, arrays=arrays, plan=plan)

        def gen():
            buf = []
            for e in src._gen():
                buf.append(e)
                if len(buf) == batch_size:
                    yield _stack(buf, single)
                    buf = []
            if buf and not drop_remainder:
                yield _stack(buf, single)
        return Dataset(gen, self.output_types, _batched_shapes(self.output_shapes, single))

    def prefetch(self, buffer_size: int) -> "Dataset":
        return Dataset(self._gen, self.output_types, self.output_shapes, self._arrays, self._plan)

    # -- iteration ----------------------------------------------------------------------------------
    def make_one_shot_iterator(self) -> "Iterator":
        it = Iterator(self.output_types, self.output_shapes)
        it._bind(self)
        return it

    def make_initializable_iterator(self) -> "Iterator":
        it = Iterator(self.output_types, self.output_shapes)
        it._default_dataset = self
        return it

    def __iter__(self):
        return self._gen()


def _types(arrs, single):
    def t(a):
        return {"f": "float32", "i": "int32", "u": "uint8", "b": "bool"}.get(a.dtype.kind, str(a.dtype)) \
            if a.dtype != np.int64 else "int64"
    return t(arrs[0]) if single else tuple(t(a) for a in arrs)


def _batched_shapes(shapes, single):
    if single:
        return (None,) + tuple(shapes)
    return tuple((None,) + tuple(s) for s in shapes)


def _stack(buf, single):
    if single:
        return np.stack(buf)
    return tuple(np.stack([b[i] for b in buf]) for i in range(len(buf[0])))


class IteratorInitializer:
    """The op returned by ``Iterator.make_initializer``; ``session.run`` (re)binds the dataset."""
    op = "MakeIterator"

    def __init__(self, iterator: "Iterator", dataset: Dataset):
        self.iterator = iterator
        self.dataset = dataset
        self.name = "make_initializer"


class Iterator:
    def __init__(self, output_types, output_shapes):
        self.output_types = output_types
        self.output_shapes = output_shapes
        self._it: Optional[PyIterator] = None
        self._outputs: Optional[Tuple[G.Tensor, ...]] = None
        self._default_dataset: Optional[Dataset] = None

    @staticmethod
    def from_structure(output_types, output_shapes=None) -> "Iterator":
        return Iterator(output_types, output_shapes)

    def make_initializer(self, dataset: Dataset) -> IteratorInitializer:
        return IteratorInitializer(self, dataset)

    @property
    def initializer(self) -> IteratorInitializer:
        if self._default_dataset is None:
            raise ValueError("Iterator does not have an initializer.")
        return IteratorInitializer(self, self._default_dataset)

    def _bind(self, dataset: Dataset) -> None:
        self._it = iter(dataset)

    def get_next(self):
        """Symbolic outputs; every ``session.run`` that needs them consumes ONE element."""
        if self._outputs is None:
            single = not isinstance(self.output_types, tuple)
            types = (self.output_types,) if single else self.output_types
            shapes = (self.output_shapes,) if single else (self.output_shapes or (None,) * len(types))
            outs = []
            for i, (t, s) in enumerate(zip(types, shapes)):
                node = G.Tensor("IteratorGetNext", (), {"iterator": self, "index": i},
                                tuple(s) if s is not None else None, t, "IteratorGetNext")
                outs.append(node)
            self._outputs = tuple(outs)
            self._single = single
        return self._outputs[0] if self._single else self._outputs

    def next_element(self) -> Tuple[np.ndarray, ...]:
        if self._it is None:
            raise RuntimeError("Iterator has not been initialized (run its initializer first).")
        try:
            e = next(self._it)
        except StopIteration:
            from ..train.session import OutOfRangeError
            raise OutOfRangeError("End of sequence")
        return e if isinstance(e, tuple) else (e,)
