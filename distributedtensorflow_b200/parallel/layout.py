"""Where every variable (slice) lives: the deterministic map every task derives from its graph.

Between-graph replication means no process ships its graph to another one; each task builds the
same variables in the same order and therefore computes the same placement
(run_mnist_distributed.py:107-116). ``build_layout`` turns the graph's variables into
``VarLayout`` records: one slice on one shard for TF-style whole-variable placement
(round-robin / greedy), or row-aligned slices on every shard for ``FlatSlicedStrategy``.
Optimizer slots and Adam's beta powers follow their variable (TF colocation rule).
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, List, Sequence

from .. import placement as P


@dataclass(frozen=True)
class VarSlice:
    shard: int        # PS task index
    row_begin: int
    row_end: int
    first: int        # flat element offset of the slice inside the variable
    n: int            # elements


@dataclass
class VarLayout:
    name: str
    shape: tuple
    dtype: str
    trainable: bool
    slices: List[VarSlice]
    index: int  # creation order

    @property
    def num_elements(self) -> int:
        return int(math.prod(self.shape)) if self.shape else 1

    @property
    def cols(self) -> int:
        """Innermost dimension (row width of the bf16 serving copy)."""
        return int(self.shape[-1]) if self.shape else 1

    @property
    def rows(self) -> int:
        return self.num_elements // self.cols

    @property
    def row_elems(self) -> int:
        """Elements per leading-dimension row (slicing granularity)."""
        return self.num_elements // (int(self.shape[0]) if self.shape else 1)

    @property
    def home_shard(self) -> int:
        return self.slices[0].shard

    def shards(self) -> List[int]:
        return sorted({s.shard for s in self.slices if s.n > 0})


def build_layout(variables: Sequence, num_ps: int, row_align: int = 64) -> Dict[str, VarLayout]:
    """variables: graph ``Variable`` objects in creation order."""
    out: Dict[str, VarLayout] = {}
    n_ps = max(1, num_ps)
    for idx, v in enumerate(variables):
        shape = tuple(int(d) for d in v.shape)
        n = int(math.prod(shape)) if shape else 1
        task = v.ps_task
        if task == P.SLICED and n_ps > 1 and len(shape) >= 1:
            lead = shape[0]
            row_elems = n // lead
            slices = []
            for s, (b, e) in enumerate(P.slice_rows(shape, n_ps, row_align)):
                slices.append(VarSlice(s, b, e, b * row_elems, (e - b) * row_elems))
        else:
            shard = 0 if task is None or task < 0 else task % n_ps
            lead = shape[0] if shape else 1
            slices = [VarSlice(shard, 0, lead, 0, n)]
        out[v.name] = VarLayout(v.name, shape, v.dtype, bool(v.trainable), slices, idx)
    return out


def bytes_per_shard(layout: Dict[str, VarLayout], num_ps: int, elem_bytes: int = 4) -> List[int]:
    loads = [0] * max(1, num_ps)
    for vl in layout.values():
        eb = 8 if vl.dtype == "int64" else elem_bytes
        for s in vl.slices:
            loads[s.shard] += s.n * eb
    return loads


def powers_shard(layout: Dict[str, VarLayout], var_names: Sequence[str]) -> int:
    """Shard holding Adam's beta powers: colocated with the first variable by name
    (TF: ``min(var_list, key=lambda x: x.name)``)."""
    first = min(var_names)
    return layout[first].home_shard
