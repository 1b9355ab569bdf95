"""`narrow` / `select` (reference: paddle_sparse/narrow.py:11-100, select.py:7-11) — composed from the
hot-path kernels: a contiguous slice is an index_select over an arange.  The reference's cache-preserving
slicing arithmetic is a "next" row of SURVEY.md 8f; results (indices, values) are identical."""
from __future__ import annotations

import torch

from .index_select import index_select
from .tensor import SparseTensor


def narrow(src: SparseTensor, dim: int, start: int, length: int) -> SparseTensor:
    if dim < 0:
        dim = src.dim() + dim
    if start < 0:
        start = src.size(dim) + start
    if dim < 2:
        idx = torch.arange(start, start + length, dtype=torch.int64, device=src.device())
        return index_select(src, dim, idx)
    value = src.storage.value()
    if value is None:
        raise ValueError
    return src.set_value(value.narrow(dim - 1, start, length), layout="coo")


def select(src: SparseTensor, dim: int, idx: int) -> SparseTensor:
    return narrow(src, dim, start=idx, length=1)


SparseTensor.narrow = lambda self, dim, start, length: narrow(self, dim, start, length)
SparseTensor.select = lambda self, dim, idx: select(self, dim, idx)
