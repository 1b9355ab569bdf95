"""Reductions of a SparseTensor (reference: paddle_sparse/reduce.py:12-93)."""
from __future__ import annotations

from typing import Optional

import torch

from . import ops
from .tensor import SparseTensor


def _check(reduce: str) -> str:
    if reduce not in ("sum", "add", "mean", "min", "max"):
        raise ValueError
    return "sum" if reduce == "add" else reduce


def reduction(src: SparseTensor, dim: Optional[int] = None, reduce: str = "sum") -> torch.Tensor:
    reduce = _check(reduce)
    value = src.storage.value()

    if dim is None:  # reduce.py:17-35
        if value is not None:
            return ops.reduce_all(value, reduce)
        fill = src.nnz() if reduce == "sum" else 1
        return torch.full([], fill, dtype=src.dtype(), device=src.device())

    if dim < 0:
        dim = src.dim() + dim

    if dim == 0:  # reduce.py:40-47
        if value is not None:
            st = src.storage
            parked = st.csc_ordered_value()
            if parked is not None and st.has_colptr():
                # fill_cache_ left value[csr2csc] behind: columns are contiguous runs there, no atomics needed
                return ops.segment_csr(parked, st.colptr(), None, reduce)
            accel = st.column_reduction_accel()
            if accel is not None:  # the second and later column reductions of one matrix (see the storage method)
                return ops.segment_csr(accel[0], accel[1], None, reduce)
            # mean / min / max need the per-column counts: a histogram computed once saves the second atomic per entry
            count = st.colcount_quiet() if reduce != "sum" else None
            return ops.scatter(value, st.col(), 0, None, src.size(1), reduce, count)
        if reduce == "sum":
            return src.storage.colcount().to(src.dtype())
        return torch.ones(src.size(1), dtype=src.dtype(), device=src.device())
    if dim == 1:  # reduce.py:50-56
        if value is not None:
            return ops.segment_csr(value, src.storage.rowptr(), None, reduce)
        if reduce == "sum":
            return src.storage.rowcount().to(src.dtype())
        return torch.ones(src.size(0), dtype=src.dtype(), device=src.device())
    if value is not None:  # reduce.py:59-69: a trailing value dimension, dense op on [nnz, ...]
        if reduce == "sum":
            return value.sum(dim=dim - 1)
        if reduce == "mean":
            return value.mean(dim=dim - 1)
        if reduce == "min":
            return value.min(dim=dim - 1)[0]
        return value.max(dim=dim - 1)[0]
    raise ValueError


def sum(src: SparseTensor, dim: Optional[int] = None) -> torch.Tensor:
    return reduction(src, dim, reduce="sum")


def mean(src: SparseTensor, dim: Optional[int] = None) -> torch.Tensor:
    return reduction(src, dim, reduce="mean")


def min(src: SparseTensor, dim: Optional[int] = None) -> torch.Tensor:
    return reduction(src, dim, reduce="min")


def max(src: SparseTensor, dim: Optional[int] = None) -> torch.Tensor:
    return reduction(src, dim, reduce="max")


SparseTensor.sum = lambda self, dim=None: sum(self, dim)
SparseTensor.mean = lambda self, dim=None: mean(self, dim)
SparseTensor.min = lambda self, dim=None: min(self, dim)
SparseTensor.max = lambda self, dim=None: max(self, dim)
