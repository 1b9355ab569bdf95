"""Addition (reference: paddle_sparse/add.py:12-100): a dense [M,1] / [1,N] operand broadcast over the
stored entries, an [nnz] operand, or another SparseTensor (union of the entry sets, values of equal
entries summed)."""
from __future__ import annotations

from typing import Optional

import torch

from . import ops
from .tensor import SparseTensor


def _broadcast_add(src: SparseTensor, other: torch.Tensor, inplace: bool):
    """gather_csr / other[col] fused with the add (add.py:14-27, 54-68): no nnz-long temporary."""
    rowptr, col, value = src.csr()
    row_wise = other.shape[0] == src.size(0) and other.shape[1] == 1
    col_wise = other.shape[0] == 1 and other.shape[1] == src.size(1)
    if not (row_wise or col_wise):
        raise ValueError(f"Size mismatch: Expected size ({src.size(0)}, 1, ...) or (1, {src.size(1)}, ...), "
                         f"but got size {list(other.shape)}.")
    vec = (other.squeeze(1) if row_wise else other.squeeze(0)).contiguous()
    if value is None:  # stored entries count as 1 (add.py:26-27)
        one = torch.ones(col.numel(), dtype=vec.dtype, device=vec.device)
        return ops.add_row(one, rowptr, vec, one) if row_wise else ops.add_col(one, col, vec, one)
    vec = vec.to(value.dtype)
    if vec.dim() != value.dim():
        raise ValueError(f"Size mismatch: operand has {vec.dim()} value dims, the matrix has {value.dim()}")
    if tuple(vec.shape[1:]) != tuple(value.shape[1:]):
        vec = vec.expand((vec.shape[0],) + tuple(value.shape[1:])).contiguous()
    out = value if inplace else None
    return ops.add_row(value, rowptr, vec, out) if row_wise else ops.add_col(value, col, vec, out)


def _sparse_add(src: SparseTensor, other: SparseTensor) -> SparseTensor:
    """add.py:30-47.  The concatenated triple goes through the constructor unsorted, so the sort it owes and
    the coalesce fuse into one pipeline (storage.py's deferred sort): histogram, partition, bucket-local
    sort + dedup + rowptr."""
    rowA, colA, valueA = src.coo()
    rowB, colB, valueB = other.coo()
    row = torch.cat([rowA, rowB], dim=0)
    col = torch.cat([colA, colB], dim=0)
    value: Optional[torch.Tensor] = None
    if valueA is not None and valueB is not None:
        if valueA.dtype != valueB.dtype or valueA.shape[1:] != valueB.shape[1:]:
            raise ValueError("sparse + sparse: value dtypes / trailing shapes differ")
        value = torch.cat([valueA, valueB], dim=0)
    M = max(src.size(0), other.size(0))
    N = max(src.size(1), other.size(1))
    out = SparseTensor(row=row, col=col, value=value, sparse_sizes=(M, N))
    return out.coalesce(reduce="sum")


def add(src: SparseTensor, other):
    if isinstance(other, torch.Tensor):
        return src.set_value(_broadcast_add(src, other, inplace=False), layout="coo")
    if isinstance(other, SparseTensor):
        return _sparse_add(src, other)
    raise NotImplementedError


def add_(src: SparseTensor, other: torch.Tensor) -> SparseTensor:
    return src.set_value_(_broadcast_add(src, other, inplace=True), layout="coo")


def add_nnz(src: SparseTensor, other: torch.Tensor, layout: Optional[str] = None) -> SparseTensor:
    value = src.storage.value()
    if value is not None:
        value = ops.add_nnz(value, other.to(value.dtype).contiguous())
    else:
        value = other + 1
    return src.set_value(value, layout=layout)


def add_nnz_(src: SparseTensor, other: torch.Tensor, layout: Optional[str] = None) -> SparseTensor:
    value = src.storage.value()
    if value is not None:
        value = ops.add_nnz(value, other.to(value.dtype).contiguous(), out=value)
    else:
        value = other + 1
    return src.set_value_(value, layout=layout)


SparseTensor.add = lambda self, other: add(self, other)
SparseTensor.add_ = lambda self, other: add_(self, other)
SparseTensor.add_nnz = lambda self, other, layout=None: add_nnz(self, other, layout)
SparseTensor.add_nnz_ = lambda self, other, layout=None: add_nnz_(self, other, layout)
SparseTensor.__add__ = SparseTensor.add
SparseTensor.__radd__ = SparseTensor.add
SparseTensor.__iadd__ = SparseTensor.add_
