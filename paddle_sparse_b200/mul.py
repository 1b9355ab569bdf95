"""Multiply (reference: paddle_sparse/mul.py:12-128): a dense [M,1] / [1,N] operand broadcast over the stored
entries, an [nnz] operand, or another coalesced SparseTensor (intersection of the entry sets, values of
equal entries multiplied; mul.py:35-73)."""
from __future__ import annotations

from typing import Optional

import torch

from . import ops
from .tensor import SparseTensor


def _broadcast_mul(src: SparseTensor, other: torch.Tensor, inplace: bool):
    """gather_csr / other[col] fused with the multiply (mul.py:77-90): no nnz-long temporary."""
    rowptr, col, value = src.csr()
    row_wise = other.shape[0] == src.size(0) and other.shape[1] == 1
    col_wise = other.shape[0] == 1 and other.shape[1] == src.size(1)
    if not (row_wise or col_wise):
        raise ValueError(f"Size mismatch: Expected size ({src.size(0)}, 1, ...) or (1, {src.size(1)}, ...), "
                         f"but got size {list(other.shape)}.")
    vec = (other.squeeze(1) if row_wise else other.squeeze(0)).contiguous()
    if value is None:  # the value becomes the expanded operand (mul.py:91-92)
        return ops.gather_csr(vec, rowptr, col.numel()) if row_wise else ops.gather_rows(vec, col)
    vec = vec.to(value.dtype)
    if vec.dim() != value.dim():
        raise ValueError(f"Size mismatch: operand has {vec.dim()} value dims, the matrix has {value.dim()}")
    if tuple(vec.shape[1:]) != tuple(value.shape[1:]):
        vec = vec.expand((vec.shape[0],) + tuple(value.shape[1:])).contiguous()
    out = value if inplace else None
    return ops.mul_row(value, rowptr, vec, out) if row_wise else ops.mul_col(value, col, vec, out)


def _sparse_mul(src: SparseTensor, other: SparseTensor) -> SparseTensor:
    """mul.py:35-73.  Both operands are coalesced, so the reference's concat + argsort + adjacent-equal scan
    is an intersection of two sorted sets: each entry of A looks its column up in B's row (ps_coo_match), the
    matches are compacted in order (they stay sorted by (row, col)) and the two values multiplied."""
    if not src.is_coalesced():
        raise ValueError("The `src` tensor is not coalesced")
    if not other.is_coalesced():
        raise ValueError("The `other` tensor is not coalesced")
    rowA, colA, valueA = src.coo()
    rowptrB, colB, valueB = other.csr()
    if valueA is None or valueB is None:
        raise ValueError("Both sparse tensors must contain values")
    if valueA.dtype != valueB.dtype or valueA.shape[1:] != valueB.shape[1:]:
        raise ValueError("sparse * sparse: value dtypes / trailing shapes differ")
    M = max(src.size(0), other.size(0))
    N = max(src.size(1), other.size(1))
    pos, mask = ops.coo_match(rowA, colA, rowptrB, colB)
    sel = ops.mask_to_index(mask)
    value = ops.mul_nnz(ops.gather_rows(valueA, sel), ops.gather_rows(valueB, ops.gather_rows(pos, sel)))
    return SparseTensor(row=ops.gather_rows(rowA, sel), col=ops.gather_rows(colA, sel), value=value,
                        sparse_sizes=(M, N), is_sorted=True, trust_data=True)


def mul(src: SparseTensor, other):
    if isinstance(other, torch.Tensor):
        return src.set_value(_broadcast_mul(src, other, inplace=False), layout="coo")
    assert isinstance(other, SparseTensor)
    return _sparse_mul(src, other)


def mul_(src: SparseTensor, other: torch.Tensor) -> SparseTensor:
    return src.set_value_(_broadcast_mul(src, other, inplace=True), layout="coo")


def mul_nnz(src: SparseTensor, other: torch.Tensor, layout: Optional[str] = None) -> SparseTensor:
    value = src.storage.value()
    value = ops.mul_nnz(value, other.to(value.dtype).contiguous()) if value is not None else other
    return src.set_value(value, layout=layout)


def mul_nnz_(src: SparseTensor, other: torch.Tensor, layout: Optional[str] = None) -> SparseTensor:
    value = src.storage.value()
    value = ops.mul_nnz(value, other.to(value.dtype).contiguous(), out=value) if value is not None else other
    return src.set_value_(value, layout=layout)


SparseTensor.mul = lambda self, other: mul(self, other)
SparseTensor.mul_ = lambda self, other: mul_(self, other)
SparseTensor.mul_nnz = lambda self, other, layout=None: mul_nnz(self, other, layout)
SparseTensor.mul_nnz_ = lambda self, other, layout=None: mul_nnz_(self, other, layout)
SparseTensor.__mul__ = SparseTensor.mul
SparseTensor.__rmul__ = SparseTensor.mul
SparseTensor.__imul__ = SparseTensor.mul_
