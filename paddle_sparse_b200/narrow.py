"""`narrow` / `__narrow_diag__` / `select` (reference: paddle_sparse/narrow.py:11-168, select.py:7-11).

A contiguous row range of a CSR matrix is a contiguous range of its entries, so `narrow(dim=0)` and
`__narrow_diag__` are slice-and-shift launches (`ops.slice_add` = one `ps_concat_add_i64` piece per array, the kernel
`cat` stitches with) that keep the caches the reference keeps: `rowcount` for a row range; `colptr` / `colcount` for
a column range; all of them plus both permutations for a diagonal block.  `narrow(dim=1)` is one compaction on the
column range (`ps_narrow_cols`: the reference builds a boolean mask and selects three times, narrow.py:56-62).
Every output owns its memory, like the reference's `paddle.slice` results, so in-place `mul_` on a slice never
reaches the source.  The two pointer reads (`rowptr[start]`, `rowptr[start + length]`) are the one host
synchronisation, as in the reference (`int(rowptr[0])`, narrow.py:111-113).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import ops
from .storage import SparseStorage
from .tensor import SparseTensor


def _entry_range(rowptr: torch.Tensor, start: int, length: int) -> Tuple[int, int]:
    """(rowptr[start], rowptr[start + length]) in one device-to-host copy."""
    if length == 0:
        first = int(rowptr[start])
        return first, first
    ends = rowptr[start:start + length + 1:length].tolist()
    return int(ends[0]), int(ends[1])


def _slice_rows(value: Optional[torch.Tensor], first: int, count: int) -> Optional[torch.Tensor]:
    return None if value is None else ops.concat_rows([value[first:first + count]])


def _check_range(size: int, start: int, length: int) -> None:
    if length < 0 or start < 0 or start + length > size:
        raise ValueError(f"narrow: range [{start}, {start + length}) outside a dimension of size {size}")


def narrow(src: SparseTensor, dim: int, start: int, length: int) -> SparseTensor:
    if dim < 0:
        dim = src.dim() + dim
    if start < 0:
        start = src.size(dim) + start
    start, length = int(start), int(length)

    if dim == 0:  # narrow.py:18-51
        _check_range(src.sparse_size(0), start, length)
        rowptr, col, value = src.csr()
        first, last = _entry_range(rowptr, start, length)
        count = last - first
        held_row, held_rowcount = src.storage._row, src.storage._rowcount
        storage = SparseStorage(
            row=None if held_row is None else ops.slice_add(held_row, first, count, -start),
            rowptr=ops.slice_add(rowptr, start, length + 1, -first),
            col=ops.slice_add(col, first, count),
            value=_slice_rows(value, first, count),
            sparse_sizes=(length, src.sparse_size(1)),
            rowcount=None if held_rowcount is None else ops.slice_add(held_rowcount, start, length),
            is_sorted=True, trust_data=True)
        return src.from_storage(storage)

    if dim == 1:  # narrow.py:53-92
        _check_range(src.sparse_size(1), start, length)
        row, col, value = src.coo()
        row, col, value = ops.narrow_cols(row, col, value, start, length)
        colptr, colcount = src.storage._colptr, src.storage._colcount
        if colptr is not None:
            colptr = ops.slice_add(colptr, start, length + 1, -int(colptr[start]))
        if colcount is not None:
            colcount = ops.slice_add(colcount, start, length)
        storage = SparseStorage(row=row, rowptr=None, col=col, value=value,
                                sparse_sizes=(src.sparse_size(0), length), colptr=colptr, colcount=colcount,
                                is_sorted=True, trust_data=True)
        return src.from_storage(storage)

    value = src.storage.value()  # narrow.py:94-100
    if value is None:
        raise ValueError
    return src.set_value(value.narrow(dim - 1, start, length).contiguous(), layout="coo")


def __narrow_diag__(src: SparseTensor, start: Tuple[int, int], length: Tuple[int, int]) -> SparseTensor:
    """narrow.py:103-168: the inverse of `cat_diag`; only meaningful on diagonally stacked matrices, whose blocks
    occupy the same entry range in CSR and in CSC order — which is why the permutations survive by a shift."""
    (r0, c0), (nr, nc) = (int(start[0]), int(start[1])), (int(length[0]), int(length[1]))
    _check_range(src.sparse_size(0), r0, nr)
    _check_range(src.sparse_size(1), c0, nc)
    rowptr, col, value = src.csr()
    first, last = _entry_range(rowptr, r0, nr)
    count = last - first
    st = src.storage

    def keep(t, lo, n, delta=0):
        return None if t is None else ops.slice_add(t, lo, n, delta)

    storage = SparseStorage(
        row=keep(st._row, first, count, -r0),
        rowptr=ops.slice_add(rowptr, r0, nr + 1, -first),
        col=ops.slice_add(col, first, count, -c0),
        value=_slice_rows(value, first, count),
        sparse_sizes=(nr, nc),
        rowcount=keep(st._rowcount, r0, nr),
        # narrow.py:133-134 subtracts colptr[c0]; on a diagonal stack that is `first`
        colptr=None if st._colptr is None else ops.slice_add(st._colptr, c0, nc + 1, -int(st._colptr[c0])),
        colcount=keep(st._colcount, c0, nc),
        csr2csc=keep(st._csr2csc, first, count, -first),
        csc2csr=keep(st._csc2csr, first, count, -first),
        is_sorted=True, trust_data=True)
    return src.from_storage(storage)


def select(src: SparseTensor, dim: int, idx: int) -> SparseTensor:
    return narrow(src, dim, start=idx, length=1)


SparseTensor.narrow = lambda self, dim, start, length: narrow(self, dim, start, length)
SparseTensor.__narrow_diag__ = lambda self, start, length: __narrow_diag__(self, start, length)
SparseTensor.select = lambda self, dim, idx: select(self, dim, idx)
