"""Concatenation of sparse matrices (reference: paddle_sparse/cat.py:12-276): along rows (dim 0), along
columns (dim 1), block-diagonal (dim (0, 1)) or along a dense value dimension.  Which structures survive is the
reference's rule, kept exactly: a cache is stitched only when EVERY operand carries it.

The reference builds each stitched array from one elementwise add per operand plus a concat; here every array is
one ps_concat_add_i64 launch (out = concat_k(x_k + delta_k), pointers without their leading 0 after the first
operand), and the values are copied once (ps_concat_bytes)."""
from __future__ import annotations

from typing import List, Optional

import torch

from . import ops
from .storage import SparseStorage
from .tensor import SparseTensor


def _stitch(parts: List[Optional[torch.Tensor]], deltas: List[int], pointer: bool = False) -> Optional[torch.Tensor]:
    """concat of the shifted parts, or None unless every operand has the part (cat.py:79-97).  `pointer`: all but
    the first part drop their leading entry (cat.py:63: `rowptr[1:] + nnz if len(rowptrs) > 0 else rowptr`)."""
    if any(p is None for p in parts):
        return None
    skip = [pointer and k > 0 for k in range(len(parts))]
    return ops.concat_add(parts, deltas, skip)


def _values(tensors: List[SparseTensor]) -> Optional[torch.Tensor]:
    vals = [t.storage._value for t in tensors]
    if any(v is None for v in vals):
        return None
    return ops.concat_rows(vals)


def _running(sizes: List[int]) -> List[int]:
    out, acc = [], 0
    for s in sizes:
        out.append(acc)
        acc += s
    return out


def cat_first(tensors: List[SparseTensor]) -> SparseTensor:
    """cat.py:49-112: stack along rows.  row / rowptr / rowcount survive; the result stays sorted."""
    st = [t.storage for t in tensors]
    row_off = _running([t.sparse_size(0) for t in tensors])
    nnz_off = _running([t.nnz() for t in tensors])
    zeros = [0] * len(tensors)
    storage = SparseStorage(
        row=_stitch([s._row for s in st], row_off),
        rowptr=_stitch([s._rowptr for s in st], nnz_off, pointer=True),
        col=_stitch([s._col for s in st], zeros),
        value=_values(tensors),
        sparse_sizes=(sum(t.sparse_size(0) for t in tensors), max(t.sparse_size(1) for t in tensors)),
        rowcount=_stitch([s._rowcount for s in st], zeros),
        is_sorted=True, trust_data=True)
    return tensors[0].from_storage(storage)


def cat_second(tensors: List[SparseTensor]) -> SparseTensor:
    """cat.py:115-172: stack along columns.  The entries interleave by row, so the constructor sorts
    (is_sorted=False); colptr / colcount describe disjoint column blocks and survive."""
    st = [t.storage for t in tensors]
    col_off = _running([t.sparse_size(1) for t in tensors])
    nnz_off = _running([t.nnz() for t in tensors])
    zeros = [0] * len(tensors)
    storage = SparseStorage(
        row=_stitch([s.row() for s in st], zeros),
        rowptr=None,
        col=_stitch([s._col for s in st], col_off),
        value=_values(tensors),
        sparse_sizes=(max(t.sparse_size(0) for t in tensors), sum(t.sparse_size(1) for t in tensors)),
        colptr=_stitch([s._colptr for s in st], nnz_off, pointer=True),
        colcount=_stitch([s._colcount for s in st], zeros),
        is_sorted=False, trust_data=True)
    return tensors[0].from_storage(storage)


def cat_diag(tensors: List[SparseTensor]) -> SparseTensor:
    """cat.py:175-276: block-diagonal.  Every structure survives: rows / cols shift by the running sizes,
    pointers and both permutations by the running nnz."""
    assert len(tensors) > 0
    st = [t.storage for t in tensors]
    row_off = _running([t.sparse_size(0) for t in tensors])
    col_off = _running([t.sparse_size(1) for t in tensors])
    nnz_off = _running([t.nnz() for t in tensors])
    zeros = [0] * len(tensors)
    storage = SparseStorage(
        row=_stitch([s._row for s in st], row_off),
        rowptr=_stitch([s._rowptr for s in st], nnz_off, pointer=True),
        col=_stitch([s._col for s in st], col_off),
        value=_values(tensors),
        sparse_sizes=(sum(t.sparse_size(0) for t in tensors), sum(t.sparse_size(1) for t in tensors)),
        rowcount=_stitch([s._rowcount for s in st], zeros),
        colptr=_stitch([s._colptr for s in st], nnz_off, pointer=True),
        colcount=_stitch([s._colcount for s in st], zeros),
        csr2csc=_stitch([s._csr2csc for s in st], nnz_off),
        csc2csr=_stitch([s._csc2csr for s in st], nnz_off),
        is_sorted=True, trust_data=True)
    return tensors[0].from_storage(storage)


def cat(tensors: List[SparseTensor], dim):
    assert len(tensors) > 0
    if isinstance(dim, int):
        dim = tensors[0].dim() + dim if dim < 0 else dim
        if dim == 0:
            return cat_first(tensors)
        if dim == 1:
            return cat_second(tensors)
        if 1 < dim < tensors[0].dim():  # a dense value dimension (cat.py:25-32)
            values = []
            for t in tensors:
                value = t.storage.value()
                assert value is not None
                values.append(value)
            return tensors[0].set_value(torch.cat(values, dim=dim - 1), layout="coo")
        raise IndexError(f"Dimension out of range: Expected to be in range of [{-tensors[0].dim()}, "
                         f"{tensors[0].dim() - 1}], but got {dim}.")
    assert isinstance(dim, (tuple, list))
    assert len(dim) == 2
    assert sorted(dim) == [0, 1]
    return cat_diag(tensors)
