"""Boolean-mask selection of rows / columns / entries (reference: paddle_sparse/masked_select.py:12-119)."""
from __future__ import annotations

from typing import Optional

import torch

from . import ops
from .index_select import _select_major, _to_row_major
from .storage import SparseStorage, get_layout
from .tensor import SparseTensor


def masked_select(src: SparseTensor, dim: int, mask: torch.Tensor) -> SparseTensor:
    """A mask is an ascending, duplicate-free index list: one stream-compaction kernel turns it into
    ids, then the index_select kernels do the rest (the reference syncs once per boolean select,
    masked_select.py:18-47 / 49-82)."""
    dim = src.dim() + dim if dim < 0 else dim
    assert mask.dim() == 1
    storage = src.storage

    if dim == 0:
        idx = ops.mask_to_index(mask)
        old_rowptr, col, value = src.csr()
        rowptr, row, col, value = _select_major(old_rowptr, storage.rowcount(), col, value, idx)
        out = SparseStorage(row=row, rowptr=None, col=col, value=value, sparse_sizes=(idx.numel(), src.sparse_size(1)),
                            rowcount=ops.ptr2count(rowptr), is_sorted=True, trust_data=True)
        out._rowptr = rowptr  # not a cache key; saves the ind2ptr the reference would redo
        return src.from_storage(out)

    if dim == 1:
        idx = ops.mask_to_index(mask)
        old_colptr, row, value = src.csc()
        colptr, ccol, crow, cvalue = _select_major(old_colptr, storage.colcount(), row, value, idx)
        M, K = src.sparse_size(0), idx.numel()
        row, col, value, csc2csr, rowptr = _to_row_major(K, ccol, crow, cvalue, M)
        out = SparseStorage(row=row, rowptr=None, col=col, value=value, sparse_sizes=(M, K),
                            colcount=ops.ptr2count(colptr), csc2csr=csc2csr, is_sorted=True, trust_data=True)
        return src.from_storage(out)

    value = src.storage.value()
    if value is None:
        raise ValueError
    idx = ops.mask_to_index(mask)
    return src.set_value(value.index_select(dim - 1, idx), layout="coo")


def masked_select_nnz(src: SparseTensor, mask: torch.Tensor, layout: Optional[str] = None) -> SparseTensor:
    """masked_select.py:93-114."""
    assert mask.dim() == 1
    if get_layout(layout) == "csc":
        idx = ops.mask_entries(mask, src.storage.csc2csr(), src.nnz())
    else:
        idx = ops.mask_to_index(mask)
    row, col, value = src.coo()
    row, col = ops.gather_rows(row, idx), ops.gather_rows(col, idx)
    if value is not None:
        value = ops.gather_rows(value, idx)
    return SparseTensor(row=row, rowptr=None, col=col, value=value, sparse_sizes=src.sparse_sizes(), is_sorted=True)


SparseTensor.masked_select = lambda self, dim, mask: masked_select(self, dim, mask)
SparseTensor.masked_select_nnz = lambda self, mask, layout=None: masked_select_nnz(self, mask, layout)
