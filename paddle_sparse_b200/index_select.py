"""Row / column / entry gathers of a SparseTensor (reference: paddle_sparse/index_select.py:13-125)."""
from __future__ import annotations

from typing import Optional

import torch

from . import ops
from .storage import SparseStorage, get_layout
from .tensor import SparseTensor


def _select_major(old_ptr, old_count, minor, value, idx):
    """Shared by dim 0 (on CSR) and dim 1 (on CSC): new pointer by a one-pass gather+scan of the
    old counts, then ONE load-balanced fill kernel (index_select.py:21-33: cumsum,
    repeat_interleave, gather_csr, 2-3 gathers)."""
    new_ptr, E = ops.count2ptr(old_count, idx, want_total=True)
    major, new_minor, new_value, _ = ops.index_select_fill(old_ptr, minor, value, idx, new_ptr, E)
    return new_ptr, major, new_minor, new_value


def _to_row_major(major_ptr_len: int, cmajor, cminor, cvalue, M: int):
    """CSC-ordered selection -> row-major: a stable radix sort on the row bits alone stands in for
    argsort(K*row + col) (index_select.py:75-76); its by-product is the new rowptr."""
    csc2csr, rowptr, _ = ops.csr2csc(cminor, M, True, False)
    row, col, value = ops.transpose_gather(csc2csr, cmajor, cminor, cvalue, rowptr, M)
    return row, col, value, csc2csr, rowptr


def index_select(src: SparseTensor, dim: int, idx: torch.Tensor) -> SparseTensor:
    dim = src.dim() + dim if dim < 0 else dim
    assert idx.dim() == 1
    K = idx.numel()

    if dim == 0:  # index_select.py:17-51
        old_rowptr, col, value = src.csr()
        rowptr, row, col, value = _select_major(old_rowptr, src.storage.rowcount(), col, value, idx)
        storage = SparseStorage(row=row, rowptr=rowptr, col=col, value=value, sparse_sizes=(K, src.sparse_size(1)),
                                rowcount=ops.ptr2count(rowptr), is_sorted=True, trust_data=True)
        return src.from_storage(storage)

    if dim == 1:  # index_select.py:53-89
        old_colptr, row, value = src.csc()
        colptr, ccol, crow, cvalue = _select_major(old_colptr, src.storage.colcount(), row, value, idx)
        M = src.sparse_size(0)
        row, col, value, csc2csr, rowptr = _to_row_major(K, ccol, crow, cvalue, M)
        storage = SparseStorage(row=row, rowptr=None, col=col, value=value, sparse_sizes=(M, K), colptr=colptr,
                                colcount=ops.ptr2count(colptr), csc2csr=csc2csr, is_sorted=True, trust_data=True)
        return src.from_storage(storage)

    value = src.storage.value()
    if value is None:
        raise ValueError
    return src.set_value(value.index_select(dim - 1, idx), layout="coo")


def index_select_nnz(src: SparseTensor, idx: torch.Tensor, layout: Optional[str] = None) -> SparseTensor:
    """index_select.py:99-120."""
    assert idx.dim() == 1
    if get_layout(layout) == "csc":
        idx = ops.gather_rows(src.storage.csc2csr(), idx)
    row, col, value = src.coo()
    row, col = ops.gather_rows(row, idx), ops.gather_rows(col, idx)
    if value is not None:
        value = ops.gather_rows(value, idx)
    return SparseTensor(row=row, rowptr=None, col=col, value=value, sparse_sizes=src.sparse_sizes(), is_sorted=True)


SparseTensor.index_select = lambda self, dim, idx: index_select(self, dim, idx)
SparseTensor.index_select_nnz = lambda self, idx, layout=None: index_select_nnz(self, idx, layout)
