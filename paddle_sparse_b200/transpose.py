"""`t` / `transpose` (reference: paddle_sparse/transpose.py:9-65)."""
from __future__ import annotations

import torch

from . import ops
from .storage import SparseStorage
from .tensor import SparseTensor


def t(src: SparseTensor) -> SparseTensor:
    """transpose.py:9-33: (row, col, value) in CSC order, caches swapped.  The transposed arrays are by-products
    of the sort that builds csr2csc (SparseStorage.take_transposed_parts); when csr2csc was cached without them
    they come from one fused gather."""
    st = src.storage
    M, N = st.sparse_sizes()
    new_row, new_col, new_value = st.take_transposed_parts()
    csr2csc = st.csr2csc()
    storage = SparseStorage(row=new_row, rowptr=st._colptr, col=new_col, value=new_value, sparse_sizes=(N, M),
                            rowcount=st._colcount, colptr=st._rowptr, colcount=st._rowcount, csr2csc=st._csc2csr,
                            csc2csr=csr2csc, is_sorted=True, trust_data=True)
    return src.from_storage(storage)


SparseTensor.t = lambda self: t(self)


def transpose(index, value, m: int, n: int, coalesced: bool = True):
    """transpose.py:41-65: transposes dimensions 0 and 1 of a COO matrix."""
    row, col = index[1], index[0]
    if coalesced:
        row, col, value = ops.coalesce(row, col, value, n, m, "add")
    return torch.stack([row, col], dim=0), value
