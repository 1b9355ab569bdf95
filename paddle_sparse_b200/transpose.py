"""`t` / `transpose` (reference: paddle_sparse/transpose.py:9-65)."""
from __future__ import annotations

import torch

from . import ops
from .storage import SparseStorage
from .tensor import SparseTensor


def t(src: SparseTensor) -> SparseTensor:
    """transpose.py:9-33: gather (row, col, value) through the cached csr2csc and swap the caches.
    One fused gather kernel instead of three; when colptr is cached the new row array is streamed
    from it instead of being gathered."""
    st = src.storage
    csr2csc = st.csr2csc()
    row, col, value = src.coo()
    M, N = st.sparse_sizes()
    new_row, new_col, new_value = ops.transpose_gather(csr2csc, row, col, value, st._colptr, N)
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
