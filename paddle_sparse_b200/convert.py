"""Interop at the edge of the path (reference: paddle_sparse/convert.py:9-28 and tensor.py:110-155, 564-591,
796-856).  Host-side glue: scipy matrices live on the host, so these converters are the place where data enters
or leaves the GPU; nothing here computes.  Where the reference hands out `paddle.sparse` tensors this package -
hosted on torch for device memory - hands out `torch.sparse` COO / CSR tensors under the torch names, and keeps the
reference's names (`to_paddle_sparse`, ...) as aliases so that call sites written for the reference still resolve."""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch

from .storage import SparseStorage, get_layout
from .tensor import SparseTensor


# ------------------------------------------------------------------ free functions (convert.py:9-28)
def to_torch_sparse(index: torch.Tensor, value: torch.Tensor, m: int, n: int) -> torch.Tensor:
    return torch.sparse_coo_tensor(index.detach(), value, (m, n) + tuple(value.shape[1:]))


def from_torch_sparse(A: torch.Tensor):
    return A.indices().detach(), A.values()


def to_scipy(index: torch.Tensor, value: torch.Tensor, m: int, n: int):
    import scipy.sparse  # imported where it is used: scipy is not needed by, and not loaded for, the GPU path
    assert not index.is_cuda and not value.is_cuda  # convert.py:18
    (row, col), data = index.detach().numpy(), value.detach().numpy()
    return scipy.sparse.coo_matrix((data, (row, col)), (m, n))


def from_scipy(A):
    A = A.tocoo()
    row, col, value = A.row.astype(np.int64), A.col.astype(np.int64), A.data
    index = torch.stack([torch.from_numpy(row), torch.from_numpy(col)], dim=0)
    return index, torch.from_numpy(value)


to_paddle_sparse, from_paddle_sparse = to_torch_sparse, from_torch_sparse


# ------------------------------------------------------------------ SparseTensor <-> scipy (tensor.py:796-856)
def tensor_from_scipy(mat, has_value: bool = True, device="cuda") -> SparseTensor:
    """tensor.py:796-825: rowptr from the CSR view, (row, col, data) from the COO view of it, and the colptr kept
    when the input already was CSC.  The arrays are uploaded once; nothing is sorted again (is_sorted=True)."""
    import scipy.sparse
    dev = torch.device(device)
    colptr = None
    if isinstance(mat, scipy.sparse.csc_matrix):
        colptr = torch.from_numpy(mat.indptr.astype(np.int64)).to(dev)
    mat = mat.tocsr()
    rowptr = torch.from_numpy(mat.indptr.astype(np.int64)).to(dev)
    mat = mat.tocoo()
    row = torch.from_numpy(mat.row.astype(np.int64)).to(dev)
    col = torch.from_numpy(mat.col.astype(np.int64)).to(dev)
    value = torch.from_numpy(mat.data).to(dev) if has_value else None
    storage = SparseStorage(row=row, rowptr=rowptr, col=col, value=value, sparse_sizes=tuple(mat.shape[:2]),
                            colptr=colptr, is_sorted=True)
    return SparseTensor.from_storage(storage)


def tensor_to_scipy(self: SparseTensor, layout: Optional[str] = None, dtype=None):
    """tensor.py:828-856."""
    import scipy.sparse
    assert self.dim() == 2
    layout = get_layout(layout)

    def np_(t):
        return t.detach().cpu().numpy()
    ones = None
    if not self.has_value():
        ones = torch.ones(self.nnz(), dtype=dtype).numpy()
    if layout == "coo":
        row, col, value = self.coo()
        return scipy.sparse.coo_matrix((np_(value) if self.has_value() else ones, (np_(row), np_(col))), self.sizes())
    if layout == "csr":
        rowptr, col, value = self.csr()
        return scipy.sparse.csr_matrix((np_(value) if self.has_value() else ones, np_(col), np_(rowptr)), self.sizes())
    colptr, row, value = self.csc()
    return scipy.sparse.csc_matrix((np_(value) if self.has_value() else ones, np_(row), np_(colptr)), self.sizes())


# ------------------------------------------------------------------ SparseTensor <-> torch.sparse (tensor.py:110-155, 564-591)
def tensor_from_torch_sparse_coo(mat: torch.Tensor, has_value: bool = True) -> SparseTensor:
    mat = mat.coalesce()
    index = mat.indices()
    return SparseTensor(row=index[0].contiguous(), rowptr=None, col=index[1].contiguous(),
                        value=mat.values() if has_value else None, sparse_sizes=(mat.shape[0], mat.shape[1]),
                        is_sorted=True, trust_data=True)


def tensor_from_torch_sparse_csr(mat: torch.Tensor, has_value: bool = True) -> SparseTensor:
    return SparseTensor(row=None, rowptr=mat.crow_indices().to(torch.int64), col=mat.col_indices().to(torch.int64),
                        value=mat.values() if has_value else None, sparse_sizes=(mat.shape[0], mat.shape[1]),
                        is_sorted=True, trust_data=True)


def tensor_to_torch_sparse_coo(self: SparseTensor, dtype=None) -> torch.Tensor:
    row, col, value = self.coo()
    if value is None:
        value = torch.ones(self.nnz(), dtype=dtype, device=self.device())
    return torch.sparse_coo_tensor(torch.stack([row, col], dim=0), value, self.sizes())


def tensor_to_torch_sparse_csr(self: SparseTensor, dtype=None) -> torch.Tensor:
    rowptr, col, value = self.csr()
    if value is None:
        value = torch.ones(self.nnz(), dtype=dtype, device=self.device())
    return torch.sparse_csr_tensor(rowptr, col, value, self.sizes())


def _no_csc(self, dtype=None):
    raise NotImplementedError("Paddle not support sparse tensor with csc layout.")  # tensor.py:587-591


SparseTensor.from_scipy = staticmethod(tensor_from_scipy)
SparseTensor.to_scipy = tensor_to_scipy
SparseTensor.from_torch_sparse_coo_tensor = staticmethod(tensor_from_torch_sparse_coo)
SparseTensor.from_torch_sparse_csr_tensor = staticmethod(tensor_from_torch_sparse_csr)
SparseTensor.to_torch_sparse_coo_tensor = tensor_to_torch_sparse_coo
SparseTensor.to_torch_sparse_csr_tensor = tensor_to_torch_sparse_csr
# the reference's names, for call sites written against it
SparseTensor.from_paddle_sparse_coo_tensor = staticmethod(tensor_from_torch_sparse_coo)
SparseTensor.from_paddle_sparse_csr_tensor = staticmethod(tensor_from_torch_sparse_csr)
SparseTensor.to_paddle_sparse_coo_tensor = tensor_to_torch_sparse_coo
SparseTensor.to_paddle_sparse_csr_tensor = tensor_to_torch_sparse_csr
SparseTensor.to_paddle_sparse_csc_tensor = _no_csc
