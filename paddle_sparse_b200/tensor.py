"""SparseTensor — the thin façade over SparseStorage, same public surface as the reference's
paddle_sparse/tensor.py:20-592 for everything that reaches the SparseStorage hot path
(constructors, coo/csr/csc, sizes and statistics, set_value, coalesce, caches, dtype moves,
`__getitem__`, to_dense).  Functions that live in their own modules in the reference
(`t`, reductions, `mul`, `index_select`, `masked_select`, `sample_adj` …) are bound onto the
class by those modules, as the reference does (e.g. reduce.py:90-93, mul.py:122-128).
Interop converters (paddle.sparse / scipy, tensor.py:110-155, 564-591, 796-856) are host-side
glue bound onto the class by convert.py.
"""
from __future__ import annotations

from typing import Any, List, Optional, Tuple

import torch

from . import ops
from .storage import SparseStorage, get_layout


class SparseTensor(object):
    storage: SparseStorage

    def __init__(self, row=None, rowptr=None, col=None, value=None, sparse_sizes=None, is_sorted: bool = False,
                 trust_data: bool = False):
        self.storage = SparseStorage(row=row, rowptr=rowptr, col=col, value=value, sparse_sizes=sparse_sizes,
                                     is_sorted=is_sorted, trust_data=trust_data)

    @classmethod
    def from_storage(cls, storage: SparseStorage):
        """tensor.py:48-64: adopts the storage (and every cache it carries) as is."""
        out = cls.__new__(cls)
        out.storage = storage
        return out

    @classmethod
    def from_edge_index(cls, edge_index, edge_attr=None, sparse_sizes=None, is_sorted: bool = False,
                        trust_data: bool = False):
        return cls(row=edge_index[0], col=edge_index[1], value=edge_attr, sparse_sizes=sparse_sizes,
                   is_sorted=is_sorted, trust_data=trust_data)

    @classmethod
    def from_dense(cls, mat: torch.Tensor, has_value: bool = True):
        """tensor.py:85-108."""
        if mat.dim() > 2:
            index = (mat.abs().sum(dim=tuple(range(2, mat.dim()))) != 0).nonzero()
        else:
            index = (mat != 0).nonzero()
        row, col = index[:, 0].contiguous(), index[:, 1].contiguous()
        value = mat[row, col] if has_value else None
        return cls(row=row, col=col, value=value, sparse_sizes=(mat.shape[0], mat.shape[1]), is_sorted=True,
                   trust_data=True)

    @classmethod
    def eye(cls, M: int, N: Optional[int] = None, has_value: bool = True, dtype=None, device="cuda",
            fill_cache: bool = False):
        """tensor.py:157-212."""
        N = M if N is None else N
        k = min(M, N)
        row = torch.arange(k, dtype=torch.int64, device=device)
        col = row.clone()
        rowptr = torch.arange(M + 1, dtype=torch.int64, device=device).clamp_(max=k)
        value = torch.ones(k, dtype=dtype or torch.float32, device=device) if has_value else None
        rowcount = colptr = colcount = csr2csc = csc2csr = None
        if fill_cache:
            rowcount = (rowptr[1:] - rowptr[:-1]).contiguous()
            colptr = torch.arange(N + 1, dtype=torch.int64, device=device).clamp_(max=k)
            colcount = (colptr[1:] - colptr[:-1]).contiguous()
            csr2csc = csc2csr = row
        storage = SparseStorage(row=row, rowptr=rowptr, col=col, value=value, sparse_sizes=(M, N), rowcount=rowcount,
                                colptr=colptr, colcount=colcount, csr2csc=csr2csc, csc2csr=csc2csr, is_sorted=True,
                                trust_data=True)
        return cls.from_storage(storage)

    def copy(self):
        return self.from_storage(self.storage)

    def clone(self):
        st = self.storage

        def c(t):
            return None if t is None else t.clone()

        return self.from_storage(SparseStorage(
            row=c(st._row), rowptr=c(st._rowptr), col=c(st._col), value=c(st._value), sparse_sizes=st._sparse_sizes,
            rowcount=c(st._rowcount), colptr=c(st._colptr), colcount=c(st._colcount), csr2csc=c(st._csr2csc),
            csc2csr=c(st._csc2csr), is_sorted=True, trust_data=True))

    def type(self, dtype, non_blocking: bool = False):
        value = self.storage.value()
        if value is None or dtype == value.dtype:
            return self
        return self.from_storage(self.storage.set_value(value.to(dtype), layout="coo"))

    def type_as(self, tensor: torch.Tensor, non_blocking: bool = False):
        return self.type(tensor.dtype)

    # ------------------------------------------------------------------ formats (tensor.py:246-257)
    def coo(self) -> Tuple[torch.Tensor, torch.Tensor, Optional[torch.Tensor]]:
        return self.storage.row(), self.storage.col(), self.storage.value()

    def csr(self) -> Tuple[torch.Tensor, torch.Tensor, Optional[torch.Tensor]]:
        return self.storage.rowptr(), self.storage.col(), self.storage.value()

    def csc(self) -> Tuple[torch.Tensor, torch.Tensor, Optional[torch.Tensor]]:
        perm = self.storage.csr2csc()
        value = self.storage.value()
        if value is not None:
            value = ops.gather_rows(value, perm)
        return self.storage.colptr(), ops.gather_rows(self.storage.row(), perm), value

    # ------------------------------------------------------------------ storage pass-throughs
    def has_value(self) -> bool:
        return self.storage.has_value()

    def set_value_(self, value, layout: Optional[str] = None):
        self.storage.set_value_(value, layout)
        return self

    def set_value(self, value, layout: Optional[str] = None):
        return self.from_storage(self.storage.set_value(value, layout))

    def sparse_sizes(self) -> Tuple[int, int]:
        return self.storage.sparse_sizes()

    def sparse_size(self, dim: int) -> int:
        return self.storage.sparse_sizes()[dim]

    def sparse_resize(self, sparse_sizes: Tuple[int, int]):
        return self.from_storage(self.storage.sparse_resize(sparse_sizes))

    def sparse_reshape(self, num_rows: int, num_cols: int):
        return self.from_storage(self.storage.sparse_reshape(num_rows, num_cols))

    def is_coalesced(self) -> bool:
        return self.storage.is_coalesced()

    def coalesce(self, reduce: str = "sum"):
        return self.from_storage(self.storage.coalesce(reduce))

    def fill_cache_(self):
        self.storage.fill_cache_()
        return self

    def clear_cache_(self):
        self.storage.clear_cache_()
        return self

    def __eq__(self, other) -> bool:
        """tensor.py:305-327."""
        if not isinstance(other, self.__class__):
            return False
        if self.sizes() != other.sizes():
            return False
        rowA, colA, valueA = self.coo()
        rowB, colB, valueB = other.coo()
        if rowA.numel() != rowB.numel() or not torch.equal(rowA, rowB) or not torch.equal(colA, colB):
            return False
        if (valueA is None) != (valueB is None):
            return False
        return valueA is None or bool(torch.equal(valueA, valueB))

    __hash__ = None

    def fill_value_(self, fill_value: float, dtype=None):
        value = torch.full((self.nnz(),), fill_value, dtype=dtype or torch.float32, device=self.device())
        return self.set_value_(value, layout="coo")

    def fill_value(self, fill_value: float, dtype=None):
        value = torch.full((self.nnz(),), fill_value, dtype=dtype or torch.float32, device=self.device())
        return self.set_value(value, layout="coo")

    # ------------------------------------------------------------------ sizes and statistics (tensor.py:346-413)
    def sizes(self) -> List[int]:
        value = self.storage.value()
        extra = [] if value is None else list(value.shape)[1:]
        return list(self.sparse_sizes()) + extra

    def size(self, dim: int) -> int:
        return self.sizes()[dim]

    def dim(self) -> int:
        return len(self.sizes())

    def nnz(self) -> int:
        return self.storage.col().numel()

    def numel(self) -> int:
        value = self.storage.value()
        return self.nnz() if value is None else value.numel()

    def density(self) -> float:
        if self.sparse_size(0) == 0 or self.sparse_size(1) == 0:
            return 0.0
        return self.nnz() / (self.sparse_size(0) * self.sparse_size(1))

    def sparsity(self) -> float:
        return 1 - self.density()

    def avg_row_length(self) -> float:
        return self.nnz() / self.sparse_size(0)

    def avg_col_length(self) -> float:
        return self.nnz() / self.sparse_size(1)

    def bandwidth(self) -> int:
        row, col, _ = self.coo()
        return int((row - col).abs_().max())

    def avg_bandwidth(self) -> float:
        row, col, _ = self.coo()
        return float((row - col).abs_().to(torch.float32).mean())

    def bandwidth_proportion(self, bandwidth: int) -> float:
        row, col, _ = self.coo()
        return int(((row - col).abs_() <= bandwidth).sum()) / self.nnz()

    def is_quadratic(self) -> bool:
        return self.sparse_size(0) == self.sparse_size(1)

    def is_symmetric(self) -> bool:
        if not self.is_quadratic():
            return False
        rowptr, col, value1 = self.csr()
        colptr, row, value2 = self.csc()
        if not torch.equal(rowptr, colptr) or not torch.equal(col, row):
            return False
        if value1 is None or value2 is None:
            return True
        return bool(torch.equal(value1, value2))

    def to_symmetric(self, reduce: str = "sum"):
        """tensor.py:415-451: the entries plus their mirror images on an (N, N) grid, N = max(M, N), equal
        positions reduced.  The reference sorts 2*nnz keys and runs segment_csr over the runs; here the doubled
        triple goes through the constructor unsorted and coalesce(reduce) - one fused sort + dedup-reduce."""
        N = max(self.size(0), self.size(1))
        row, col, value = self.coo()
        new_row = torch.cat([row, col], dim=0)
        new_col = torch.cat([col, row], dim=0)
        if value is not None:
            value = torch.cat([value, value], dim=0)
        out = self.__class__(row=new_row, col=new_col, value=value, sparse_sizes=(N, N), trust_data=True)
        return out.coalesce(reduce)

    # ------------------------------------------------------------------ autograd / device / dtype
    def detach_(self):
        value = self.storage.value()
        if value is not None:
            value.detach_()
        return self

    def detach(self):
        value = self.storage.value()
        return self if value is None else self.set_value(value.detach(), layout="coo")

    def requires_grad(self) -> bool:
        value = self.storage.value()
        return False if value is None else value.requires_grad

    def requires_grad_(self, requires_grad: bool = True, dtype=None):
        """tensor.py:472-483."""
        if requires_grad and not self.has_value():
            self.fill_value_(1.0, dtype)
        value = self.storage.value()
        if value is not None:
            value.requires_grad_(requires_grad)
        return self

    def to_device(self, device, non_blocking: bool = False):
        """tensor.py:232-239 (a CPU target is refused, see `to`)."""
        return self.to(device=device, non_blocking=non_blocking)

    def device_as(self, tensor: torch.Tensor, non_blocking: bool = False):
        """tensor.py:241-242."""
        return self.to_device(tensor.device, non_blocking=non_blocking)

    def cpu(self):
        """tensor.py:692-693 — refused: this package holds no CPU representation (no CPU fallback)."""
        return self.to_device(torch.device("cpu"))

    def pin_memory(self):
        """tensor.py:485-486 pins host arrays; device arrays have nothing to pin."""
        raise RuntimeError("paddle_sparse_b200.SparseTensor is GPU-only: pin_memory() applies to host tensors")

    def is_pinned(self) -> bool:
        return False  # tensor.py:488-489 on device-resident arrays

    def share_memory_(self):
        """tensor.py:597-599 moves host arrays to shared memory; a no-op for device arrays, as in the framework."""
        return self

    def is_shared(self) -> bool:
        return False  # tensor.py:602-603 on device-resident arrays

    def device(self):
        return self.storage.col().device

    def is_cuda(self) -> bool:
        return self.storage.col().is_cuda

    def cuda(self):
        return self  # GPU-only package: a SparseTensor always lives on a CUDA device

    def to(self, *args, **kwargs):
        """tensor.py:606-700: `to(device, dtype, non_blocking)`, `to(dtype, non_blocking)` or `to(other, non_blocking)`.
        The dtype applies to the values; a CUDA device moves every stored array; a CPU device is refused - this
        package has no CPU representation (the reference moves the arrays to the host here)."""
        if not 0 < len(args) + len(kwargs) <= 3:
            raise TypeError("to() expects (device, dtype, non_blocking), (dtype, non_blocking) or (other, non_blocking)")
        if not set(kwargs) <= {"device", "dtype", "blocking", "non_blocking", "other"}:
            raise TypeError(f"to() got unexpected keyword arguments {sorted(set(kwargs))}")
        device, dtype = kwargs.get("device"), kwargs.get("dtype")
        other = kwargs.get("other")
        for a in args:
            if isinstance(a, torch.Tensor):
                other = a
            elif isinstance(a, torch.dtype):
                dtype = a
            elif isinstance(a, (str, torch.device, int)) and not isinstance(a, bool):
                device = a
        if other is not None:
            device, dtype = other.device, other.dtype
        out = self
        if device is not None:
            device = torch.device("cuda", device) if isinstance(device, int) else torch.device(device)
            if device.type != "cuda":
                raise RuntimeError("paddle_sparse_b200.SparseTensor is GPU-only (no CPU representation): "
                                   f"cannot move to {device}")
            if device.index is not None and device != self.device():
                st = self.storage

                def mv(t):
                    return None if t is None else t.to(device)
                out = self.from_storage(SparseStorage(
                    row=mv(st._row), rowptr=mv(st._rowptr), col=mv(st._col), value=mv(st._value),
                    sparse_sizes=st._sparse_sizes, rowcount=mv(st._rowcount), colptr=mv(st._colptr),
                    colcount=mv(st._colcount), csr2csc=mv(st._csr2csc), csc2csr=mv(st._csc2csr), is_sorted=True,
                    trust_data=True))
        return out.type(dtype) if dtype is not None else out

    def dtype(self):
        value = self.storage.value()
        return torch.float32 if value is None else value.dtype  # tensor.py:503-505

    def is_floating_point(self) -> bool:
        value = self.storage.value()
        return True if value is None else value.is_floating_point()

    def bool(self):
        return self.type(torch.bool)

    def byte(self):
        return self.type(torch.uint8)

    def char(self):
        return self.type(torch.int8)

    def short(self):
        return self.type(torch.int16)

    def bfloat16(self):
        return self.type(torch.bfloat16)

    def half(self):
        return self.type(torch.float16)

    def float(self):
        return self.type(torch.float32)

    def double(self):
        return self.type(torch.float64)

    def int(self):
        return self.type(torch.int32)

    def long(self):
        return self.type(torch.int64)

    # ------------------------------------------------------------------ conversions
    def to_dense(self, dtype=None) -> torch.Tensor:
        """tensor.py:543-562."""
        row, col, value = self.coo()
        if value is not None:
            mat = torch.zeros(self.sizes(), dtype=value.dtype, device=self.device())
            mat[row, col] = value
        else:
            mat = torch.zeros(self.sizes(), dtype=dtype or torch.float32, device=self.device())
            mat[row, col] = 1
        return mat

    def __getitem__(self, index: Any):
        """tensor.py:704-757."""
        index = list(index) if isinstance(index, tuple) else [index]
        if len([i for i in index if not isinstance(i, torch.Tensor) and i is Ellipsis]) > 1:
            raise SyntaxError
        dim = 0
        out = self
        while len(index) > 0:
            item = index.pop(0)
            if isinstance(item, (list, tuple)):
                item = torch.as_tensor(item, device=self.device())
            elif hasattr(item, "__array__") and not isinstance(item, torch.Tensor):
                item = torch.as_tensor(item).to(self.device())
            if isinstance(item, int):
                out = out.select(dim, item)
                dim += 1
            elif isinstance(item, slice):
                if item.step is not None:
                    raise ValueError("Step parameter not yet supported.")
                start = 0 if item.start is None else item.start
                start = self.size(dim) + start if start < 0 else start
                stop = self.size(dim) if item.stop is None else item.stop
                stop = self.size(dim) + stop if stop < 0 else stop
                out = out.narrow(dim, start, max(stop - start, 0))
                dim += 1
            elif isinstance(item, torch.Tensor):
                if item.dtype == torch.bool:
                    out = out.masked_select(dim, item)
                    dim += 1
                elif item.dtype == torch.int64:
                    out = out.index_select(dim, item)
                    dim += 1
            elif item is Ellipsis:
                if self.dim() - len(index) < dim:
                    raise SyntaxError
                dim = self.dim() - len(index)
            else:
                raise SyntaxError
        return out

    def __repr__(self) -> str:
        row, col, value = self.coo()
        parts = [f"row={row.tolist() if row.numel() <= 32 else row}", f"col={col.tolist() if col.numel() <= 32 else col}"]
        if value is not None:
            parts.append(f"val={value.tolist() if value.numel() <= 32 else value}")
        parts.append(f"size={tuple(self.sizes())}, nnz={self.nnz()}, density={100 * self.density():.02f}%")
        pad = " " * (len(self.__class__.__name__) + 1)
        return f"{self.__class__.__name__}(" + (",\n" + pad).join(parts) + ")"


__all__ = ["SparseTensor", "SparseStorage", "get_layout"]
