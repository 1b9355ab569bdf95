"""TEST INFRASTRUCTURE ONLY (oracle).

NumPy restatement of the reference's Python op chains for the SparseStorage hot
path, one function per reference routine, each citing the file:line it follows
(paths relative to /root/reference).  ``ind2ptr``/``ptr2ind``/``sample_adj`` are not
re-derived here: they call the reference's own compiled CPU code through
``oracle.native`` (falling back to the plain-C port where oracle/_ref is absent).

Third-party arithmetic that is NOT under /root/reference and is restated from its
documented semantics:
  * paddle_scatter (unpinned git clone, README.md:21-24): segment_csr, gather_csr,
    scatter, scatter_add.  Empty segments / untouched outputs yield 0 for every
    reduce (upstream torch_scatter convention) — "parity unpinned" for that detail.
  * paddle (unpinned nightly): argsort (tie order unspecified -> we use a STABLE
    argsort, a valid instance), cumsum, nonzero, boolean-mask select, gather.

Pinned by tests/test_oracle.py against every golden vector the reference's tests
hold for this path.
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np

from . import native

I64 = np.int64


# =========================================================================== paddle_scatter
def _reduce_name(reduce: str) -> str:
    assert reduce in ("sum", "add", "mean", "min", "max"), reduce
    return "sum" if reduce == "add" else reduce


def _acc_dtype(dtype):
    # float16 inputs accumulate in float32 (the bf16/f16 oracle of SURVEY.md hard-part 6)
    return np.float32 if dtype == np.float16 else dtype


def segment_csr(src: np.ndarray, indptr: np.ndarray, reduce: str = "sum") -> np.ndarray:
    """paddle_scatter.segment_csr(src, indptr, None, reduce) along dim 0.
    Call sites: storage.py:471, reduce.py:51, tensor.py:437."""
    reduce = _reduce_name(reduce)
    indptr = np.asarray(indptr, dtype=I64)
    S = indptr.size - 1
    out = np.zeros((S,) + src.shape[1:], dtype=src.dtype)
    if S == 0:
        return out
    counts = indptr[1:] - indptr[:-1]
    nz = np.nonzero(counts > 0)[0]
    if nz.size == 0:
        return out
    starts = indptr[:-1][nz]
    # reduceat over the non-empty segments only: consecutive non-empty starts delimit
    # exactly the segments because empty ones contribute no elements in between.
    lo, hi = int(indptr[0]), int(indptr[-1])
    body = src[lo:hi].astype(_acc_dtype(src.dtype), copy=False)
    rel = (starts - lo).astype(np.intp)
    if reduce in ("sum", "mean"):
        red = np.add.reduceat(body, rel, axis=0)
        if reduce == "mean":
            c = counts[nz].reshape((-1,) + (1,) * (src.ndim - 1))
            if np.issubdtype(src.dtype, np.integer):
                red = np.trunc(red / c).astype(src.dtype)  # C++ integer division
            else:
                red = red / c.astype(red.dtype)
    elif reduce == "min":
        red = np.minimum.reduceat(body, rel, axis=0)
    else:
        red = np.maximum.reduceat(body, rel, axis=0)
    out[nz] = red.astype(src.dtype)
    return out


def gather_csr(src: np.ndarray, indptr: np.ndarray) -> np.ndarray:
    """paddle_scatter.gather_csr(src, indptr): out[e] = src[s] for e in [indptr[s], indptr[s+1]).
    Call sites: mul.py:16,79, index_select.py:29,65."""
    indptr = np.asarray(indptr, dtype=I64)
    return np.repeat(src, indptr[1:] - indptr[:-1], axis=0)


def scatter(src: np.ndarray, index: np.ndarray, dim_size: int, reduce: str = "sum") -> np.ndarray:
    """paddle_scatter.scatter(src, index, 0, None, dim_size, reduce).  Call site: reduce.py:42."""
    reduce = _reduce_name(reduce)
    out = np.zeros((dim_size,) + src.shape[1:], dtype=_acc_dtype(src.dtype))
    if src.shape[0] == 0:
        return out.astype(src.dtype)
    if reduce in ("sum", "mean"):
        np.add.at(out, index, src.astype(out.dtype, copy=False))
        if reduce == "mean":
            cnt = np.bincount(index, minlength=dim_size).astype(I64)
            cnt = np.maximum(cnt, 1).reshape((-1,) + (1,) * (src.ndim - 1))
            if np.issubdtype(src.dtype, np.integer):
                out = np.floor_divide(out, cnt)  # rounding_mode="floor" upstream
            else:
                out = out / cnt.astype(out.dtype)
    else:
        touched = np.zeros(dim_size, dtype=bool)
        touched[index] = True
        if reduce == "min":
            init = np.inf if np.issubdtype(out.dtype, np.floating) else np.iinfo(out.dtype).max
            out[...] = init
            np.minimum.at(out, index, src)
        else:
            init = -np.inf if np.issubdtype(out.dtype, np.floating) else np.iinfo(out.dtype).min
            out[...] = init
            np.maximum.at(out, index, src)
        out[~touched] = 0
    return out.astype(src.dtype)


def scatter_add(src: np.ndarray, index: np.ndarray, dim_size: int) -> np.ndarray:
    """paddle_scatter.scatter_add.  Call site: storage.py:414-418 (histogram of col)."""
    out = np.zeros(dim_size, dtype=src.dtype)
    np.add.at(out, index, src)
    return out


# =========================================================================== utils.py
def index_sort(keys: np.ndarray) -> np.ndarray:
    """utils.py:14-23 -> Tensor.argsort() (pyg_lib absent); only perm is consumed."""
    return np.argsort(keys, kind="stable").astype(I64)


# =========================================================================== storage.py
class SparseStorage:
    """Restates paddle_sparse/storage.py:31-521 on NumPy arrays."""

    def __init__(self, row=None, rowptr=None, col=None, value=None, sparse_sizes=None,
                 rowcount=None, colptr=None, colcount=None, csr2csc=None, csc2csr=None,
                 is_sorted: bool = False, trust_data: bool = False):
        assert row is not None or rowptr is not None
        assert col is not None
        col = np.ascontiguousarray(col)
        assert col.dtype == I64 and col.ndim == 1

        # storage.py:65-89 — infer / validate the sparse sizes
        M = 0
        if sparse_sizes is None or sparse_sizes[0] is None:
            if rowptr is not None:
                M = rowptr.size - 1
            elif row is not None and row.size > 0:
                M = int(row.max()) + 1
        else:
            M = sparse_sizes[0]
            if rowptr is not None:
                assert rowptr.size - 1 == M
            elif row is not None and row.size > 0:
                assert trust_data or int(row.max()) < M
        N = 0
        if sparse_sizes is None or sparse_sizes[1] is None:
            if col.size > 0:
                N = int(col.max()) + 1
        else:
            N = sparse_sizes[1]
            if col.size > 0:
                assert trust_data or int(col.max()) < N

        if row is not None:
            row = np.ascontiguousarray(row)
            assert row.dtype == I64 and row.ndim == 1 and row.size == col.size
        if rowptr is not None:
            rowptr = np.ascontiguousarray(rowptr)
            assert rowptr.dtype == I64 and rowptr.size - 1 == M
        if value is not None:
            value = np.ascontiguousarray(value)
            assert value.shape[0] == col.shape[0]
        for name, arr, n in (("rowcount", rowcount, M), ("colcount", colcount, N)):
            if arr is not None:
                assert arr.dtype == I64 and arr.size == n, name
        if colptr is not None:
            assert colptr.dtype == I64 and colptr.size - 1 == N
        for arr in (csr2csc, csc2csr):
            if arr is not None:
                assert arr.dtype == I64 and arr.size == col.size

        self._row, self._rowptr, self._col, self._value = row, rowptr, col, value
        self._sparse_sizes = (M, N)
        self._rowcount, self._colptr, self._colcount = rowcount, colptr, colcount
        self._csr2csc, self._csc2csr = csr2csc, csc2csr

        # storage.py:158-171 — sort on construct
        if not is_sorted and col.size > 0:
            idx = np.zeros(col.size + 1, dtype=I64)
            idx[1:] = self.row()
            idx[1:] *= N
            idx[1:] += col
            if (idx[1:] < idx[:-1]).any():
                perm = index_sort(idx[1:])
                self._row = self.row()[perm]
                self._col = col[perm]
                if value is not None:
                    self._value = value[perm]
                self._csr2csc = None
                self._csc2csr = None

    # -- accessors ---------------------------------------------------------
    def row(self):  # storage.py:195-206
        if self._row is None:
            self._row = native.ptr2ind(self._rowptr, self._col.size)
        return self._row

    def rowptr(self):  # storage.py:211-222
        if self._rowptr is None:
            self._rowptr = native.ind2ptr(self._row, self._sparse_sizes[0])
        return self._rowptr

    def col(self):
        return self._col

    def value(self):
        return self._value

    def has_value(self):
        return self._value is not None

    def sparse_resize(self, sparse_sizes):
        """storage.py:281-330."""
        assert len(sparse_sizes) == 2
        nnz = self._col.size
        d0 = sparse_sizes[0] - self._sparse_sizes[0]
        rowcount, rowptr = self._rowcount, self._rowptr
        if d0 > 0:
            if rowptr is not None:
                rowptr = np.concatenate([rowptr, np.full(d0, nnz, dtype=I64)])
            if rowcount is not None:
                rowcount = np.concatenate([rowcount, np.zeros(d0, dtype=I64)])
        elif d0 < 0:
            rowptr = None if rowptr is None else rowptr[:d0]
            rowcount = None if rowcount is None else rowcount[:d0]
        d1 = sparse_sizes[1] - self._sparse_sizes[1]
        colcount, colptr = self._colcount, self._colptr
        if d1 > 0:
            if colptr is not None:
                colptr = np.concatenate([colptr, np.full(d1, nnz, dtype=I64)])
            if colcount is not None:
                colcount = np.concatenate([colcount, np.zeros(d1, dtype=I64)])
        elif d1 < 0:
            colptr = None if colptr is None else colptr[:d1]
            colcount = None if colcount is None else colcount[:d1]
        return SparseStorage(row=self._row, rowptr=rowptr, col=self._col, value=self._value,
                             sparse_sizes=tuple(sparse_sizes), rowcount=rowcount, colptr=colptr, colcount=colcount,
                             csr2csc=self._csr2csc, csc2csr=self._csc2csr, is_sorted=True, trust_data=True)

    def sparse_reshape(self, num_rows: int, num_cols: int):
        """storage.py:332-368."""
        assert num_rows > 0 or num_rows == -1
        assert num_cols > 0 or num_cols == -1
        assert num_rows > 0 or num_cols > 0
        total = self._sparse_sizes[0] * self._sparse_sizes[1]
        if num_rows == -1:
            num_rows = total // num_cols
        if num_cols == -1:
            num_cols = total // num_rows
        assert num_rows * num_cols == total
        idx = self._sparse_sizes[1] * self.row() + self._col
        return SparseStorage(row=idx // num_cols, col=idx % num_cols, value=self._value,
                             sparse_sizes=(num_rows, num_cols), is_sorted=True, trust_data=True)

    def sparse_sizes(self):
        return self._sparse_sizes

    def sparse_size(self, dim):
        return self._sparse_sizes[dim]

    def set_value_(self, value, layout="coo"):  # storage.py:233-246
        if value is not None and layout == "csc":
            value = value[self.csc2csr()]
        self._value = None if value is None else np.ascontiguousarray(value)
        return self

    def set_value(self, value, layout="coo"):  # storage.py:248-273
        if value is not None and layout == "csc":
            value = value[self.csc2csr()]
        return SparseStorage(row=self._row, rowptr=self._rowptr, col=self._col, value=value,
                             sparse_sizes=self._sparse_sizes, rowcount=self._rowcount,
                             colptr=self._colptr, colcount=self._colcount,
                             csr2csc=self._csr2csc, csc2csr=self._csc2csr,
                             is_sorted=True, trust_data=True)

    def rowcount(self):  # storage.py:373-381
        if self._rowcount is None:
            rp = self.rowptr()
            self._rowcount = rp[1:] - rp[:-1]
        return self._rowcount

    def colptr(self):  # storage.py:386-400
        if self._colptr is None:
            if self._csr2csc is not None:
                self._colptr = native.ind2ptr(self._col[self._csr2csc], self._sparse_sizes[1])
            else:
                cp = np.zeros(self._sparse_sizes[1] + 1, dtype=I64)
                cp[1:] = np.cumsum(self.colcount())
                self._colptr = cp
        return self._colptr

    def colcount(self):  # storage.py:405-420
        if self._colcount is None:
            if self._colptr is not None:
                self._colcount = self._colptr[1:] - self._colptr[:-1]
            else:
                self._colcount = scatter_add(np.ones_like(self._col), self._col,
                                             self._sparse_sizes[1])
        return self._colcount

    def csr2csc(self):  # storage.py:425-434
        if self._csr2csc is None:
            idx = self._sparse_sizes[0] * self._col + self.row()
            self._csr2csc = index_sort(idx)
        return self._csr2csc

    def csc2csr(self):  # storage.py:439-447
        if self._csc2csr is None:
            self._csc2csr = index_sort(self.csr2csc())
        return self._csc2csr

    def is_coalesced(self) -> bool:  # storage.py:449-452
        idx = np.full(self._col.size + 1, -1, dtype=I64)
        idx[1:] = self._sparse_sizes[1] * self.row() + self._col
        return bool((idx[1:] > idx[:-1]).all())

    def coalesce(self, reduce: str = "add"):  # storage.py:454-486
        idx = np.full(self._col.size + 1, -1, dtype=I64)
        idx[1:] = self._sparse_sizes[1] * self.row() + self._col
        mask = idx[1:] > idx[:-1]
        if mask.all():
            return self
        row, col = self.row()[mask], self._col[mask]
        value = self._value
        if value is not None:
            ptr = np.concatenate([np.nonzero(mask)[0].astype(I64),
                                  np.array([value.shape[0]], dtype=I64)])
            value = segment_csr(value, ptr, reduce)
        return SparseStorage(row=row, col=col, value=value, sparse_sizes=self._sparse_sizes,
                             is_sorted=True, trust_data=True)

    def fill_cache_(self):  # storage.py:488-496
        self.row(); self.rowptr(); self.rowcount(); self.colptr(); self.colcount()
        self.csr2csc(); self.csc2csr()
        return self

    def clear_cache_(self):  # storage.py:498-504
        self._rowcount = self._colptr = self._colcount = None
        self._csr2csc = self._csc2csr = None
        return self

    def cached_keys(self):  # storage.py:506-518
        return [k for k in ("rowcount", "colptr", "colcount", "csr2csc", "csc2csr")
                if getattr(self, "_" + k) is not None]

    def num_cached_keys(self):
        return len(self.cached_keys())


# =========================================================================== tensor.py (facade)
class SparseTensor:
    """The slice of paddle_sparse/tensor.py the hot path touches (:20-64, :246-257)."""

    def __init__(self, row=None, rowptr=None, col=None, value=None, sparse_sizes=None,
                 is_sorted=False, trust_data=False):
        self.storage = SparseStorage(row=row, rowptr=rowptr, col=col, value=value,
                                     sparse_sizes=sparse_sizes, is_sorted=is_sorted,
                                     trust_data=trust_data)

    @classmethod
    def from_storage(cls, st: SparseStorage):  # tensor.py:48-64
        out = cls(row=st._row, rowptr=st._rowptr, col=st._col, value=st._value,
                  sparse_sizes=st._sparse_sizes, is_sorted=True, trust_data=True)
        for k in ("_rowcount", "_colptr", "_colcount", "_csr2csc", "_csc2csr"):
            setattr(out.storage, k, getattr(st, k))
        return out

    def coo(self):
        return self.storage.row(), self.storage.col(), self.storage.value()

    def csr(self):
        return self.storage.rowptr(), self.storage.col(), self.storage.value()

    def csc(self):  # tensor.py:252-257
        perm = self.storage.csr2csc()
        value = self.storage.value()
        if value is not None:
            value = value[perm]
        return self.storage.colptr(), self.storage.row()[perm], value

    def sparse_sizes(self):
        return self.storage.sparse_sizes()

    def sparse_size(self, d):
        return self.storage.sparse_sizes()[d]

    def sizes(self):
        v = self.storage.value()
        return list(self.sparse_sizes()) + (list(v.shape[1:]) if v is not None else [])

    def size(self, d):
        return self.sizes()[d]

    def dim(self):
        return len(self.sizes())

    def nnz(self):
        return self.storage.col().size

    def dtype(self):  # tensor.py:503-505
        v = self.storage.value()
        return v.dtype if v is not None else np.dtype(np.float32)

    def set_value(self, value, layout="coo"):
        return self.from_storage(self.storage.set_value(value, layout))

    def set_value_(self, value, layout="coo"):
        self.storage.set_value_(value, layout)
        return self

    def coalesce(self, reduce="sum"):
        return self.from_storage(self.storage.coalesce(reduce))

    def is_coalesced(self):
        return self.storage.is_coalesced()

    def fill_cache_(self):
        self.storage.fill_cache_()
        return self

    def to_dense(self):
        row, col, value = self.coo()
        v = value if value is not None else np.ones(self.nnz(), dtype=np.float32)
        mat = np.zeros(tuple(self.sparse_sizes()) + v.shape[1:], dtype=v.dtype)
        mat[row, col] = v
        return mat


# =========================================================================== coalesce.py / transpose.py
def coalesce(index, value, m, n, op="add"):
    """coalesce.py:8-29."""
    st = SparseStorage(row=index[0], col=index[1], value=value, sparse_sizes=(m, n),
                       is_sorted=False)
    st = st.coalesce(reduce=op)
    return np.stack([st.row(), st.col()], axis=0), st.value()


def t(src: SparseTensor) -> SparseTensor:
    """transpose.py:9-33."""
    csr2csc = src.storage.csr2csc()
    row, col, value = src.coo()
    if value is not None:
        value = value[csr2csc]
    M, N = src.storage.sparse_sizes()
    st = SparseStorage(row=col[csr2csc], rowptr=src.storage._colptr, col=row[csr2csc],
                       value=value, sparse_sizes=(N, M), rowcount=src.storage._colcount,
                       colptr=src.storage._rowptr, colcount=src.storage._rowcount,
                       csr2csc=src.storage._csc2csr, csc2csr=csr2csc, is_sorted=True)
    return SparseTensor.from_storage(st)


def transpose(index, value, m, n, coalesced=True):
    """transpose.py:41-65."""
    row, col = index[1], index[0]
    if coalesced:
        st = SparseStorage(row=row, col=col, value=value, sparse_sizes=(n, m), is_sorted=False)
        st = st.coalesce()
        row, col, value = st.row(), st.col(), st.value()
    return np.stack([row, col], axis=0), value


# =========================================================================== reduce.py
def reduction(src: SparseTensor, dim: Optional[int] = None, reduce: str = "sum"):
    """reduce.py:12-71."""
    value = src.storage.value()
    if dim is None:
        if value is not None:
            r = _reduce_name(reduce)
            return {"sum": value.sum, "mean": value.mean, "min": value.min, "max": value.max}[r]()
        return np.array(src.nnz() if _reduce_name(reduce) == "sum" else 1, dtype=src.dtype())
    if dim < 0:
        dim = src.dim() + dim
    r = _reduce_name(reduce)
    if dim == 0 and value is not None:
        return scatter(value, src.storage.col(), src.size(1), reduce)
    if dim == 0:
        if r == "sum":
            return src.storage.colcount().astype(src.dtype())
        return np.ones(src.size(1), dtype=src.dtype())
    if dim == 1 and value is not None:
        return segment_csr(value, src.storage.rowptr(), reduce)
    if dim == 1:
        if r == "sum":
            return src.storage.rowcount().astype(src.dtype())
        return np.ones(src.size(0), dtype=src.dtype())
    if value is not None:
        f = {"sum": np.sum, "mean": np.mean, "min": np.min, "max": np.max}[r]
        return f(value, axis=dim - 1)
    raise ValueError


# =========================================================================== mul.py
def mul_(src: SparseTensor, other: np.ndarray) -> SparseTensor:
    """mul.py:76-93 (the out-of-place `mul` :12-33 intends the same arithmetic, see
    SURVEY.md 3.4)."""
    rowptr, col, value = src.csr()
    if other.shape[0] == src.size(0) and other.shape[1] == 1:
        other = gather_csr(other.squeeze(1), rowptr)
    elif other.shape[0] == 1 and other.shape[1] == src.size(1):
        other = other.squeeze(0)[col]
    else:
        raise ValueError("Size mismatch")
    if value is not None:
        value = value * other.astype(value.dtype)
    else:
        value = other
    return src.set_value_(value, layout="coo")


def mul(src: SparseTensor, other: np.ndarray) -> SparseTensor:
    out = SparseTensor.from_storage(src.storage)
    if out.storage._value is not None:
        out.storage._value = out.storage._value.copy()
    return mul_(out, other)


def mul_sparse(src: SparseTensor, other: SparseTensor) -> SparseTensor:
    """mul.py:35-73, sparse (.) sparse: both operands coalesced; concatenate, sort the keys
    row*N+col behind a -1 sentinel, keep the second entry of every adjacent equal pair and
    multiply it with its predecessor."""
    if not src.is_coalesced():
        raise ValueError("The `src` tensor is not coalesced")
    if not other.is_coalesced():
        raise ValueError("The `other` tensor is not coalesced")
    rowA, colA, valueA = src.coo()
    rowB, colB, valueB = other.coo()
    row = np.concatenate([rowA, rowB])
    col = np.concatenate([colA, colB])
    if valueA is None or valueB is None:
        raise ValueError("Both sparse tensors must contain values")
    value = np.concatenate([valueA, valueB])
    M = max(src.size(0), other.size(0))
    N = max(src.size(1), other.size(1))
    idx = np.full(col.size + 1, -1, dtype=np.int64)
    idx[1:] = row * N + col
    perm = np.argsort(idx[1:], kind="stable")
    idx[1:] = idx[1:][perm]
    row, col, value = row[perm], col[perm], value[perm]
    valid_mask = idx[1:] == idx[:-1]
    valid_idx = np.nonzero(valid_mask)[0]
    return SparseTensor(row=row[valid_mask], col=col[valid_mask],
                        value=value[valid_idx - 1] * value[valid_idx], sparse_sizes=(M, N))


# =========================================================================== add.py
def add_(src: SparseTensor, other: np.ndarray) -> SparseTensor:
    """add.py:53-69 (dense [M,1] / [1,N] operand; the out-of-place `add` :12-28 computes the same)."""
    rowptr, col, value = src.csr()
    if other.shape[0] == src.size(0) and other.shape[1] == 1:
        other = gather_csr(other.squeeze(1), rowptr)
    elif other.shape[0] == 1 and other.shape[1] == src.size(1):
        other = other.squeeze(0)[col]
    else:
        raise ValueError("Size mismatch")
    if value is not None:
        value = value + other.astype(value.dtype)
    else:
        value = other + np.ones((), dtype=other.dtype)
    return src.set_value_(value, layout="coo")


def add(src: SparseTensor, other) -> SparseTensor:
    """add.py:12-50.  Sparse + sparse (:30-47): concatenate the COO triples, construct on the
    element-wise maximum of the sizes, coalesce with "sum"; the value is dropped unless both have one."""
    if isinstance(other, SparseTensor):
        rowA, colA, valueA = src.coo()
        rowB, colB, valueB = other.coo()
        row = np.concatenate([rowA, rowB])
        col = np.concatenate([colA, colB])
        value = None
        if valueA is not None and valueB is not None:
            value = np.concatenate([valueA, valueB])
        M = max(src.size(0), other.size(0))
        N = max(src.size(1), other.size(1))
        return SparseTensor(row=row, col=col, value=value, sparse_sizes=(M, N)).coalesce(reduce="sum")
    out = SparseTensor.from_storage(src.storage)
    if out.storage._value is not None:
        out.storage._value = out.storage._value.copy()
    return add_(out, other)


# =========================================================================== tensor.py (symmetry)
def is_symmetric(src: SparseTensor) -> bool:
    """tensor.py:400-413."""
    if src.sparse_size(0) != src.sparse_size(1):
        return False
    rowptr, col, value1 = src.csr()
    colptr, row, value2 = src.csc()
    if (rowptr != colptr).any() or (col != row).any():
        return False
    if value1 is None or value2 is None:
        return True
    return bool((value1 == value2).all())


def to_symmetric(src: SparseTensor, reduce: str = "sum") -> SparseTensor:
    """tensor.py:415-451: keys of [row;col]*N + [col;row] behind a -1 sentinel, sorted; a run starts
    where the sorted key grows; values of a run are reduced with segment_csr; the index of a run is taken
    from its first element.  Result is (N, N) with N = max(M, N)."""
    N = max(src.size(0), src.size(1))
    row, col, value = src.coo()
    n = row.size
    idx = np.full(2 * n + 1, -1, dtype=np.int64)
    idx[1:n + 1] = row
    idx[n + 1:] = col
    idx[1:] *= N
    idx[1:n + 1] += col
    idx[n + 1:] += row
    perm = np.argsort(idx, kind="stable")
    idx = idx[perm]
    mask = idx[1:] > idx[:-1]
    perm = perm[1:] - 1
    sel = perm[mask]
    if value is not None:
        ptr = np.concatenate([np.nonzero(mask)[0], np.array([perm.size], dtype=np.int64)])
        value = segment_csr(np.concatenate([value, value])[perm], ptr, reduce)
    new_row = np.concatenate([row, col])[sel]
    new_col = np.concatenate([col, row])[sel]
    return SparseTensor(row=new_row, col=new_col, value=value, sparse_sizes=(N, N),
                        is_sorted=True, trust_data=True)


# =========================================================================== cat.py
def _cat_pieces(parts, shifts, pointer=False):
    """`concat([x_k + shift_k])`, None unless every operand has the piece; pointers drop their leading entry
    after the first operand (cat.py:63, :137, :201, :216)."""
    if any(p is None for p in parts):
        return None
    out = []
    for k, (p, sh) in enumerate(zip(parts, shifts)):
        out.append((p[1:] if pointer and k > 0 else p) + sh)
    return np.concatenate(out)


def cat(tensors, dim):
    """cat.py:12-276: dim 0 (cat_first :49-112), dim 1 (cat_second :115-172), dim (0, 1) (cat_diag :175-276),
    or a dense value dimension (:25-32)."""
    assert len(tensors) > 0
    st = [t.storage for t in tensors]
    rows_off = np.concatenate([[0], np.cumsum([t.sparse_size(0) for t in tensors])]).tolist()
    cols_off = np.concatenate([[0], np.cumsum([t.sparse_size(1) for t in tensors])]).tolist()
    nnz_off = np.concatenate([[0], np.cumsum([t.nnz() for t in tensors])]).tolist()
    zeros = [0] * len(tensors)
    vals = [s._value for s in st]
    value = None if any(v is None for v in vals) else np.concatenate(vals)
    if isinstance(dim, int):
        dim = tensors[0].dim() + dim if dim < 0 else dim
        if dim == 0:
            storage = SparseStorage(
                row=_cat_pieces([s._row for s in st], rows_off), rowptr=_cat_pieces([s._rowptr for s in st], nnz_off, True),
                col=np.concatenate([s._col for s in st]), value=value,
                sparse_sizes=(rows_off[-1], max(t.sparse_size(1) for t in tensors)),
                rowcount=_cat_pieces([s._rowcount for s in st], zeros), is_sorted=True)
            return SparseTensor.from_storage(storage)
        if dim == 1:
            storage = SparseStorage(
                row=np.concatenate([t.coo()[0] for t in tensors]), rowptr=None,
                col=_cat_pieces([s._col for s in st], cols_off), value=value,
                sparse_sizes=(max(t.sparse_size(0) for t in tensors), cols_off[-1]),
                colptr=_cat_pieces([s._colptr for s in st], nnz_off, True),
                colcount=_cat_pieces([s._colcount for s in st], zeros), is_sorted=False)
            return SparseTensor.from_storage(storage)
        if 1 < dim < tensors[0].dim():
            values = [t.storage.value() for t in tensors]
            assert all(v is not None for v in values)
            return tensors[0].set_value(np.concatenate(values, axis=dim - 1), layout="coo")
        raise IndexError("Dimension out of range")
    assert isinstance(dim, (tuple, list)) and len(dim) == 2 and sorted(dim) == [0, 1]
    storage = SparseStorage(
        row=_cat_pieces([s._row for s in st], rows_off), rowptr=_cat_pieces([s._rowptr for s in st], nnz_off, True),
        col=_cat_pieces([s._col for s in st], cols_off), value=value, sparse_sizes=(rows_off[-1], cols_off[-1]),
        rowcount=_cat_pieces([s._rowcount for s in st], zeros),
        colptr=_cat_pieces([s._colptr for s in st], nnz_off, True),
        colcount=_cat_pieces([s._colcount for s in st], zeros),
        csr2csc=_cat_pieces([s._csr2csc for s in st], nnz_off), csc2csr=_cat_pieces([s._csc2csr for s in st], nnz_off),
        is_sorted=True)
    return SparseTensor.from_storage(storage)


# =========================================================================== narrow.py / select.py
def narrow(src: SparseTensor, dim: int, start: int, length: int) -> SparseTensor:
    """narrow.py:11-100.  No reference test exercises it (parity unpinned): this restatement follows the source
    line by line and is cross-checked against index_select over an arange, which the golden vectors do pin."""
    if dim < 0:
        dim = src.dim() + dim
    if start < 0:
        start = src.size(dim) + start
    st = src.storage
    if dim == 0:  # narrow.py:18-51
        rowptr, col, value = src.csr()
        rowptr = rowptr[start:start + length + 1]
        row_start = int(rowptr[0])
        rowptr = rowptr - row_start
        row_length = int(rowptr[-1])
        row = st._row
        if row is not None:
            row = row[row_start:row_start + row_length] - start
        col = col[row_start:row_start + row_length].copy()
        if value is not None:
            value = value[row_start:row_start + row_length].copy()
        rowcount = st._rowcount
        if rowcount is not None:
            rowcount = rowcount[start:start + length].copy()
        storage = SparseStorage(row=row, rowptr=rowptr, col=col, value=value,
                                sparse_sizes=(length, src.sparse_size(1)), rowcount=rowcount, is_sorted=True)
        return SparseTensor.from_storage(storage)
    if dim == 1:  # narrow.py:53-92
        row, col, value = src.coo()
        mask = (col >= start) & (col < start + length)
        row = row[mask]
        col = col[mask] - start
        if value is not None:
            value = value[mask]
        colptr = st._colptr
        if colptr is not None:
            colptr = colptr[start:start + length + 1]
            colptr = colptr - colptr[0]
        colcount = st._colcount
        if colcount is not None:
            colcount = colcount[start:start + length].copy()
        storage = SparseStorage(row=row, rowptr=None, col=col, value=value,
                                sparse_sizes=(src.sparse_size(0), length), colptr=colptr, colcount=colcount,
                                is_sorted=True)
        return SparseTensor.from_storage(storage)
    value = st.value()  # narrow.py:94-100
    if value is None:
        raise ValueError
    index = [slice(None)] * value.ndim
    index[dim - 1] = slice(start, start + length)
    return src.set_value(value[tuple(index)].copy(), layout="coo")


def narrow_diag(src: SparseTensor, start, length) -> SparseTensor:
    """`__narrow_diag__`, narrow.py:103-168: the inverse of cat_diag."""
    st = src.storage
    rowptr, col, value = src.csr()
    rowptr = rowptr[start[0]:start[0] + length[0] + 1]
    row_start = int(rowptr[0])
    rowptr = rowptr - row_start
    row_length = int(rowptr[-1])
    sl = slice(row_start, row_start + row_length)
    row = st._row
    if row is not None:
        row = row[sl] - start[0]
    col = col[sl] - start[1]
    if value is not None:
        value = value[sl].copy()
    rowcount = st._rowcount
    if rowcount is not None:
        rowcount = rowcount[start[0]:start[0] + length[0]].copy()
    colptr = st._colptr
    if colptr is not None:
        colptr = colptr[start[1]:start[1] + length[1] + 1]
        colptr = colptr - int(colptr[0])
    colcount = st._colcount
    if colcount is not None:
        colcount = colcount[start[1]:start[1] + length[1]].copy()
    csr2csc = st._csr2csc
    if csr2csc is not None:
        csr2csc = csr2csc[sl] - row_start
    csc2csr = st._csc2csr
    if csc2csr is not None:
        csc2csr = csc2csr[sl] - row_start
    storage = SparseStorage(row=row, rowptr=rowptr, col=col, value=value, sparse_sizes=tuple(length),
                            rowcount=rowcount, colptr=colptr, colcount=colcount, csr2csc=csr2csc, csc2csr=csc2csr,
                            is_sorted=True)
    return SparseTensor.from_storage(storage)


def select(src: SparseTensor, dim: int, idx: int) -> SparseTensor:  # select.py:7-8
    return narrow(src, dim, start=idx, length=1)


# =========================================================================== index_select.py
def index_select(src: SparseTensor, dim: int, idx: np.ndarray) -> SparseTensor:
    """index_select.py:13-96."""
    dim = src.dim() + dim if dim < 0 else dim
    idx = np.asarray(idx, dtype=I64)
    K = idx.shape[0]
    if dim == 0:
        old_rowptr, col, value = src.csr()
        rowcount = src.storage.rowcount()[idx]
        rowptr = np.zeros(K + 1, dtype=I64)
        rowptr[1:] = np.cumsum(rowcount)
        row = np.repeat(np.arange(K, dtype=I64), rowcount)
        perm = np.arange(row.shape[0], dtype=I64)
        perm += gather_csr(old_rowptr[idx] - rowptr[:-1], rowptr)
        col = col[perm]
        if value is not None:
            value = value[perm]
        st = SparseStorage(row=row, rowptr=rowptr, col=col, value=value,
                           sparse_sizes=(K, src.sparse_size(1)), rowcount=rowcount,
                           is_sorted=True)
        return SparseTensor.from_storage(st)
    if dim == 1:
        old_colptr, row, value = src.csc()
        colcount = src.storage.colcount()[idx]
        colptr = np.zeros(K + 1, dtype=I64)
        colptr[1:] = np.cumsum(colcount)
        col = np.repeat(np.arange(K, dtype=I64), colcount)
        perm = np.arange(col.shape[0], dtype=I64)
        perm += gather_csr(old_colptr[idx] - colptr[:-1], colptr)
        row = row[perm]
        csc2csr = index_sort(K * row + col)
        row, col = row[csc2csr], col[csc2csr]
        if value is not None:
            value = value[perm][csc2csr]
        st = SparseStorage(row=row, col=col, value=value,
                           sparse_sizes=(src.sparse_size(0), K), colptr=colptr,
                           colcount=colcount, csc2csr=csc2csr, is_sorted=True)
        return SparseTensor.from_storage(st)
    value = src.storage.value()
    if value is None:
        raise ValueError
    return src.set_value(np.take(value, idx, axis=dim - 1), layout="coo")


def index_select_nnz(src: SparseTensor, idx: np.ndarray, layout: str = "coo") -> SparseTensor:
    """index_select.py:99-120."""
    if layout == "csc":
        idx = src.storage.csc2csr()[idx]
    row, col, value = src.coo()
    row, col = row[idx], col[idx]
    if value is not None:
        value = value[idx]
    return SparseTensor(row=row, col=col, value=value, sparse_sizes=src.sparse_sizes(),
                        is_sorted=True)


# =========================================================================== masked_select.py
def masked_select(src: SparseTensor, dim: int, mask: np.ndarray) -> SparseTensor:
    """masked_select.py:12-90."""
    dim = src.dim() + dim if dim < 0 else dim
    mask = np.asarray(mask, dtype=bool)
    if dim == 0:
        row, col, value = src.coo()
        rowcount = src.storage.rowcount()[mask]
        emask = mask[row]
        row = np.repeat(np.arange(rowcount.shape[0], dtype=I64), rowcount)
        col = col[emask]
        if value is not None:
            value = value[emask]
        st = SparseStorage(row=row, col=col, value=value,
                           sparse_sizes=(rowcount.shape[0], src.sparse_size(1)),
                           rowcount=rowcount, is_sorted=True)
        return SparseTensor.from_storage(st)
    if dim == 1:
        row, col, value = src.coo()
        csr2csc = src.storage.csr2csc()
        row, col = row[csr2csc], col[csr2csc]
        colcount = src.storage.colcount()[mask]
        emask = mask[col]
        K = colcount.shape[0]
        col = np.repeat(np.arange(K, dtype=I64), colcount)
        row = row[emask]
        csc2csr = index_sort(K * row + col)
        row, col = row[csc2csr], col[csc2csr]
        if value is not None:
            value = value[csr2csc][emask][csc2csr]
        st = SparseStorage(row=row, col=col, value=value,
                           sparse_sizes=(src.sparse_size(0), K), colcount=colcount,
                           csc2csr=csc2csr, is_sorted=True)
        return SparseTensor.from_storage(st)
    value = src.storage.value()
    if value is None:
        raise ValueError
    return src.set_value(np.take(value, np.nonzero(mask)[0], axis=dim - 1), layout="coo")


def masked_select_nnz(src: SparseTensor, mask: np.ndarray, layout: str = "coo") -> SparseTensor:
    """masked_select.py:93-114."""
    mask = np.asarray(mask, dtype=bool)
    if layout == "csc":
        mask = mask[src.storage.csc2csr()]
    row, col, value = src.coo()
    row, col = row[mask], col[mask]
    if value is not None:
        value = value[mask]
    return SparseTensor(row=row, col=col, value=value, sparse_sizes=src.sparse_sizes(),
                        is_sorted=True)


# =========================================================================== sample.py
def sample_adj(src: SparseTensor, subset: np.ndarray, num_neighbors: int,
               replace: bool = False, seed: int = 0, which: str = "port"
               ) -> Tuple[SparseTensor, np.ndarray]:
    """sample.py:31-53.  which="ref" runs the reference's own sample_adj_cpu,
    which="port" the plain-C restatement with the fixed n_id order."""
    rowptr, col, value = src.csr()
    fn = native.sample_adj_ref if which == "ref" else native.sample_adj_port
    out_rowptr, out_col, n_id, e_id = fn(rowptr, col, subset, num_neighbors, replace, seed)
    if value is not None:
        value = value[e_id]
    out = SparseTensor(rowptr=out_rowptr, col=out_col, value=value,
                       sparse_sizes=(len(subset), n_id.size), is_sorted=True)
    return out, n_id


def sample(src: SparseTensor, num_neighbors: int, subset: Optional[np.ndarray] = None, seed: int = 0) -> np.ndarray:
    """sample.py:10-28: rand([S, k]) * rowcount -> int64 -> + rowptr -> col[...].  The reference draws from the
    framework's generator; the draw of (position i, draw k) is pinned here as the top 24 bits of
    Philox4x32-10(seed; (i, k)) * 2^-24 (a float32 in [0, 1)), the mapping ps_sample_uniform uses; the product in
    float32, truncated, clamped to the row (the reference would index past it when rand * count rounds up)."""
    rowptr, col, _ = src.csr()
    rows = np.arange(rowptr.size - 1) if subset is None else np.asarray(subset)
    out = np.empty((rows.size, num_neighbors), dtype=np.int64)
    for i, r in enumerate(rows):
        lo, deg = int(rowptr[r]), int(rowptr[r + 1] - rowptr[r])
        for k in range(num_neighbors):
            u = np.float32(native.philox_u64(seed, i, k) >> 40) * np.float32(1.0 / 16777216.0)
            off = int(np.float32(u) * np.float32(deg))
            if off >= deg:
                off = max(deg - 1, 0)
            out[i, k] = col[min(lo + off, col.size - 1)]
    return out


# =========================================================================== helpers for tests/bench
def round_bf16(x: np.ndarray) -> np.ndarray:
    """Round-to-nearest-even float32 -> bfloat16 -> float32 (NumPy has no bf16)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    u = x.view(np.uint32).astype(np.uint64)
    rounded = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16) << 16
    return rounded.astype(np.uint32).view(np.float32).reshape(x.shape)


def build_pipeline(row, col, value, M, N):
    """The BASELINE.json headline pipeline on the CPU: coalesce -> rowptr -> csr2csc ->
    colptr -> csc2csr -> t().  Returns a dict of every array the GPU path must match."""
    idx, val = coalesce(np.stack([row, col]), value, M, N, "add")
    A = SparseTensor(row=idx[0], col=idx[1], value=val, sparse_sizes=(M, N),
                     is_sorted=True, trust_data=True)
    st = A.storage
    out = dict(row=st.row(), col=st.col(), value=st.value(), rowptr=st.rowptr().copy())
    out["csr2csc"] = st.csr2csc().copy()
    out["colptr"] = st.colptr().copy()
    out["csc2csr"] = st.csc2csr().copy()
    st.rowcount(); st.colcount()
    At = t(A)
    out["t_row"], out["t_col"], out["t_value"] = At.coo()
    return out
