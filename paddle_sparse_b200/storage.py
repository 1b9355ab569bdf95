"""SparseStorage — same public surface as the reference's paddle_sparse/storage.py:31-762,
rebuilt on the fused sm_100a kernels of libpaddle_sparse_b200.so.

What differs underneath (never in the API):
  * construction validates with ONE kernel + one sync (`ops.index_max`) instead of two
    `int(x.max())` round trips (storage.py:65-89), and decides sortedness with one
    boundary kernel (`ops.coo_order`) instead of a 4-pass key build (storage.py:158-163);
  * sort-on-construct is one packed-key radix sort that carries the value as payload
    (storage.py:165-169: argsort + three gathers);
  * `coalesce` is one dedup + segmented-reduce kernel (storage.py:454-486: key rebuild,
    two mask selects, nonzero, segment_csr);
  * `csr2csc` is a stable radix sort on the column bits only and produces `colptr` /
    `colcount` as by-products (storage.py:386-434: a second 64-bit argsort, a histogram, a
    cumsum); `csc2csr` is one inverse-permutation scatter (storage.py:439-447: a third sort).
Caches stay lazy and observable exactly as in the reference (`cached_keys`, test_storage.py:53-108).
"""
from __future__ import annotations

import warnings
from typing import List, Optional, Tuple

import torch

from . import ops

layouts: List[str] = ["coo", "csr", "csc"]


def get_layout(layout: Optional[str] = None) -> str:
    """storage.py:20-28."""
    if layout is None:
        layout = "coo"
        warnings.warn('`layout` argument unset, using default layout "coo". This may lead to unexpected behaviour.')
    assert layout in ("coo", "csr", "csc")
    return layout


def _check_index(t: torch.Tensor, like: torch.Tensor, numel: Optional[int] = None) -> torch.Tensor:
    assert t.dtype == torch.int64
    assert t.device == like.device
    assert t.dim() == 1
    if numel is not None:
        assert t.numel() == numel
    return t.contiguous()


class SparseStorage(object):
    def __init__(self, row=None, rowptr=None, col=None, value=None, sparse_sizes=None, rowcount=None,
                 colptr=None, colcount=None, csr2csc=None, csc2csr=None, is_sorted: bool = False,
                 trust_data: bool = False):
        assert row is not None or rowptr is not None
        assert col is not None
        assert col.dtype == torch.int64
        assert col.dim() == 1
        if not col.is_cuda and not (is_sorted and trust_data):
            # Host tensors are accepted only as an inert holder (`pin_memory()` staging copies of an already built
            # storage, storage.py:693-733): nothing is validated, sorted or computed for them, and every operator
            # raises on a CPU tensor.  There is no CPU compute path.
            raise RuntimeError("paddle_sparse_b200.SparseStorage is GPU-only (no CPU fallback): `col` is on the CPU")
        col = col.contiguous()
        nnz = col.numel()

        # ---- sizes: one fused max kernel and one sync where the reference has two (storage.py:65-89)
        need_m = rowptr is None and row is not None and row.numel() > 0 and (
            sparse_sizes is None or sparse_sizes[0] is None or not trust_data)
        need_n = nnz > 0 and (sparse_sizes is None or sparse_sizes[1] is None or not trust_data)
        row_max = col_max = -1
        inspection = None
        if need_m or need_n:
            if (need_m and need_n and not is_sorted and sparse_sizes is not None and sparse_sizes[0] is not None
                    and sparse_sizes[1] is not None and nnz >= (1 << 16)):
                # the validation read doubles as the first read of the coalesce / sort that an unsorted input owes
                row_max, col_max, inspection = ops.coo_inspect(row, col, int(sparse_sizes[0]), int(sparse_sizes[1]))
            elif need_m and need_n:
                row_max, col_max = ops.index_max(row, col)
            elif need_m:
                row_max, _ = ops.index_max(row)
            else:
                col_max, _ = ops.index_max(col)

        M = 0
        if sparse_sizes is None or sparse_sizes[0] is None:
            if rowptr is not None:
                M = rowptr.numel() - 1
            elif need_m:
                M = row_max + 1
        else:
            M = int(sparse_sizes[0])
            if rowptr is not None:
                assert rowptr.numel() - 1 == M
            elif need_m:
                assert row_max < M

        N = 0
        if sparse_sizes is None or sparse_sizes[1] is None:
            if need_n:
                N = col_max + 1
        else:
            N = int(sparse_sizes[1])
            if need_n:
                assert col_max < N

        if row is not None:
            row = _check_index(row, col, nnz)
        if rowptr is not None:
            rowptr = _check_index(rowptr, col, M + 1)
        if value is not None:
            assert value.device == col.device
            assert value.shape[0] == nnz
            value = value.contiguous()
        if rowcount is not None:
            rowcount = _check_index(rowcount, col, M)
        if colptr is not None:
            colptr = _check_index(colptr, col, N + 1)
        if colcount is not None:
            colcount = _check_index(colcount, col, N)
        if csr2csc is not None:
            csr2csc = _check_index(csr2csc, col, nnz)
        if csc2csr is not None:
            csc2csr = _check_index(csc2csr, col, nnz)

        self._row = row
        self._rowptr = rowptr
        self._col = col
        self._value = value
        self._sparse_sizes = (M, N)
        self._rowcount = rowcount
        self._colptr = colptr
        self._colcount = colcount
        self._csr2csc = csr2csc
        self._csc2csr = csc2csr

        # ---- sort-on-construct (storage.py:158-171), DEFERRED to the first access of row / col / value: a
        # `coalesce()` that follows the constructor then runs sort + dedup + rowptr as one fused pipeline
        # (ops.coalesce) instead of sort -> write -> re-read -> dedup.  Every accessor sees sorted data.
        self._pending_sort = (not is_sorted) and nnz > 0
        # the sort is deferred, so `row` / `col` are still the caller's tensors: remember their version counters and
        # validate again at materialisation time if they were edited in place in between
        d0 = self.__dict__
        self._seen = (None if d0["_row_raw"] is None else d0["_row_raw"]._version, d0["_col_raw"]._version,
                      not trust_data)
        # what ops.coo_inspect left on the device, valid while row / col are the tensors it read, unedited
        self._inspection = None
        if inspection is not None and self._pending_sort:
            d = self.__dict__
            self._inspection = inspection + (d["_row_raw"], d["_row_raw"]._version, d["_col_raw"], d["_col_raw"]._version)

    def _take_inspection(self):
        """The constructor's inspection if row / col are still the tensors it read, with no in-place edit since."""
        ins, self._inspection = self.__dict__.get("_inspection"), None
        if ins is None:
            return None
        state, max_bucket, row_t, row_v, col_t, col_v = ins
        d = self.__dict__
        if d["_row_raw"] is row_t and d["_col_raw"] is col_t and row_t._version == row_v and col_t._version == col_v:
            return state, max_bucket
        return None

    def _revalidate_if_edited(self) -> None:
        """Deferred sort only: an in-place edit of the caller's row / col between construction and first use is
        checked against the sparse sizes again (the reference had copied them by then; here they are read late)."""
        d = self.__dict__
        row_v, col_v, checked = d.get("_seen", (None, None, False))
        row, col = d["_row_raw"], d["_col_raw"]
        if not checked or ((row is None or row._version == row_v) and col._version == col_v):
            return
        self._inspection = None
        row_max, col_max = ops.index_max(row, col) if row is not None else (-1, ops.index_max(col)[0])
        M, N = self._sparse_sizes
        if row_max >= M or col_max >= N:
            raise AssertionError("row / col were edited in place after construction and no longer fit sparse_sizes")

    def _ensure_sorted(self) -> None:
        if not self.__dict__.get("_pending_sort", False):
            return
        self._revalidate_if_edited()
        self._pending_sort = False
        self._inspection = None  # only the fused constructor + coalesce pipeline consumes it
        d = self.__dict__
        if d["_row_raw"] is None:
            d["_row_raw"] = ops.ptr2ind(self._rowptr, d["_col_raw"].numel())  # storage.py:202
        in_order, _ = ops.coo_order(d["_row_raw"], d["_col_raw"])
        if not in_order:
            M, N = self._sparse_sizes
            d["_row_raw"], d["_col_raw"], d["_value_raw"], _ = ops.sort_coo(d["_row_raw"], d["_col_raw"],
                                                                           d["_value_raw"], M, N)
            self._csr2csc = None
            self._csc2csr = None

    def _sorted_field(name):  # noqa: N805  (property factory, evaluated in the class body)
        raw = name + "_raw"

        def get(self):
            self._ensure_sorted()
            return self.__dict__[raw]

        def set_(self, v):
            self.__dict__[raw] = v

        return property(get, set_)

    _row = _sorted_field("_row")
    _col = _sorted_field("_col")
    _value = _sorted_field("_value")
    del _sorted_field

    @classmethod
    def empty(cls, device="cuda"):
        e = torch.empty(0, dtype=torch.int64, device=device)
        return cls(row=e, col=e.clone(), sparse_sizes=(0, 0), is_sorted=True, trust_data=True)

    # ------------------------------------------------------------------ row / rowptr / col / value
    def has_row(self) -> bool:
        return self._row is not None

    def row(self) -> torch.Tensor:
        if self._row is None:
            if self._rowptr is None:
                raise ValueError
            self._row = ops.ptr2ind(self._rowptr, self._col.numel())  # storage.py:202
        return self._row

    def has_rowptr(self) -> bool:
        return self._rowptr is not None

    def rowptr(self) -> torch.Tensor:
        if self._rowptr is None:
            if self._row is None:
                raise ValueError
            self._rowptr = ops.ind2ptr(self._row, self._sparse_sizes[0])  # storage.py:218
        return self._rowptr

    def col(self) -> torch.Tensor:
        return self._col

    def has_value(self) -> bool:
        return self._value is not None

    def value(self) -> Optional[torch.Tensor]:
        return self._value

    def _prepare_value(self, value, layout):
        if value is not None:
            if get_layout(layout) == "csc":
                value = ops.gather_rows(value.contiguous(), self.csc2csr())  # storage.py:238-241
            value = value.contiguous()
            assert value.device == self._col.device
            assert value.shape[0] == self._col.numel()
        return value

    def set_value_(self, value, layout: Optional[str] = None):
        self._value = self._prepare_value(value, layout)
        parts = getattr(self, "_t_parts", None)
        if parts is not None:  # the parked value[csr2csc] is stale, also when `value` is the same tensor rewritten in place
            self._t_parts = parts[:3] + (None, -1)
        return self

    def set_value(self, value, layout: Optional[str] = None):
        return self._rewrap(value=self._prepare_value(value, layout))

    def _rewrap(self, **changes):
        kw = dict(row=self._row, rowptr=self._rowptr, col=self._col, value=self._value,
                  sparse_sizes=self._sparse_sizes, rowcount=self._rowcount, colptr=self._colptr,
                  colcount=self._colcount, csr2csc=self._csr2csc, csc2csr=self._csc2csr,
                  is_sorted=True, trust_data=True)
        kw.update(changes)
        return SparseStorage(**kw)

    def sparse_sizes(self) -> Tuple[int, int]:
        return self._sparse_sizes

    def sparse_size(self, dim: int) -> int:
        return self._sparse_sizes[dim]

    # ------------------------------------------------------------------ resize / reshape (storage.py:281-368)
    def sparse_resize(self, sparse_sizes: Tuple[int, int]):
        """storage.py:281-330: pointers grow by entries equal to nnz, counts by zeros; shrinking cuts; row / col /
        value and both permutations are shared."""
        assert len(sparse_sizes) == 2
        (old_m, old_n), nnz = self._sparse_sizes, self._col.numel()

        def ptr(t, size):
            return t if t is None or t.numel() == size + 1 else ops.pad_i64(t, size + 1, nnz)

        def count(t, size):
            return t if t is None or t.numel() == size else ops.pad_i64(t, size, 0)

        m, n = int(sparse_sizes[0]), int(sparse_sizes[1])
        return self._rewrap(sparse_sizes=(m, n), rowptr=ptr(self._rowptr, m), rowcount=count(self._rowcount, m),
                            colptr=ptr(self._colptr, n), colcount=count(self._colcount, n))

    def sparse_reshape(self, num_rows: int, num_cols: int):
        """storage.py:332-368: the same entries read row-major on a (num_rows, num_cols) grid; every cache is dropped."""
        assert num_rows > 0 or num_rows == -1
        assert num_cols > 0 or num_cols == -1
        assert num_rows > 0 or num_cols > 0
        total = self.sparse_size(0) * self.sparse_size(1)
        if num_rows == -1:
            num_rows = total // num_cols
        if num_cols == -1:
            num_cols = total // num_rows
        assert num_rows * num_cols == total
        row, col = ops.reshape_index(self.row(), self._col, self.sparse_size(1), num_cols)
        return SparseStorage(row=row, col=col, value=self._value, sparse_sizes=(num_rows, num_cols), is_sorted=True,
                             trust_data=True)

    # ------------------------------------------------------------------ lazy caches
    def has_rowcount(self) -> bool:
        return self._rowcount is not None

    def rowcount(self) -> torch.Tensor:
        if self._rowcount is None:
            self._rowcount = ops.ptr2count(self.rowptr())  # storage.py:373-381
        return self._rowcount

    def has_colptr(self) -> bool:
        return self._colptr is not None

    def _build_csc(self) -> None:
        """One stable radix sort on the column bits yields csr2csc and, as a by-product, colptr.
        The by-product is parked in `_colptr_spare` (not a cache key) and promoted by colptr(),
        so `cached_keys()` evolves exactly as in the reference."""
        need_ptr = self._colptr is None
        perm, colptr, _ = ops.csr2csc(self._col, self._sparse_sizes[1], need_ptr, False)
        self._csr2csc = perm
        if need_ptr:
            self._colptr_spare = colptr

    def _build_csc_carry(self, everything: bool) -> None:
        """The CSC side in ONE call (ops.csc_build): the stable sort by col carries (entry index, row, value) so
        its last pass also writes the transposed arrays, which are parked in `_t_parts` for `t()` (not a cache
        key; dropped by clear_cache_ and ignored once the value tensor has been replaced).  `everything` also
        builds colcount and csc2csr and promotes all CSC caches (fill_cache_, storage.py:488-496)."""
        r = ops.csc_build(self.row(), self._col, self._value, self._sparse_sizes[1], everything, everything)
        self._csr2csc = r["csr2csc"]
        self._t_parts = (r["t_row"], r["t_col"], r["t_value"], self._value,
                         -1 if self._value is None else self._value._version)
        if everything:
            if self._colptr is None:
                self._colptr = r["colptr"]
            if self._colcount is None:
                self._colcount = r["colcount"]
            self._csc2csr = r["csc2csr"]
        elif self._colptr is None:
            self._colptr_spare = r["colptr"]

    def _parked_value_valid(self, parts) -> bool:
        """The parked value[csr2csc] still describes `_value`: same tensor object and no in-place edit since
        (torch's version counter; edits through this package's kernels go through set_value_, which unparks)."""
        value_then, version_then = parts[3], parts[4]
        if self._value is None:
            return value_then is None and parts[2] is None
        return value_then is self._value and version_then == self._value._version

    def csc_ordered_value(self) -> Optional[torch.Tensor]:
        """value[csr2csc] when the carrying CSC sort left it behind and it is still current, else None (never
        computes anything): lets a column reduction run as a segment reduction over colptr."""
        parts = getattr(self, "_t_parts", None)
        if parts is None or parts[2] is None or self._value is None or not self._parked_value_valid(parts):
            return None
        return parts[2]

    def colcount_quiet(self) -> torch.Tensor:
        """The histogram of col for internal use: the cached `colcount` when there is one, else a private copy computed
        once - `cached_keys()` keeps evolving exactly as in the reference (whose scatter needs no counts)."""
        if self._colcount is not None:
            return self._colcount
        spare = getattr(self, "_colcount_spare", None)
        if spare is None:
            spare = self._colcount_spare = ops.scatter_add_ones(self._col, self._sparse_sizes[1])
        return spare

    def column_reduction_accel(self):
        """(value[csr2csc], colptr) for `reduction(dim=0)` (reduce.py:40-42), or None.  One column reduction runs as an
        atomic scatter (a third of HBM speed); from the SECOND one on the same matrix and value tensor, the carrying
        CSC sort is run once and its two by-products are kept PRIVATELY (no cache key appears), so that later column
        reductions are segment reductions over contiguous columns (0.65 -> 0.10 ms on the configs[2] shape).
        One 4-byte value per entry, no gradient; dropped by clear_cache_ and when the value tensor changes."""
        value = self._value
        if value is None or value.dim() != 1 or value.element_size() != 4 or value.requires_grad or value.numel() == 0:
            return None
        acc = getattr(self, "_dim0_accel", None)
        if acc is not None and acc[2] is value and acc[3] == value._version:
            return acc[0], acc[1]
        self._dim0_calls = getattr(self, "_dim0_calls", 0) + 1
        if self._dim0_calls < 2:
            return None
        r = ops.csc_build(self.row(), self._col, value, self._sparse_sizes[1], False, False)
        self._dim0_accel = (r["t_value"], r["colptr"], value, value._version)
        return r["t_value"], r["colptr"]

    def take_transposed_parts(self):
        """(col[csr2csc], row[csr2csc], value[csr2csc]) for transpose.py:9-33, building csr2csc if needed.
        Uses the by-products of the carrying sort when they are there, else one fused gather."""
        if self._csr2csc is None:
            self._build_csc_carry(False)
        parts, self._t_parts = getattr(self, "_t_parts", None), None
        if parts is not None:
            t_row, t_col, t_value = parts[:3]
            if not self._parked_value_valid(parts):  # set_value_ / in-place edit since: re-gather the values only
                t_value = None if self._value is None else ops.gather_rows(self._value, self._csr2csc)
            return t_row, t_col, t_value
        return ops.transpose_gather(self._csr2csc, self.row(), self._col, self._value, self._colptr,
                                    self._sparse_sizes[1])

    def colptr(self) -> torch.Tensor:
        if self._colptr is None:
            spare = getattr(self, "_colptr_spare", None)
            if spare is not None:
                self._colptr, self._colptr_spare = spare, None
            elif self._csr2csc is not None:
                # storage.py:392-393: ind2ptr(col[csr2csc], N)
                self._colptr = ops.ind2ptr(ops.gather_rows(self._col, self._csr2csc), self._sparse_sizes[1])
            else:
                # storage.py:395-398: [0, cumsum(colcount)] — the histogram is scanned in one pass
                self._colptr = ops.count2ptr(self.colcount())
        return self._colptr

    def has_colcount(self) -> bool:
        return self._colcount is not None

    def colcount(self) -> torch.Tensor:
        if self._colcount is None:
            if self._colptr is not None:
                self._colcount = ops.ptr2count(self._colptr)  # storage.py:408-412
            else:
                self._colcount = ops.scatter_add_ones(self._col, self._sparse_sizes[1])  # storage.py:413-418
        return self._colcount

    def has_csr2csc(self) -> bool:
        return self._csr2csc is not None

    def csr2csc(self) -> torch.Tensor:
        if self._csr2csc is None:
            self._build_csc()  # storage.py:425-434 (+ colptr for free when it is not cached yet)
        return self._csr2csc

    def has_csc2csr(self) -> bool:
        return self._csc2csr is not None

    def csc2csr(self) -> torch.Tensor:
        if self._csc2csr is None:
            self._csc2csr = ops.invert_perm(self.csr2csc())  # storage.py:439-447
        return self._csc2csr

    def is_coalesced(self) -> bool:
        """storage.py:449-452."""
        if self._col.numel() <= 1:
            return True
        return ops.coo_order(self.row(), self._col)[1]

    def coalesce(self, reduce: str = "add"):
        """storage.py:454-486; returns `self` when nothing has to be merged (:459-460)."""
        d = self.__dict__
        if (d.get("_pending_sort", False) and d["_row_raw"] is not None and self._rowptr is None
                and self._csr2csc is None and self._csc2csr is None and self._colptr is None):
            # constructor immediately followed by coalesce(): one fused sort + dedup + rowptr pipeline
            self._revalidate_if_edited()
            M, N = self._sparse_sizes
            row, col, value, rowptr = ops.coalesce(d["_row_raw"], d["_col_raw"], d["_value_raw"], M, N, reduce, True,
                                                   self._take_inspection())
            if row.numel() == d["_col_raw"].numel():  # nothing merged: this object simply becomes sorted
                d["_row_raw"], d["_col_raw"], d["_value_raw"] = row, col, value
                self._rowptr = rowptr
                self._pending_sort = False
                return self
            return SparseStorage(row=row, rowptr=rowptr, col=col, value=value, sparse_sizes=self._sparse_sizes,
                                 is_sorted=True, trust_data=True)
        if self._col.numel() <= 1:
            return self
        row, col, value = ops.coalesce_sorted(self.row(), self._col, self._value, reduce, self._sparse_sizes)
        if row.numel() == self._col.numel():
            return self
        return SparseStorage(row=row, col=col, value=value, sparse_sizes=self._sparse_sizes, is_sorted=True,
                             trust_data=True)

    def fill_cache_(self, carry: bool = True):
        """storage.py:488-496 (same set of caches; csr2csc is built first so colptr comes with it).  `carry` (extension,
        default on): the CSC sort also carries row and value so that a following `t()` costs nothing; a caller that
        will not transpose (the row-sharded build) turns it off and saves 20 bytes per entry of writes."""
        self.row()
        self.rowptr()
        # the CSC side first: its kernels are the long ones, the host prepares the small launches behind them
        if self._csr2csc is None and self._col.numel() > 0:
            if carry:
                self._build_csc_carry(True)
            else:
                r = ops.csc_build(None, self._col, None, self._sparse_sizes[1], True, True, carry=False)
                self._csr2csc, self._csc2csr = r["csr2csc"], r["csc2csr"]
                if self._colptr is None:
                    self._colptr = r["colptr"]
                if self._colcount is None:
                    self._colcount = r["colcount"]
        self.rowcount()
        self.csr2csc()
        self.colptr()
        self.colcount()
        self.csc2csr()
        return self

    def clear_cache_(self):
        self._colptr_spare = None
        self._colcount_spare = None
        self._dim0_accel = None
        self._dim0_calls = 0
        self._t_parts = None
        self._rowcount = None
        self._colptr = None
        self._colcount = None
        self._csr2csc = None
        self._csc2csr = None
        return self

    def cached_keys(self) -> List[str]:
        keys: List[str] = []
        if self.has_rowcount():
            keys.append("rowcount")
        if self.has_colptr():
            keys.append("colptr")
        if self.has_colcount():
            keys.append("colcount")
        if self.has_csr2csc():
            keys.append("csr2csc")
        if self.has_csc2csr():
            keys.append("csc2csr")
        return keys

    def num_cached_keys(self) -> int:
        return len(self.cached_keys())

    # ------------------------------------------------------------------ copies, dtype and placement (storage.py:523-776)
    def _fields(self):
        return dict(row=self._row, rowptr=self._rowptr, col=self._col, value=self._value, rowcount=self._rowcount,
                    colptr=self._colptr, colcount=self._colcount, csr2csc=self._csr2csc, csc2csr=self._csc2csr)

    def _rebuilt(self, fn):
        """A new storage whose every present tensor is fn(tensor); sizes kept, nothing re-validated or re-sorted."""
        kw = {k: (None if t is None else fn(t)) for k, t in self._fields().items()}
        return SparseStorage(sparse_sizes=self._sparse_sizes, is_sorted=True, trust_data=True, **kw)

    def copy(self):
        """storage.py:523-537: a new storage object SHARING every tensor (test_storage.py:148-150)."""
        return self._rebuilt(lambda t: t)

    def clone(self):
        """storage.py:539-580: a new storage owning copies of every tensor (test_storage.py:152-154)."""
        return self._rebuilt(lambda t: t.clone())

    def type(self, dtype: torch.dtype, non_blocking: bool = False):
        """storage.py:582-594: value cast; `self` when there is no value or the dtype already matches."""
        value = self._value
        if value is None or dtype == value.dtype:
            return self
        return self.set_value(value.to(dtype=dtype, non_blocking=non_blocking), layout="coo")

    def type_as(self, tensor: torch.Tensor, non_blocking: bool = False):
        return self.type(dtype=tensor.dtype, non_blocking=non_blocking)

    def to_device(self, device, non_blocking: bool = False):
        """storage.py:599-641: every tensor moved, caches kept; `self` when already there.  A CUDA target gives a
        working storage; a CPU target gives an inert host copy (see the constructor), as pin_memory() does."""
        device = torch.device(device)
        here = self._col.device
        if device == here or (device.type == "cuda" and device.index is None and here.type == "cuda"):
            return self
        return self._rebuilt(lambda t: t.to(device, non_blocking=non_blocking))

    def device_as(self, tensor: torch.Tensor, non_blocking: bool = False):
        return self.to_device(device=tensor.device, non_blocking=non_blocking)

    def cuda(self):
        """storage.py:646-691: `self` when the tensors already live on a CUDA device."""
        if self._col.is_cuda:
            return self
        return self._rebuilt(lambda t: t.cuda())

    def pin_memory(self):
        """storage.py:693-733: page-locked host copies of every tensor (a staging copy for asynchronous uploads:
        `st.pin_memory().to_device("cuda", non_blocking=True)`); the result holds data only."""
        return self._rebuilt(lambda t: t.cpu().pin_memory() if t.is_cuda else t.pin_memory())

    def is_pinned(self) -> bool:
        """storage.py:735-762 as intended (the reference body shadows its own accumulator): true when
        every present tensor is page-locked host memory."""
        return all(t.is_pinned() for t in self._fields().values() if t is not None)

    def share_memory_(self):
        """storage.py:765-766: unsupported in the reference too (warns, returns None)."""
        warnings.warn("`share_memory_` is not supported (device tensors are not shared between processes this way). "
                      "Please avoid calling this api.")

    def is_shared(self) -> bool:
        """storage.py:769-772: always False, with the reference's warning."""
        warnings.warn("`is_shared` is not supported. The return value is always `False`.")
        return False
