"""Operator layer: the callables the reference gets from ``paddle_sparse_ops`` (csrc/convert.cpp,
csrc/sample.cpp, csrc/version.cpp), from ``paddle_scatter`` (segment_csr / gather_csr / scatter /
scatter_add) and from ``index_sort`` (paddle_sparse/utils.py:14-23), plus the fused entry points
SparseStorage uses instead of the reference's chains of stock tensor ops.

Every function takes device tensors, allocates its outputs on the same device, launches on the
current stream and calls straight into libpaddle_sparse_b200.so through ctypes with raw pointers
(include/ps_abi.h).  torch is used only as the array library: allocation, ``data_ptr()`` and the
current stream (the production adapter hands Paddle tensors over by DLPack).
GPU only: a CPU tensor raises, exactly opposite to csrc/sample.cpp:13-15 which throws on GPU ones.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Tuple

import torch

from . import _lib
from . import autograd as _ag

PS_F32, PS_BF16, PS_F16, PS_F64, PS_I32, PS_I64, PS_U8 = range(7)
_DTYPES = {
    torch.float32: PS_F32, torch.bfloat16: PS_BF16, torch.float16: PS_F16, torch.float64: PS_F64,
    torch.int32: PS_I32, torch.int64: PS_I64,
}
_REDUCE = {"sum": 0, "add": 0, "mean": 1, "min": 2, "max": 3}  # testing.py:10; "add" == "sum"


def reduce_code(reduce: str) -> int:
    try:
        return _REDUCE[reduce]
    except KeyError:
        raise ValueError(f"unknown reduce '{reduce}'") from None


def dtype_code(t: torch.Tensor) -> int:
    try:
        return _DTYPES[t.dtype]
    except KeyError:
        raise TypeError(f"value dtype {t.dtype} is not supported by paddle_sparse_b200") from None


def _dev(t: torch.Tensor, name: str = "tensor") -> torch.Tensor:
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise RuntimeError(f"paddle_sparse_b200 is GPU-only (no CPU fallback): `{name}` is not a CUDA tensor")
    return t.contiguous()


def _idx(t: torch.Tensor, name: str = "index") -> torch.Tensor:
    t = _dev(t, name)
    if t.dtype != torch.int64:
        raise TypeError(f"`{name}` must be int64, got {t.dtype}")
    return t


def _p(t: Optional[torch.Tensor]):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _stream(t: torch.Tensor):
    return ctypes.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


def _width(value: torch.Tensor) -> int:
    n = value.shape[0]
    return (value.numel() // n) if n > 0 else int(torch.Size(value.shape[1:]).numel())


def _row_bytes(value: torch.Tensor) -> int:
    return _width(value) * value.element_size()


def _new(ref: torch.Tensor, n, dtype=torch.int64) -> torch.Tensor:
    return torch.empty(n, dtype=dtype, device=ref.device)


def _host_i64(n: int = 1):
    return (ctypes.c_int64 * n)()


# ------------------------------------------------------------------ paddle_sparse_ops (csrc/)
def sparse_cuda_version() -> torch.Tensor:
    """csrc/version.cpp:8-33: [1] int64 CPU tensor with the CUDA_VERSION of the build."""
    return torch.tensor([int(_lib.lib().ps_cuda_version())], dtype=torch.int64)


def ind2ptr(ind: torch.Tensor, M: int) -> torch.Tensor:
    """csrc/convert.cpp:13-23."""
    ind = _idx(ind, "ind")
    out = _new(ind, M + 1)
    with torch.cuda.device(ind.device):
        _lib.call("ps_ind2ptr", _p(ind), ind.numel(), M, _p(out), _stream(ind))
    return out


def ptr2ind(ptr: torch.Tensor, E: int) -> torch.Tensor:
    """csrc/convert.cpp:46-56."""
    ptr = _idx(ptr, "ptr")
    out = _new(ptr, E)
    with torch.cuda.device(ptr.device):
        _lib.call("ps_ptr2ind", _p(ptr), ptr.numel() - 1, E, _p(out), _stream(ptr))
    return out


def sample_uniform(rowptr, col, subset, num_neighbors: int, seed: int = 0) -> torch.Tensor:
    """sample.py:10-28 in one launch: [S, num_neighbors] neighbours drawn uniformly with replacement (subset None: every
    row).  Draw (i, k) is a counter-based function of `seed`."""
    rowptr, col = _idx(rowptr, "rowptr"), _idx(col, "col")
    subset = None if subset is None else _idx(subset, "subset")
    S = rowptr.numel() - 1 if subset is None else subset.numel()
    out = torch.empty((S, int(num_neighbors)), dtype=torch.int64, device=col.device)
    with torch.cuda.device(col.device):
        _lib.call("ps_sample_uniform", _p(rowptr), _p(col), col.numel(), _p(subset), S, int(num_neighbors),
                  ctypes.c_uint64(int(seed) & (2**64 - 1)), _p(out), _stream(col))
    return out


def sample_adj(rowptr, col, idx, num_neighbors: int, replace: bool, seed: int = 0):
    """csrc/sample.cpp:8-24 (the reference throws on GPU tensors; this is the GPU path).
    Returns [out_rowptr, out_col, n_id, e_id]."""
    rowptr, col, idx = _idx(rowptr, "rowptr"), _idx(col, "col"), _idx(idx, "idx")
    if idx.dim() != 1:
        raise RuntimeError("sample_adj: idx must be 1-D")  # CHECK_INPUT, sample_cpu.cpp:19
    S = idx.numel()
    st = _stream(rowptr)
    with torch.cuda.device(rowptr.device):
        hb = _host_i64()
        _lib.call("ps_sample_adj_bound", _p(rowptr), _p(idx), S, int(num_neighbors), int(bool(replace)), hb, st)
        bound = int(hb[0])
        out_rowptr = _new(rowptr, S + 1)
        out_col = _new(rowptr, bound)
        out_e_id = _new(rowptr, bound)
        out_n_id = _new(rowptr, S + bound)
        sizes = _host_i64(2)
        _lib.call("ps_sample_adj", _p(rowptr), _p(col), rowptr.numel() - 1, col.numel(), _p(idx), S,
                  int(num_neighbors), int(bool(replace)), ctypes.c_uint64(seed & (2**64 - 1)), bound,
                  _p(out_rowptr), _p(out_col), _p(out_n_id), _p(out_e_id), sizes, st)
    E, Nn = int(sizes[0]), int(sizes[1])
    return [out_rowptr, out_col[:E], out_n_id[:Nn], out_e_id[:E]]


# ------------------------------------------------------------------ pointers and counts
def ptr2count(ptr: torch.Tensor) -> torch.Tensor:
    """storage.py:373-381 / 405-412."""
    ptr = _idx(ptr, "ptr")
    M = ptr.numel() - 1
    out = _new(ptr, M)
    with torch.cuda.device(ptr.device):
        _lib.call("ps_ptr2count", _p(ptr), M, _p(out), _stream(ptr))
    return out


def count2ptr(count: torch.Tensor, idx: Optional[torch.Tensor] = None, want_total: bool = False):
    """[0, cumsum(count[idx])] (storage.py:395-398, index_select.py:21-23)."""
    count = _idx(count, "count")
    idx = None if idx is None else _idx(idx, "idx")
    n = count.numel() if idx is None else idx.numel()
    out = _new(count, n + 1)
    tot = _host_i64() if want_total else None
    with torch.cuda.device(count.device):
        _lib.call("ps_count2ptr", _p(count), _p(idx), n, _p(out), tot, _stream(count))
    return (out, int(tot[0])) if want_total else out


def scatter_add_ones(index: torch.Tensor, dim_size: int) -> torch.Tensor:
    """storage.py:413-418: scatter_add(ones_like(col), col, dim_size=N) == histogram."""
    index = _idx(index)
    out = _new(index, dim_size)
    with torch.cuda.device(index.device):
        _lib.call("ps_histogram", _p(index), index.numel(), dim_size, _p(out), _stream(index))
    return out


def index_max(a: torch.Tensor, b: Optional[torch.Tensor] = None) -> Tuple[int, int]:
    """storage.py:65-89: (int(a.max()), int(b.max())) in one launch and one sync; -1 when empty."""
    a = _idx(a, "a")
    b = None if b is None else _idx(b, "b")
    h = _host_i64(2)
    with torch.cuda.device(a.device):
        _lib.call("ps_index_max", _p(a), a.numel(), _p(b), 0 if b is None else b.numel(), h, _stream(a))
    return int(h[0]), int(h[1])


def coo_order(row: torch.Tensor, col: torch.Tensor) -> Tuple[bool, bool]:
    """(is_sorted, is_coalesced) of the (row, col) sequence (storage.py:158-163, 449-452)."""
    row, col = _idx(row, "row"), _idx(col, "col")
    h = (ctypes.c_int32 * 2)()
    with torch.cuda.device(row.device):
        _lib.call("ps_coo_order", _p(row), _p(col), row.numel(), h, _stream(row))
    return bool(h[0]), bool(h[1])


# ------------------------------------------------------------------ construct / coalesce
def _value_args(value: Optional[torch.Tensor]):
    if value is None:
        return None, 0, 1
    return value, dtype_code(value), max(_width(value), 1)


def sort_coo(row, col, value, M: int, N: int, want_perm: bool = False):
    """Sort-on-construct (storage.py:158-171): stable sort by row*N + col.
    Returns (row, col, value, perm|None)."""
    if _ag.needs_grad(value):  # indices by the kernel, values by a differentiable gather
        r, c, _, perm = sort_coo(row, col, None, M, N, True)
        return r, c, gather_rows(value, perm), (perm if want_perm else None)
    row, col = _idx(row, "row"), _idx(col, "col")
    value = None if value is None else _dev(value, "value")
    v, dt, w = _value_args(value)
    nnz = row.numel()
    out_row, out_col = _new(row, nnz), _new(row, nnz)
    out_value = None if value is None else torch.empty_like(value)
    out_perm = _new(row, nnz) if want_perm else None
    with torch.cuda.device(row.device):
        _lib.call("ps_sort_coo", _p(row), _p(col), nnz, M, N, _p(v), dt, w, _p(out_row), _p(out_col),
                  _p(out_value), _p(out_perm), _stream(row))
    return out_row, out_col, out_value, out_perm


def coo_inspect(row: torch.Tensor, col: torch.Tensor, M: int, N: int):
    """The constructor's validation pass (storage.py:65-89) fused with the first read `coalesce` needs: one pass
    over (row, col) gives (max row, max col) and, where the hybrid plan applies, the bucket histogram left on the
    device.  Returns (max_row, max_col, inspection | None); hand the inspection to `coalesce` on the same arrays."""
    row, col = _idx(row, "row"), _idx(col, "col")
    state = torch.empty(int(_lib.lib().ps_coo_inspect_bytes()), dtype=torch.uint8, device=row.device)
    h = _host_i64(3)
    with torch.cuda.device(row.device):
        _lib.call("ps_coo_inspect", _p(row), _p(col), row.numel(), M, N, _p(state), h, _stream(row))
    return int(h[0]), int(h[1]), ((state, int(h[2])) if h[2] >= 0 else None)


def coalesce(row, col, value, M: int, N: int, reduce: str = "add", want_rowptr: bool = False, inspection=None):
    """coalesce.py:8-29 fused: packed-key radix sort + run detection + segmented reduce.
    Returns (row', col', value'), plus rowptr' (= ind2ptr(row', M), written by the same pass) when asked.
    `inspection`: what coo_inspect left behind for these very arrays (saves a read and a host sync)."""
    if _ag.needs_grad(value):  # sort (indices), run starts, then gather + segment reduction: both differentiable
        r, c, _, perm = sort_coo(row, col, None, M, N, True)
        ptr = _ag.run_starts(r, c)
        heads = ptr[:-1]
        out = (r[heads].contiguous(), c[heads].contiguous(), segment_csr(gather_rows(value, perm), ptr, None, reduce))
        return out + (ind2ptr(out[0], M),) if want_rowptr else out
    row, col = _idx(row, "row"), _idx(col, "col")
    value = None if value is None else _dev(value, "value")
    v, dt, w = _value_args(value)
    nnz = row.numel()
    out_row, out_col = _new(row, nnz), _new(row, nnz)
    out_value = None if value is None else torch.empty_like(value)
    rowptr = _new(row, M + 1) if want_rowptr else None
    h = _host_i64()
    with torch.cuda.device(row.device):
        _lib.call("ps_coalesce", _p(row), _p(col), nnz, M, N, _p(v), dt, w, reduce_code(reduce), _p(out_row),
                  _p(out_col), _p(out_value), _p(rowptr), h, _p(inspection[0]) if inspection else None,
                  inspection[1] if inspection else -1, _stream(row))
    k = int(h[0])
    res = (out_row[:k], out_col[:k], (None if out_value is None else out_value[:k]))
    return res + (rowptr,) if want_rowptr else res


def coalesce_u32(row32, col32, row_offset: int, value, M: int, N: int, reduce: str = "add", want_rowptr: bool = False):
    """`coalesce` on tuples held as 32-bit words: entry e is (row32[e] - row_offset, col32[e]).  This is what the
    fused peer exchange delivers (GLOBAL row ids, 12 bytes per tuple on the wire); row32 / col32 are int32 tensors
    whose bits are read as unsigned.  Same results as `coalesce` on the widened, shifted arrays."""
    for name, t in (("row32", row32), ("col32", col32)):
        if not (t.is_cuda and t.dtype == torch.int32 and t.is_contiguous()):
            raise ValueError(f"coalesce_u32: {name} must be a contiguous CUDA int32 tensor")
    value = None if value is None else _dev(value, "value")
    v, dt, w = _value_args(value)
    nnz = row32.numel()
    out_row = torch.empty(nnz, dtype=torch.int64, device=row32.device)
    out_col = torch.empty(nnz, dtype=torch.int64, device=row32.device)
    out_value = None if value is None else torch.empty_like(value)
    rowptr = torch.empty(M + 1, dtype=torch.int64, device=row32.device) if want_rowptr else None
    h = _host_i64()
    with torch.cuda.device(row32.device):
        _lib.call("ps_coalesce_u32", _p(row32), _p(col32), int(row_offset), nnz, M, N, _p(v), dt, w, reduce_code(reduce),
                  _p(out_row), _p(out_col), _p(out_value), _p(rowptr), h, _stream(row32))
    k = int(h[0])
    res = (out_row[:k], out_col[:k], (None if out_value is None else out_value[:k]))
    return res + (rowptr,) if want_rowptr else res


def coalesce_sorted(row, col, value, reduce: str = "add", sparse_sizes=None):
    """SparseStorage.coalesce on sorted input (storage.py:454-486).  Returns (row', col', value').
    `sparse_sizes` (M, N): taken from the largest ids when not given (one more launch and sync); ids of 2^32 and
    beyond are refused by the kernel side."""
    if _ag.needs_grad(value):
        r, c = _idx(row, "row"), _idx(col, "col")
        ptr = _ag.run_starts(r, c)
        heads = ptr[:-1]
        return r[heads].contiguous(), c[heads].contiguous(), segment_csr(value, ptr, None, reduce)
    row, col = _idx(row, "row"), _idx(col, "col")
    value = None if value is None else _dev(value, "value")
    v, dt, w = _value_args(value)
    nnz = row.numel()
    out_row, out_col = _new(row, nnz), _new(row, nnz)
    out_value = None if value is None else torch.empty_like(value)
    h = _host_i64()
    if sparse_sizes is None:
        mr, mc = index_max(row, col)
        sparse_sizes = (mr + 1, mc + 1)
    with torch.cuda.device(row.device):
        _lib.call("ps_coalesce_sorted", _p(row), _p(col), nnz, int(sparse_sizes[0]), int(sparse_sizes[1]), _p(v), dt, w,
                  reduce_code(reduce), _p(out_row), _p(out_col), _p(out_value), h, _stream(row))
    k = int(h[0])
    return out_row[:k], out_col[:k], (None if out_value is None else out_value[:k])


# ------------------------------------------------------------------ CSC structures
def csr2csc(col: torch.Tensor, N: int, want_colptr: bool = True, want_colcount: bool = False):
    """storage.py:425-434 (+ colptr :386-400, colcount :405-420 as by-products).
    Returns (csr2csc, colptr|None, colcount|None)."""
    col = _idx(col, "col")
    nnz = col.numel()
    perm = _new(col, nnz)
    colptr = _new(col, N + 1) if want_colptr else None
    colcount = _new(col, N) if want_colcount else None
    with torch.cuda.device(col.device):
        _lib.call("ps_csr2csc", _p(col), nnz, N, _p(perm), _p(colptr), _p(colcount), _stream(col))
    return perm, colptr, colcount


def csc_build(row: Optional[torch.Tensor], col: torch.Tensor, value: Optional[torch.Tensor], N: int,
              want_colcount: bool = True, want_csc2csr: bool = True, carry: bool = True):
    """storage.py:386-447 + transpose.py:9-33 in one call: csr2csc, colptr, colcount, csc2csr and the transposed
    arrays (col[csr2csc], row[csr2csc], value[csr2csc]); the values ride through the sort when they are one 4-byte
    element per entry, else they are gathered once at the end.
    Returns dict(csr2csc, colptr, colcount, csc2csr, t_row, t_col, t_value)."""
    col = _idx(col, "col")
    if not carry:  # csr2csc, colptr, colcount, csc2csr only: nothing rides along, no transposed arrays
        nnz = col.numel()
        perm, colptr = _new(col, nnz), _new(col, N + 1)
        colcount = _new(col, N) if want_colcount else None
        inv = _new(col, nnz) if want_csc2csr else None
        with torch.cuda.device(col.device):
            _lib.call("ps_csc_build", None, _p(col), None, 0, nnz, N, _p(perm), _p(colptr), _p(colcount), _p(inv), None,
                      None, None, _stream(col))
        return dict(csr2csc=perm, colptr=colptr, colcount=colcount, csc2csr=inv, t_row=None, t_col=None, t_value=None)
    row = _idx(row, "row")
    value = None if value is None else _dev(value, "value")
    nnz = col.numel()
    carry = value is not None and nnz > 0 and _row_bytes(value) == 4 and not _ag.needs_grad(value)
    perm, t_row, t_col = _new(col, nnz), _new(col, nnz), _new(col, nnz)
    colptr = _new(col, N + 1)
    colcount = _new(col, N) if want_colcount else None
    inv = _new(col, nnz) if want_csc2csr else None
    t_value = torch.empty_like(value) if carry else None
    with torch.cuda.device(col.device):
        _lib.call("ps_csc_build", _p(row), _p(col), _p(value) if carry else None, 4 if carry else 0, nnz, N, _p(perm),
                  _p(colptr), _p(colcount), _p(inv), _p(t_row), _p(t_col), _p(t_value), _stream(col))
    if value is not None and not carry:
        t_value = gather_rows(value, perm)
    return dict(csr2csc=perm, colptr=colptr, colcount=colcount, csc2csr=inv, t_row=t_row, t_col=t_col, t_value=t_value)


def invert_perm(perm: torch.Tensor) -> torch.Tensor:
    """storage.py:439-447: argsort of a permutation == its inverse."""
    perm = _idx(perm, "perm")
    out = _new(perm, perm.numel())
    with torch.cuda.device(perm.device):
        _lib.call("ps_invert_perm", _p(perm), perm.numel(), _p(out), _stream(perm))
    return out


def gather_rows(src: torch.Tensor, index: torch.Tensor) -> torch.Tensor:
    """src[index] along dim 0 (storage.py:166-169, 238-241; transpose.py:15; index_select.py:104-108)."""
    if _ag.needs_grad(src):
        return _ag.GatherRows.apply(_dev(src, "src"), _idx(index, "index"))
    src, index = _dev(src, "src"), _idx(index, "index")
    out = torch.empty((index.numel(),) + tuple(src.shape[1:]), dtype=src.dtype, device=src.device)
    with torch.cuda.device(src.device):
        _lib.call("ps_gather_rows", _p(src), _row_bytes(src) if src.shape[0] else src.element_size() * _width(src),
                  _p(index), index.numel(), _p(out), _stream(src))
    return out


def transpose_gather(csr2csc_, row, col, value, colptr=None, N: int = 0):
    """transpose.py:9-33: (col[csr2csc], row[csr2csc], value[csr2csc]) in one pass."""
    if _ag.needs_grad(value):
        out_row, out_col, _ = transpose_gather(csr2csc_, row, col, None, colptr, N)
        return out_row, out_col, gather_rows(value, csr2csc_)
    perm, row, col = _idx(csr2csc_, "csr2csc"), _idx(row, "row"), _idx(col, "col")
    value = None if value is None else _dev(value, "value")
    nnz = row.numel()
    out_row, out_col = _new(row, nnz), _new(row, nnz)
    out_value = None if value is None else torch.empty_like(value)
    colptr = None if colptr is None else _idx(colptr, "colptr")
    with torch.cuda.device(row.device):
        _lib.call("ps_transpose_gather", _p(perm), _p(row), _p(col), _p(value),
                  0 if value is None else _row_bytes(value), nnz, _p(colptr), N, _p(out_row), _p(out_col),
                  _p(out_value), _stream(row))
    return out_row, out_col, out_value


# ------------------------------------------------------------------ paddle_scatter replacements
def segment_csr(src: torch.Tensor, indptr: torch.Tensor, out=None, reduce: str = "sum") -> torch.Tensor:
    """paddle_scatter.segment_csr(src, indptr, None, reduce) along dim 0 (storage.py:471, reduce.py:51)."""
    assert out is None, "segment_csr: `out` is not supported"
    if _ag.needs_grad(src):
        return _ag.SegmentCSR.apply(_dev(src, "src"), _idx(indptr, "indptr"), reduce)
    src, indptr = _dev(src, "src"), _idx(indptr, "indptr")
    S = indptr.numel() - 1
    res = torch.empty((S,) + tuple(src.shape[1:]), dtype=src.dtype, device=src.device)
    with torch.cuda.device(src.device):
        _lib.call("ps_segment_reduce", _p(src), dtype_code(src), max(_width(src), 1), src.shape[0], _p(indptr), S,
                  reduce_code(reduce), _p(res), _stream(src))
    return res


def gather_csr(src: torch.Tensor, indptr: torch.Tensor, E: Optional[int] = None) -> torch.Tensor:
    """paddle_scatter.gather_csr(src, indptr) (mul.py:16,79; index_select.py:29,65).  E = indptr[-1]
    (pass it when known: it saves the device->host read)."""
    src, indptr = _dev(src, "src"), _idx(indptr, "indptr")
    S = indptr.numel() - 1
    if E is None:
        E = int(indptr[-1].item())
    if _ag.needs_grad(src):
        return _ag.GatherCSR.apply(src, indptr, E)
    out = torch.empty((E,) + tuple(src.shape[1:]), dtype=src.dtype, device=src.device)
    with torch.cuda.device(src.device):
        _lib.call("ps_gather_csr", _p(src), src.element_size() * max(_width(src), 1), _p(indptr), S, E, _p(out),
                  _stream(src))
    return out


def scatter(src: torch.Tensor, index: torch.Tensor, dim: int = 0, out=None, dim_size: Optional[int] = None,
            reduce: str = "sum", count: Optional[torch.Tensor] = None) -> torch.Tensor:
    """paddle_scatter.scatter(src, index, 0, None, dim_size, reduce) (reduce.py:42).  `count` (optional): the
    histogram of `index` when the caller holds it (a cached colcount) - saves one atomic per entry."""
    assert dim == 0 and out is None and dim_size is not None
    if _ag.needs_grad(src):
        return _ag.Scatter.apply(_dev(src, "src"), _idx(index, "index"), dim_size, reduce, count)
    src, index = _dev(src, "src"), _idx(index, "index")
    count = None if count is None else _idx(count, "count")
    assert count is None or count.numel() == dim_size
    res = torch.empty((dim_size,) + tuple(src.shape[1:]), dtype=src.dtype, device=src.device)
    with torch.cuda.device(src.device):
        _lib.call("ps_scatter_reduce", _p(src), dtype_code(src), max(_width(src), 1), _p(index), index.numel(),
                  dim_size, reduce_code(reduce), _p(res), _p(count), _stream(src))
    return res


def reduce_all(src: torch.Tensor, reduce: str = "sum") -> torch.Tensor:
    """reduce.py:17-27 dim=None: 0-d result of the dense reduction over every element."""
    src = _dev(src, "src")
    if _ag.needs_grad(src):  # a dense reduction of the value tensor: the framework's own differentiable op
        reduce = "sum" if reduce == "add" else reduce
        return {"sum": src.sum, "mean": src.mean, "min": src.min, "max": src.max}[reduce]()
    res = torch.empty(1, dtype=src.dtype, device=src.device)
    with torch.cuda.device(src.device):
        _lib.call("ps_reduce_all", _p(src), dtype_code(src), src.numel(), reduce_code(reduce), _p(res), _stream(src))
    return res[0]


# ------------------------------------------------------------------ broadcast multiply / add (mul.py:76-119, add.py:12-81)
def _no_inplace_grad(out, *tensors):
    if out is not None and _ag.needs_grad(*tensors):
        raise RuntimeError("in-place operation on values that require grad: use the out-of-place form")


def _bc_row(sym, value, rowptr, other, out=None):
    _no_inplace_grad(out, value, other)
    if _ag.needs_grad(value, other):
        return _ag.BroadcastRow.apply(_dev(value, "value"), _idx(rowptr, "rowptr"), _dev(other, "other"), sym == "ps_mul_row")
    value, rowptr, other = _dev(value, "value"), _idx(rowptr, "rowptr"), _dev(other, "other")
    out = torch.empty_like(value) if out is None else out
    with torch.cuda.device(value.device):
        _lib.call(sym, _p(value), dtype_code(value), max(_width(value), 1), _p(rowptr), rowptr.numel() - 1,
                  value.shape[0], _p(other), _p(out), _stream(value))
    return out


def _bc_col(sym, value, col, other, out=None):
    _no_inplace_grad(out, value, other)
    if _ag.needs_grad(value, other):
        return _ag.BroadcastCol.apply(_dev(value, "value"), _idx(col, "col"), _dev(other, "other"), sym == "ps_mul_col")
    value, col, other = _dev(value, "value"), _idx(col, "col"), _dev(other, "other")
    out = torch.empty_like(value) if out is None else out
    with torch.cuda.device(value.device):
        _lib.call(sym, _p(value), dtype_code(value), max(_width(value), 1), _p(col), value.shape[0],
                  _p(other), _p(out), _stream(value))
    return out


def _bc_nnz(sym, value, other, out=None):
    _no_inplace_grad(out, value, other)
    if _ag.needs_grad(value, other):  # element-wise on [nnz, ...]: the framework's own differentiable op
        return value * other if sym == "ps_mul_nnz" else value + other
    value, other = _dev(value, "value"), _dev(other, "other")
    out = torch.empty_like(value) if out is None else out
    with torch.cuda.device(value.device):
        _lib.call(sym, _p(value), dtype_code(value), value.numel(), _p(other), _p(out), _stream(value))
    return out


def mul_row(value, rowptr, other, out=None):
    return _bc_row("ps_mul_row", value, rowptr, other, out)


def mul_col(value, col, other, out=None):
    return _bc_col("ps_mul_col", value, col, other, out)


def mul_nnz(value, other, out=None):
    return _bc_nnz("ps_mul_nnz", value, other, out)


def add_row(value, rowptr, other, out=None):
    return _bc_row("ps_add_row", value, rowptr, other, out)


def add_col(value, col, other, out=None):
    return _bc_col("ps_add_col", value, col, other, out)


def add_nnz(value, other, out=None):
    return _bc_nnz("ps_add_nnz", value, other, out)


# ------------------------------------------------------------------ set operations (mul.py:35-73)
def coo_match(row, col, rowptr_b, col_b):
    """Position in the coalesced CSR (rowptr_b, col_b) of every entry (row[e], col[e]), -1 when absent,
    and the byte mask of the matched entries."""
    row, col, rowptr_b, col_b = _idx(row, "row"), _idx(col, "col"), _idx(rowptr_b, "rowptr"), _idx(col_b, "col")
    pos = _new(row, row.numel())
    mask = _new(row, row.numel(), torch.uint8)
    with torch.cuda.device(row.device):
        _lib.call("ps_coo_match", _p(row), _p(col), row.numel(), _p(rowptr_b), _p(col_b), rowptr_b.numel() - 1,
                  _p(pos), _p(mask), _stream(row))
    return pos, mask


# ------------------------------------------------------------------ resize / reshape (storage.py:281-368)
def pad_i64(src: torch.Tensor, total: int, fill: int) -> torch.Tensor:
    """src grown to `total` entries with `fill` in the new tail, or cut to `total` (sparse_resize)."""
    src = _idx(src, "src")
    out = _new(src, total)
    with torch.cuda.device(src.device):
        _lib.call("ps_pad_i64", _p(src), src.numel(), int(fill), int(total), _p(out), _stream(src))
    return out


def reshape_index(row: torch.Tensor, col: torch.Tensor, old_n: int, new_n: int):
    """(idx // new_n, idx % new_n) for idx = old_n * row + col (sparse_reshape)."""
    row, col = _idx(row, "row"), _idx(col, "col")
    out_row, out_col = _new(row, row.numel()), _new(row, row.numel())
    with torch.cuda.device(row.device):
        _lib.call("ps_reshape_index", _p(row), _p(col), row.numel(), int(old_n), int(new_n), _p(out_row), _p(out_col),
                  _stream(row))
    return out_row, out_col


# ------------------------------------------------------------------ cat (cat.py:49-276)
def concat_add(pieces, deltas, skip_first=None) -> torch.Tensor:
    """concat_k(pieces[k][skip_k:] + deltas[k]) in one launch (per 32 pieces); int64, 1-D."""
    pieces = [_idx(t, "piece") for t in pieces]
    K = len(pieces)
    skip = [0] * K if skip_first is None else [int(bool(x)) for x in skip_first]
    lens = [max(t.numel() - sk, 0) for t, sk in zip(pieces, skip)]
    out = _new(pieces[0], sum(lens))
    srcs = (ctypes.c_void_p * K)(*[t.data_ptr() + 8 * sk for t, sk in zip(pieces, skip)])
    c_lens = (ctypes.c_int64 * K)(*lens)
    c_deltas = (ctypes.c_int64 * K)(*[int(d) for d in deltas])
    with torch.cuda.device(out.device):
        _lib.call("ps_concat_add_i64", srcs, c_lens, c_deltas, K, _p(out), _stream(out))
    return out


def concat_rows(pieces) -> torch.Tensor:
    """Concatenation along dim 0 of same-dtype tensors with equal trailing shape (the value concat of cat.py)."""
    pieces = [_dev(t, "piece") for t in pieces]
    first = pieces[0]
    for t in pieces:
        if t.dtype != first.dtype or t.shape[1:] != first.shape[1:]:
            raise ValueError("concat_rows: dtypes / trailing shapes differ")
    if _ag.needs_grad(*pieces):
        return torch.cat(pieces, dim=0)
    K = len(pieces)
    out = torch.empty((sum(t.shape[0] for t in pieces),) + tuple(first.shape[1:]), dtype=first.dtype,
                      device=first.device)
    srcs = (ctypes.c_void_p * K)(*[t.data_ptr() for t in pieces])
    nbytes = (ctypes.c_int64 * K)(*[t.numel() * t.element_size() for t in pieces])
    with torch.cuda.device(out.device):
        _lib.call("ps_concat_bytes", srcs, nbytes, K, _p(out), _stream(out))
    return out


# ------------------------------------------------------------------ select building blocks
def _mask_u8(mask: torch.Tensor) -> torch.Tensor:
    mask = _dev(mask, "mask")
    if mask.dtype == torch.bool:
        return mask.view(torch.uint8)
    if mask.dtype == torch.uint8:
        return mask
    raise TypeError(f"mask must be bool, got {mask.dtype}")


def mask_to_index(mask: torch.Tensor) -> torch.Tensor:
    """Ascending indices of the set entries (masked_select.py:18-24)."""
    m = _mask_u8(mask)
    out = _new(m, m.numel())
    h = _host_i64()
    with torch.cuda.device(m.device):
        _lib.call("ps_mask_to_index", _p(m), m.numel(), _p(out), h, _stream(m))
    return out[: int(h[0])]


def mask_entries(mask: torch.Tensor, index: Optional[torch.Tensor], nnz: int) -> torch.Tensor:
    """Entry ids e with mask[index[e]] (index None: mask[e]) (masked_select.py:28-30, 60-62, 93-114)."""
    m = _mask_u8(mask)
    index = None if index is None else _idx(index, "index")
    out = _new(m, nnz)
    h = _host_i64()
    with torch.cuda.device(m.device):
        _lib.call("ps_mask_entries", _p(m), _p(index), nnz, _p(out), h, _stream(m))
    return out[: int(h[0])]


def slice_add(x: torch.Tensor, start: int, length: int, delta: int = 0) -> torch.Tensor:
    """x[start:start+length] + delta as a fresh tensor (the `paddle.slice(...) - offset` pairs of narrow.py)."""
    return concat_add([_idx(x, "x")[start:start + length]], [delta])


def narrow_cols(row: torch.Tensor, col: torch.Tensor, value: Optional[torch.Tensor], start: int, length: int):
    """narrow.py:53-62: the entries with start <= col < start+length as (row, col - start, value), order kept."""
    if _ag.needs_grad(value):
        out_row, out_col, _ = narrow_cols(row, col, None, start, length)
        sel = ((col >= int(start)) & (col < int(start) + int(length))).nonzero().squeeze(1)
        return out_row, out_col, gather_rows(value, sel)
    row, col = _idx(row, "row"), _idx(col, "col")
    value = None if value is None else _dev(value, "value")
    nnz = col.numel()
    out_row, out_col = _new(col, nnz), _new(col, nnz)
    out_value = None if value is None else torch.empty_like(value)
    h = _host_i64()
    with torch.cuda.device(col.device):
        _lib.call("ps_narrow_cols", _p(row), _p(col), _p(value),
                  0 if value is None else value.element_size() * max(_width(value), 1), nnz, int(start),
                  int(start) + int(length), _p(out_row), _p(out_col), _p(out_value), h, _stream(col))
    k = int(h[0])
    return out_row[:k], out_col[:k], None if out_value is None else out_value[:k]


def index_select_fill(old_ptr, minor, value, idx, out_ptr, E: int, want_perm: bool = False):
    """index_select.py:17-33 / 53-67 in one launch.  Returns (major, minor', value', perm|None)."""
    if _ag.needs_grad(value):
        major, new_minor, _, perm = index_select_fill(old_ptr, minor, None, idx, out_ptr, E, True)
        return major, new_minor, gather_rows(value, perm), (perm if want_perm else None)
    old_ptr, minor, idx, out_ptr = _idx(old_ptr), _idx(minor), _idx(idx), _idx(out_ptr)
    value = None if value is None else _dev(value, "value")
    K = idx.numel()
    out_major, out_minor = _new(minor, E), _new(minor, E)
    out_value = None if value is None else torch.empty((E,) + tuple(value.shape[1:]), dtype=value.dtype,
                                                       device=value.device)
    out_perm = _new(minor, E) if want_perm else None
    with torch.cuda.device(minor.device):
        _lib.call("ps_index_select_fill", _p(old_ptr), _p(minor), _p(value),
                  0 if value is None else value.element_size() * max(_width(value), 1), _p(idx), K, _p(out_ptr), E,
                  _p(out_major), _p(out_minor), _p(out_value), _p(out_perm), _stream(minor))
    return out_major, out_minor, out_value, out_perm


# ------------------------------------------------------------------ row-range partitioning
def dist_row_histogram(row: torch.Tensor, shift: int, buckets: int) -> torch.Tensor:
    row = _idx(row, "row")
    out = _new(row, buckets)
    with torch.cuda.device(row.device):
        _lib.call("ps_dist_row_histogram", _p(row), row.numel(), shift, buckets, _p(out), _stream(row))
    return out


def dist_partition(row, col, value, splitters: torch.Tensor):
    """Stable partition of the local tuples by destination rank.  Returns (row, col, value, counts[G])."""
    row, col, splitters = _idx(row, "row"), _idx(col, "col"), _idx(splitters, "splitters")
    value = None if value is None else _dev(value, "value")
    G = splitters.numel() - 1
    nnz = row.numel()
    out_row, out_col = _new(row, nnz), _new(row, nnz)
    out_value = None if value is None else torch.empty_like(value)
    counts = _new(row, G)
    with torch.cuda.device(row.device):
        _lib.call("ps_dist_partition", _p(row), _p(col), _p(value),
                  0 if value is None else value.element_size() * max(_width(value), 1), nnz, _p(splitters), G,
                  _p(out_row), _p(out_col), _p(out_value), _p(counts), _stream(row))
    return out_row, out_col, out_value, counts


def dist_dest(row: torch.Tensor, splitters: torch.Tensor):
    """Destination rank of every tuple (one byte each: at most 64 ranks) and the per-destination counts (device)."""
    row, splitters = _idx(row, "row"), _idx(splitters, "splitters")
    G = splitters.numel() - 1
    dest = torch.empty(row.numel(), dtype=torch.uint8, device=row.device)
    counts = _new(row, G)
    with torch.cuda.device(row.device):
        _lib.call("ps_dist_dest", _p(row), row.numel(), _p(splitters), G, _p(dest), _p(counts), _stream(row))
    return dest, counts


def dist_scatter_peer(dest, row, col, value, bases: torch.Tensor, peer_row: torch.Tensor, peer_col: torch.Tensor,
                      peer_value: Optional[torch.Tensor], narrow: bool = False) -> None:
    """The stable partition pass writing destination d's run at element bases[d] of the d-th buffer whose base
    address peer_row / peer_col / peer_value (int64 device tensors of addresses) hold: partition and exchange in
    one kernel over peer memory.  narrow: row / col are stored as 32-bit words (12 bytes per tuple on the wire)."""
    row, col = _idx(row, "row"), _idx(col, "col")
    value = None if value is None else _dev(value, "value")
    G = bases.numel()
    with torch.cuda.device(row.device):
        _lib.call("ps_dist_scatter_peer", _p(dest), _p(row), _p(col), _p(value),
                  0 if value is None else value.element_size() * max(_width(value), 1), row.numel(), G, _p(bases),
                  _p(peer_row), _p(peer_col), _p(peer_value), int(bool(narrow)), _stream(row))


def add_scalar(x: torch.Tensor, delta: int) -> torch.Tensor:
    x = _idx(x, "x")
    out = torch.empty_like(x)
    with torch.cuda.device(x.device):
        _lib.call("ps_add_scalar", _p(x), x.numel(), int(delta), _p(out), _stream(x))
    return out


# ------------------------------------------------------------------ synthetic inputs (bench / tests)
def synth_coo(seed: int, first: int, n: int, M: int, N: int, device, with_value: bool = True, rmat: bool = False):
    """SURVEY.md 8d counter-hash generator, on the device (uniform, or the R-MAT skewed variant).
    Returns (row, col, value|None)."""
    device = torch.device(device)
    row = torch.empty(n, dtype=torch.int64, device=device)
    col = torch.empty(n, dtype=torch.int64, device=device)
    value = torch.empty(n, dtype=torch.float32, device=device) if with_value else None
    with torch.cuda.device(device):
        _lib.call("ps_synth_rmat" if rmat else "ps_synth_coo", ctypes.c_uint64(seed & (2**64 - 1)), first, n, M, N,
                  _p(row), _p(col), _p(value), _stream(row))
    return row, col, value
