"""GPU parity tests of the operator layer (through the C ABI) against the CPU oracle.

Index outputs are compared bit-exactly; fp32 sums at 1e-6 relative, bf16 at 1e-3 relative,
min/max exactly (BASELINE.json north_star).  Golden vectors are the reference's own
(tests/golden/reference_vectors.py cites test/*.py)."""
import numpy as np
import pytest
import torch

from oracle import native, np_ref
from tests.golden import reference_vectors as G
from tests._util import DT, I64, REDUCES, _random_coo, assert_values, dev, host

pytestmark = pytest.mark.gpu


# ----------------------------------------------------------------------------- ind2ptr / ptr2ind
def test_ind2ptr_ptr2ind_golden(ops):
    g = G.IND2PTR
    rowptr = ops.ind2ptr(dev(np.array(g["row"], I64)), g["M"])
    assert host(rowptr).tolist() == g["rowptr"]
    assert host(ops.ptr2ind(rowptr, 6)).tolist() == g["row"]
    rowptr = ops.ind2ptr(dev(np.array([], I64)), 8)
    assert host(rowptr).tolist() == [0] * 9
    assert host(ops.ptr2ind(rowptr, 0)).tolist() == []


@pytest.mark.parametrize("M,nnz", [(1, 5), (7, 3), (100, 1000), (5000, 300), (3, 10000), (100000, 1000003),
                                   (1 << 20, 17), (50, 200001)])
def test_ind2ptr_ptr2ind_random(ops, M, nnz):
    rng = np.random.default_rng(M * 31 + nnz)
    ind = np.sort(rng.integers(0, M, nnz)).astype(I64)
    if nnz > 1000:  # a hub row and a long gap
        ind[: nnz // 3] = ind[nnz // 3]
    want = native.ind2ptr(ind, M)
    got = ops.ind2ptr(dev(ind), M)
    assert np.array_equal(host(got), want)
    assert np.array_equal(host(ops.ptr2ind(got, nnz)), native.ptr2ind(want, nnz))
    # unaligned base pointer (a view at an odd element offset)
    buf = dev(np.concatenate([[0], ind]).astype(I64))
    assert np.array_equal(host(ops.ind2ptr(buf[1:], M)), want)


@pytest.mark.parametrize("gap", [3, 40, 1500, 2040, 2047, 2048, 5000])
def test_ptr2ind_runs_of_empty_rows(ops, gap):
    """A run of `gap` empty rows inside a tile: up to a tile's worth of segment starts the expansion searches a
    sorted table (walking it, or searching per entry when the run lands inside one thread's entries); beyond that it
    falls back to scatter-of-starts + max-scan.  Same answers as the reference's ptr2ind_cpu either way."""
    rng = np.random.default_rng(gap)
    for lead in (0, 1000, 2047, 4100):
        row = np.concatenate([np.zeros(lead, I64), np.full(3000, gap + 1, I64),
                              np.sort(rng.integers(gap + 1, gap + 60, 7000)).astype(I64)])
        M = gap + 70
        ptr = native.ind2ptr(row, M)
        assert np.array_equal(host(ops.ptr2ind(dev(ptr), row.size)), row)
        v = dev(rng.integers(1, 9, row.size).astype(np.float32))
        o = dev(rng.integers(1, 9, M).astype(np.float32))
        assert np.array_equal(host(ops.mul_row(v, dev(ptr), o)), host(v) * host(o)[row])
        assert np.array_equal(host(ops.gather_csr(o, dev(ptr), row.size)), host(o)[row])


def test_ptr_count_helpers(ops):
    rng = np.random.default_rng(5)
    cnt = rng.integers(0, 9, 100001).astype(I64)
    ptr, tot = ops.count2ptr(dev(cnt), want_total=True)
    want = np.concatenate([[0], np.cumsum(cnt)])
    assert np.array_equal(host(ptr), want) and tot == want[-1]
    assert np.array_equal(host(ops.ptr2count(ptr)), cnt)
    idx = rng.integers(0, cnt.size, 5000).astype(I64)
    assert np.array_equal(host(ops.count2ptr(dev(cnt), dev(idx))), np.concatenate([[0], np.cumsum(cnt[idx])]))
    col = rng.integers(0, 777, 123457).astype(I64)
    assert np.array_equal(host(ops.scatter_add_ones(dev(col), 777)), np.bincount(col, minlength=777))
    assert ops.index_max(dev(col), dev(cnt)) == (int(col.max()), int(cnt.max()))
    assert ops.index_max(dev(np.array([], I64))) == (-1, -1)


def test_coo_order(ops):
    r = dev(np.array([0, 0, 1, 1], I64))
    assert ops.coo_order(r, dev(np.array([0, 1, 0, 1], I64))) == (True, True)
    assert ops.coo_order(r, dev(np.array([0, 0, 0, 1], I64))) == (True, False)
    assert ops.coo_order(r, dev(np.array([1, 0, 0, 1], I64))) == (False, False)
    assert ops.coo_order(dev(np.array([], I64)), dev(np.array([], I64))) == (True, True)


# ----------------------------------------------------------------------------- sort / coalesce
@pytest.mark.parametrize("M,N,nnz", [(2, 2, 4), (10, 7, 50), (1000, 1000, 20000), (10000, 10000, 100000),
                                     (70000, 3, 50000), (3000000, 3000000, 200000)])
def test_sort_coo_matches_oracle(ops, M, N, nnz):
    rng = np.random.default_rng(nnz)
    row, col = _random_coo(rng, M, N, nnz)
    value = rng.standard_normal(nnz).astype(np.float32)
    st = np_ref.SparseStorage(row=row, col=col, value=value, sparse_sizes=(M, N))
    r, c, v, perm = ops.sort_coo(dev(row), dev(col), dev(value), M, N, want_perm=True)
    assert np.array_equal(host(r), st.row()) and np.array_equal(host(c), st.col())
    assert np.array_equal(host(v), st.value())  # stable sort: identical tie order
    assert np.array_equal(host(perm), np.argsort(row * N + col, kind="stable"))
    r2, c2, v2, _ = ops.sort_coo(dev(row), dev(col), dev(value), M, N)
    assert np.array_equal(host(r2), st.row()) and np.array_equal(host(v2), st.value())
    # wide values ride through the permutation path
    wide = rng.standard_normal((nnz, 3)).astype(np.float64)
    _, _, v3, _ = ops.sort_coo(dev(row), dev(col), dev(wide), M, N)
    assert np.array_equal(host(v3), wide[np.argsort(row * N + col, kind="stable")])


def test_sort_on_construct_golden(ops):
    g = G.SORT_ON_CONSTRUCT
    r, c, v, _ = ops.sort_coo(dev(np.array(g["row"], I64)), dev(np.array(g["col"], I64)),
                              dev(np.array(g["value"], np.float32)), 2, 2)
    assert host(r).tolist() == g["out_row"] and host(c).tolist() == g["out_col"]
    assert host(v).tolist() == g["out_value"]


@pytest.mark.parametrize("dtype", DT)
@pytest.mark.parametrize("reduce", REDUCES)
def test_coalesce_matches_oracle(ops, dtype, reduce):
    rng = np.random.default_rng(11)
    M, N, nnz = 10000, 10000, 100000  # BASELINE.json configs[0]
    row, col = _random_coo(rng, M, N, nnz, 0.2)
    if dtype in (torch.int32, torch.int64):
        val = rng.integers(-50, 50, nnz)
    else:
        val = rng.random(nnz).astype(np.float32)
    tv = dev(val, dtype)
    ref_val = host(tv)  # the values as the device sees them (bf16/f16 rounded)
    idx, want = np_ref.coalesce(np.stack([row, col]), ref_val.astype(np.float64) if dtype in (
        torch.float16, torch.bfloat16) else ref_val, M, N, reduce)
    r, c, v = ops.coalesce(dev(row), dev(col), tv, M, N, reduce)
    assert np.array_equal(host(r), idx[0]) and np.array_equal(host(c), idx[1])
    exact = reduce in ("min", "max") or dtype in (torch.int32, torch.int64)
    if dtype in (torch.int32, torch.int64) and reduce == "mean":
        cnt = np_ref.coalesce(np.stack([row, col]), np.ones(nnz, I64), M, N, "sum")[1]
        tot = np_ref.coalesce(np.stack([row, col]), ref_val, M, N, "sum")[1]
        want = np.trunc(tot / cnt).astype(tot.dtype)  # segment_csr mean: C++ integer division
    assert_values(host(v), want, dtype, exact)


def test_coalesce_golden_2d(ops):
    g = G.COALESCE
    row, col = dev(np.array(g["row"], I64)), dev(np.array(g["col"], I64))
    r, c, v = ops.coalesce(row, col, None, g["m"], g["n"])
    assert [host(r).tolist(), host(c).tolist()] == g["index"]
    val = dev(np.array(g["value"], np.float32))
    r, c, v = ops.coalesce(row, col, val, g["m"], g["n"], "add")
    assert [host(r).tolist(), host(c).tolist()] == g["index"] and host(v).tolist() == g["add"]
    r, c, v = ops.coalesce(row, col, val, g["m"], g["n"], "max")
    assert host(v).tolist() == g["max"]


@pytest.mark.parametrize("reduce", ["sum", "mean", "min", "max"])
def test_coalesce_sorted_and_heavy_duplicates(ops, reduce):
    rng = np.random.default_rng(3)
    # few distinct keys: runs of thousands of duplicates cross tile boundaries
    row = rng.integers(0, 3, 50000).astype(I64)
    col = rng.integers(0, 3, 50000).astype(I64)
    val = rng.random(50000).astype(np.float32)
    st = np_ref.SparseStorage(row=row, col=col, value=val, sparse_sizes=(3, 3))
    want = st.coalesce(reduce)
    r, c, v = ops.coalesce_sorted(dev(st.row()), dev(st.col()), dev(st.value()), reduce)
    assert np.array_equal(host(r), want.row()) and np.array_equal(host(c), want.col())
    np.testing.assert_allclose(host(v), want.value(), rtol=2e-5 if reduce in ("sum", "mean") else 0)
    r, c, v = ops.coalesce(dev(row), dev(col), dev(val), 3, 3, reduce)
    assert np.array_equal(host(r), want.row())
    np.testing.assert_allclose(host(v), want.value(), rtol=2e-5 if reduce in ("sum", "mean") else 0)


def test_coalesce_sorted_refuses_ids_beyond_the_packed_comparison(ops):
    """Entries are compared as row << 32 | col: sparse sizes beyond 2^32 are refused, not truncated; with explicit
    sizes the result is the same as with sizes taken from the data."""
    row, col = dev(np.array([0, 0, 1, 1, 1], I64)), dev(np.array([2, 2, 0, 3, 3], I64))
    val = dev(np.array([1, 2, 3, 4, 5], np.float32))
    a = ops.coalesce_sorted(row, col, val, "add")
    b = ops.coalesce_sorted(row, col, val, "add", (1 << 32, 1 << 32))
    assert host(a[0]).tolist() == host(b[0]).tolist() == [0, 1, 1]
    assert host(a[1]).tolist() == host(b[1]).tolist() == [2, 0, 3]
    assert host(a[2]).tolist() == host(b[2]).tolist() == [3.0, 3.0, 9.0]
    for sizes in (((1 << 32) + 1, 4), (4, 1 << 33)):
        with pytest.raises(RuntimeError, match=r"2\^32"):
            ops.coalesce_sorted(row, col, val, "add", sizes)


@pytest.mark.parametrize("reduce", ["add", "mean", "max"])
@pytest.mark.parametrize("dtype", [torch.float32, torch.int32])
def test_coalesce_long_runs_and_row_gaps(ops, reduce, dtype):
    """The LSD pipeline's dedup kernel (32-bit keys): one key repeated 300,000 times - a run that crosses a hundred
    tiles is reduced tile by tile and folded by dedup_cross_kernel, not walked by one thread - runs that end exactly at
    tile boundaries, and rows tens of thousands apart (long rowptr gaps are filled by the whole CTA)."""
    rng = np.random.default_rng(77)
    M, N = 60000, 50000
    row = np.concatenate([np.full(300000, 7), np.full(3072 * 3, 41000), np.full(5000, M - 1),
                          rng.integers(100, 200, 20000)]).astype(I64)
    col = np.concatenate([np.full(300000, 11), np.full(3072 * 3, 3), rng.integers(0, 4, 5000),
                          rng.integers(0, N, 20000)]).astype(I64)
    p = rng.permutation(row.size)
    row, col = row[p], col[p]
    val = rng.integers(-3, 4, row.size)
    r, c, v, rowptr = ops.coalesce(dev(row), dev(col), dev(val, dtype), M, N, reduce, True)
    idx, want = np_ref.coalesce(np.stack([row, col]), val.astype(np.float64), M, N, reduce)
    assert np.array_equal(host(r), idx[0]) and np.array_equal(host(c), idx[1])
    assert np.array_equal(host(rowptr), np.searchsorted(idx[0], np.arange(M + 1), side="left"))
    got = host(v.to(torch.float64))
    if dtype == torch.int32 and reduce == "mean":
        cnt = np_ref.coalesce(np.stack([row, col]), np.ones(row.size), M, N, "sum")[1]
        tot = np_ref.coalesce(np.stack([row, col]), val.astype(np.float64), M, N, "sum")[1]
        want = np.trunc(tot / cnt)
    if reduce == "mean" and dtype == torch.float32:
        np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-7)
    else:
        assert np.array_equal(got, want)
    # the same through the sorted-input entry point
    st = np_ref.SparseStorage(row=row, col=col, value=val.astype(np.float64), sparse_sizes=(M, N))
    r2, c2, v2 = ops.coalesce_sorted(dev(st.row()), dev(st.col()), dev(st.value(), dtype), reduce, (M, N))
    assert np.array_equal(host(r2), idx[0]) and np.array_equal(host(c2), idx[1])
    assert np.array_equal(host(v2.to(torch.float64)), got)


def test_coalesce_empty_and_single(ops):
    e = dev(np.array([], I64))
    r, c, v = ops.coalesce(e, e, dev(np.array([], np.float32)), 5, 5)
    assert r.numel() == 0 and c.numel() == 0 and v.numel() == 0
    one = dev(np.array([3], I64))
    r, c, v = ops.coalesce(one, one, dev(np.array([2.5], np.float32)), 5, 5)
    assert host(r).tolist() == [3] and host(v).tolist() == [2.5]


@pytest.mark.parametrize("M,N,nnz", [(1, 1, 1), (7, 5, 3), (10000, 10000, 100000), (3, 3, 50000),
                                     (2500000, 1000, 40000), (300, 2500000, 200000)])
@pytest.mark.parametrize("with_value", [True, False])
def test_coalesce_emits_rowptr(ops, M, N, nnz, with_value):
    """ps_coalesce's out_rowptr == ind2ptr(row', M) (storage.py:211-222), including empty leading / trailing /
    interior rows (M far larger than the number of occupied rows) and rows spanning many tiles."""
    rng = np.random.default_rng(M + nnz)
    row, col = _random_coo(rng, M, N, nnz, 0.2)
    val = rng.random(nnz).astype(np.float32) if with_value else None
    r, c, v, rowptr = ops.coalesce(dev(row), dev(col), None if val is None else dev(val), M, N, "add", True)
    idx, want = np_ref.coalesce(np.stack([row, col]), val, M, N, "add")
    assert np.array_equal(host(r), idx[0]) and np.array_equal(host(c), idx[1])
    assert np.array_equal(host(rowptr), np.searchsorted(idx[0], np.arange(M + 1), side="left"))
    if with_value:
        np.testing.assert_allclose(host(v), want, rtol=2e-5)
    e = dev(np.array([], I64))
    assert host(ops.coalesce(e, e, None, 4, 4, "add", True)[3]).tolist() == [0] * 5


def test_lazy_sort_then_coalesce_is_fused_and_equivalent():
    """SparseStorage defers sort-on-construct (storage.py:158-171) to the first accessor; coalesce() right after the
    constructor takes the fused path.  Both orders of use must give the reference's results, and coalesce() must
    return the SAME object when nothing is merged (storage.py:459-460)."""
    import paddle_sparse_b200 as ps
    rng = np.random.default_rng(21)
    M = N = 500
    row, col = _random_coo(rng, M, N, 4000, 0.3)
    val = rng.random(4000).astype(np.float32)
    want = np_ref.SparseStorage(row=row, col=col, value=val, sparse_sizes=(M, N))
    wc = want.coalesce("add")
    a = ps.SparseStorage(row=dev(row), col=dev(col), value=dev(val), sparse_sizes=(M, N))
    ca = a.coalesce("add")  # fused: the constructor has not sorted yet
    assert ca is not a
    assert np.array_equal(host(ca.row()), wc.row()) and np.array_equal(host(ca.col()), wc.col())
    np.testing.assert_allclose(host(ca.value()), wc.value(), rtol=2e-6)
    assert np.array_equal(host(ca.rowptr()), wc.rowptr()) and ca.cached_keys() == []
    # the original object still behaves like the reference's sorted, uncoalesced storage
    assert np.array_equal(host(a.row()), want.row()) and np.array_equal(host(a.col()), want.col())
    assert np.array_equal(host(a.value()), want.value())
    b = ps.SparseStorage(row=dev(row), col=dev(col), value=dev(val), sparse_sizes=(M, N))
    assert np.array_equal(host(b.col()), want.col())  # accessor first: sorts, then the sorted-input path
    cb = b.coalesce("add")
    assert np.array_equal(host(cb.col()), wc.col())
    # nothing to merge: same object comes back, now sorted
    key = np.unique(row * N + col)
    perm = rng.permutation(key.size)
    r2, c2 = (key // N)[perm].astype(I64), (key % N)[perm].astype(I64)
    d = ps.SparseStorage(row=dev(r2), col=dev(c2), sparse_sizes=(M, N))
    assert d.coalesce() is d
    assert np.array_equal(host(d.row()), key // N) and np.array_equal(host(d.col()), key % N)
    assert np.array_equal(host(d.rowptr()), np.searchsorted(key // N, np.arange(M + 1)))


@pytest.mark.parametrize("M,N,nnz,dup", [(2449029, 2449029, 70000, 0.2), (2449029, 2449029, 1500000, 0.3),
                                         (3000000, 1 << 22, 400000, 0.0), (1 << 23, 1 << 23, 3000000, 0.05),
                                         (70000, 2000000, 250000, 0.5),
                                         # 51- and 53-bit keys (row shards of papers100M): the wide hybrid
                                         (13882495, 111059956, 2500000, 0.2), (1 << 26, 1 << 27, 300000, 0.3)])
@pytest.mark.parametrize("dtype,reduce", [(torch.float32, "add"), (torch.float32, "max"), (torch.int32, "mean"),
                                          (torch.bfloat16, "add"), (None, "add")])
def test_hybrid_coalesce_matches_oracle(ops, M, N, nnz, dup, dtype, reduce):
    """keys wider than 32 bits and >= 2^16 entries take the hybrid path (local_sort.cuh: top-bit buckets by
    one-sweep passes, bucket-local sort + dedup + rowptr in shared memory); same results as the oracle."""
    rng = np.random.default_rng(nnz)
    row, col = _random_coo(rng, M, N, nnz, dup)
    val = None if dtype is None else rng.integers(-40, 40, nnz)
    tv = None if dtype is None else dev(val, dtype)
    r, c, v, rowptr = ops.coalesce(dev(row), dev(col), tv, M, N, reduce, True)
    idx, want = np_ref.coalesce(np.stack([row, col]), None if val is None else val.astype(np.float64), M, N, reduce)
    assert np.array_equal(host(r), idx[0]) and np.array_equal(host(c), idx[1])
    assert np.array_equal(host(rowptr), np.searchsorted(idx[0], np.arange(M + 1), side="left"))
    if dtype is not None:
        if dtype == torch.int32 and reduce == "mean":
            cnt = np_ref.coalesce(np.stack([row, col]), np.ones(nnz), M, N, "sum")[1]
            tot = np_ref.coalesce(np.stack([row, col]), val.astype(np.float64), M, N, "sum")[1]
            want = np.trunc(tot / cnt)
        got = host(v.to(torch.float64))
        if dtype == torch.bfloat16:
            np.testing.assert_allclose(got, want, rtol=8e-3, atol=1e-6)
        else:
            assert np.array_equal(got, want)


@pytest.mark.parametrize("M,N,nnz,dup", [(5000, 60000, 30000, 0.3),           # 29-bit keys: 32-bit LSD sort
                                         (2449029, 2449029, 40000, 0.2),      # 44-bit keys, small: 64-bit LSD sort
                                         (2449029, 2449029, 900000, 0.3),     # 44-bit keys: the wide hybrid
                                         (13882495, 111059956, 1500000, 0.2)])  # 51-bit keys (a papers100M row shard)
@pytest.mark.parametrize("dtype,reduce", [(torch.float32, "add"), (torch.int32, "max"), (None, "add")])
def test_coalesce_u32_matches_coalesce(ops, M, N, nnz, dup, dtype, reduce):
    """The form the fused peer exchange feeds: 32-bit GLOBAL row ids with the shard's first row as offset, 32-bit
    columns.  Same outputs, bit for bit, as `coalesce` on the widened local arrays (which the tests above hold against
    the oracle)."""
    rng = np.random.default_rng(nnz + 7)
    row, col = _random_coo(rng, M, N, nnz, dup)
    off = 1234567
    val = None if dtype is None else dev(rng.integers(-40, 40, nnz), dtype)
    want = ops.coalesce(dev(row), dev(col), val, M, N, reduce, True)
    row32 = torch.from_numpy((row + off).astype(np.uint32).view(np.int32)).cuda()
    col32 = torch.from_numpy(col.astype(np.uint32).view(np.int32)).cuda()
    got = ops.coalesce_u32(row32, col32, off, val, M, N, reduce, True)
    assert torch.equal(got[0], want[0]) and torch.equal(got[1], want[1]) and torch.equal(got[3], want[3])
    if dtype is not None:
        assert torch.equal(got[2], want[2])
    idx, _ = np_ref.coalesce(np.stack([row, col]), None, M, N, reduce)
    assert np.array_equal(host(got[0]), idx[0]) and np.array_equal(host(got[1]), idx[1])
    with pytest.raises(RuntimeError, match=r"2\^32"):
        ops.coalesce_u32(row32, col32, (1 << 32) - 5, val, M, N, reduce, True)


@pytest.mark.parametrize("nnz", [300000, 3000000])
@pytest.mark.parametrize("dtype,reduce", [(torch.float32, "add"), (torch.int32, "max"), (torch.bfloat16, "add"),
                                          (None, "add")])
def test_hybrid_coalesce_skewed_rows_are_refined(ops, nnz, dtype, reduce):
    """Power-law rows (the R-MAT stream of the bench, and hand-made hubs on top of it): buckets beyond a CTA's
    capacity are re-partitioned level by level (refine.cuh) instead of sending the whole input to the LSD sort.  Row,
    col, rowptr bit-exact against the oracle, values exact (small integers)."""
    M = N = 2449029
    row, col, _ = ops.synth_coo(0x5EED0002, 0, nnz, M, N, "cuda", with_value=False, rmat=True)
    row, col = host(row).copy(), host(col).copy()
    rng = np.random.default_rng(nnz)
    row[: nnz // 10] = 77                      # one row holding 10 % of the entries ...
    col[: nnz // 40] = rng.integers(5, 9, nnz // 40)   # ... a quarter of them on four columns (thousands of duplicates,
    row[nnz // 10: nnz // 10 + 5000] = M - 1   # but below the 4,096 a single key may have); the last row is a hub too
    col[nnz // 10: nnz // 10 + 5000] = rng.integers(0, N, 5000)
    p = rng.permutation(nnz)
    row, col = row[p], col[p]
    val = None if dtype is None else rng.integers(-40, 40, nnz)
    tv = None if dtype is None else dev(val, dtype)
    r, c, v, rowptr = ops.coalesce(dev(row), dev(col), tv, M, N, reduce, True)
    idx, want = np_ref.coalesce(np.stack([row, col]), None if val is None else val.astype(np.float64), M, N, reduce)
    assert np.array_equal(host(r), idx[0]) and np.array_equal(host(c), idx[1])
    assert np.array_equal(host(rowptr), np.searchsorted(idx[0], np.arange(M + 1), side="left"))
    if dtype is not None:
        got = host(v.to(torch.float64))
        if dtype == torch.bfloat16:
            np.testing.assert_allclose(got, want, rtol=8e-3, atol=1e-6)
        else:
            assert np.array_equal(got, want)
    # sort-on-construct on the same stream: stable (duplicate keys carry distinct values)
    if dtype == torch.float32:
        order = np.argsort(row * N + col, kind="stable")
        sv = (np.arange(nnz) % 251).astype(np.float32)
        r2, c2, v2, _ = ops.sort_coo(dev(row), dev(col), dev(sv), M, N)
        assert np.array_equal(host(r2), row[order]) and np.array_equal(host(c2), col[order])
        assert np.array_equal(host(v2), sv[order])


def test_hybrid_coalesce_one_key_beyond_a_bucket_falls_back(ops):
    """A single (row, col) with more duplicates than a bucket holds cannot be split by key bits: the LSD sort takes
    over, with the same results."""
    rng = np.random.default_rng(12)
    M = N = 2449029
    row = np.concatenate([rng.integers(0, M, 200000), np.full(9000, 1234)]).astype(I64)
    col = np.concatenate([rng.integers(0, N, 200000), np.full(9000, 4321)]).astype(I64)
    p = rng.permutation(row.size)
    row, col = row[p], col[p]
    val = rng.integers(-5, 6, row.size).astype(np.float32)
    r, c, v, rowptr = ops.coalesce(dev(row), dev(col), dev(val), M, N, "add", True)
    idx, want = np_ref.coalesce(np.stack([row, col]), val.astype(np.float64), M, N, "add")
    assert np.array_equal(host(r), idx[0]) and np.array_equal(host(c), idx[1])
    assert np.array_equal(host(v).astype(np.float64), want)
    assert np.array_equal(host(rowptr), np.searchsorted(idx[0], np.arange(M + 1), side="left"))


def test_hybrid_coalesce_hub_rows_fall_back(ops):
    """one bucket beyond the shared-memory capacity (a hub row range): the plain LSD pipeline takes over"""
    rng = np.random.default_rng(9)
    M = N = 2449029
    row = np.concatenate([rng.integers(0, M, 100000), rng.integers(5, 9, 60000)]).astype(I64)
    col = rng.integers(0, N, row.size).astype(I64)
    val = rng.random(row.size).astype(np.float32)
    r, c, v, rowptr = ops.coalesce(dev(row), dev(col), dev(val), M, N, "add", True)
    idx, want = np_ref.coalesce(np.stack([row, col]), val, M, N, "add")
    assert np.array_equal(host(r), idx[0]) and np.array_equal(host(c), idx[1])
    np.testing.assert_allclose(host(v), want, rtol=2e-6)
    assert np.array_equal(host(rowptr), np.searchsorted(idx[0], np.arange(M + 1), side="left"))


@pytest.mark.parametrize("M,N,nnz,dup", [(2449029, 2449029, 500000, 0.3), (1 << 23, 1 << 22, 2000000, 0.1),
                                         (13882495, 111059956, 1000000, 0.3)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, None])
def test_hybrid_sort_coo_is_stable(ops, M, N, nnz, dup, dtype):
    """sort-on-construct (storage.py:158-171) on the hybrid path: duplicate keys carry distinct values, so any
    instability of the bucket-local passes or of the atomic ranking shows up as a value mismatch."""
    rng = np.random.default_rng(nnz + 1)
    row, col = _random_coo(rng, M, N, nnz, dup)
    order = np.argsort(row * N + col, kind="stable")
    val = None if dtype is None else np.arange(nnz) % 251
    tv = None if dtype is None else dev(val, dtype)
    r, c, v, _ = ops.sort_coo(dev(row), dev(col), tv, M, N)
    assert np.array_equal(host(r), row[order]) and np.array_equal(host(c), col[order])
    if dtype is not None:
        assert np.array_equal(host(v.to(torch.float64)), val[order].astype(np.float64))


# ----------------------------------------------------------------------------- broadcast mul / add on [nnz] values
@pytest.mark.parametrize("dtype", DT)
@pytest.mark.parametrize("shape", ["long rows", "short rows", "many empty rows"])
def test_broadcast_row_col_vector_paths(ops, dtype, shape):
    """mul.py:76-93 / add.py:50-70 on width-1 values: the 128-bit kernels (aligned), the element-wise ones (a view
    at an odd element offset), the partial last tile, in place and out of place.  Small integers keep every dtype
    exact, so the comparison with NumPy (gather_csr / other[col] then * or +) is bit for bit."""
    rng = np.random.default_rng(len(shape) + 7)
    M, N = {"long rows": (90, 4000), "short rows": (30000, 700), "many empty rows": (400000, 50)}[shape]
    nnz = 100003
    row = np.sort(rng.integers(0, M, nnz)).astype(I64)
    if shape == "many empty rows":
        row = np.sort(rng.choice(np.arange(0, M, 97), nnz)).astype(I64)
    col = rng.integers(0, N, nnz).astype(I64)
    rowptr = native.ind2ptr(row, M)
    val = rng.integers(1, 12, nnz + 1).astype(np.float64)
    orow = rng.integers(1, 12, M).astype(np.float64)
    ocol = rng.integers(1, 12, N).astype(np.float64)
    d_ptr, d_col, d_orow, d_ocol = dev(rowptr), dev(col), dev(orow, dtype), dev(ocol, dtype)
    for off in (0, 1):  # off = 1: base pointers that are not 16-byte aligned
        n = nnz - off
        buf = dev(val, dtype)
        v = buf[off:off + n]
        r, c = row[off:off + n], col[off:off + n]
        ptr_np = native.ind2ptr(r, M)
        p = dev(ptr_np)
        cc = dev(np.concatenate([[0], c]).astype(I64))[1:] if off else dev(c)
        x = val[off:off + n]
        assert np.array_equal(host(ops.mul_row(v, p, d_orow)).astype(np.float64), x * orow[r])
        assert np.array_equal(host(ops.add_row(v, p, d_orow)).astype(np.float64), x + orow[r])
        assert np.array_equal(host(ops.mul_col(v, cc, d_ocol)).astype(np.float64), x * ocol[c])
        assert np.array_equal(host(ops.add_col(v, cc, d_ocol)).astype(np.float64), x + ocol[c])
        assert np.array_equal(host(v).astype(np.float64), x)  # inputs untouched
        w = v.clone() if off == 0 else dev(val, dtype)[off:off + n]
        ops.mul_row(w, p, d_orow, out=w)
        assert np.array_equal(host(w).astype(np.float64), x * orow[r])
        w = v.clone() if off == 0 else dev(val, dtype)[off:off + n]
        ops.mul_col(w, cc, d_ocol, out=w)
        assert np.array_equal(host(w).astype(np.float64), x * ocol[c])
    del d_ptr, d_col


def test_atomic_rank_order_selftest(ops):
    """Stability of every sort here rests on ATOMS serving the colliding lanes of one warp instruction in lane order
    (radix_sort.cuh phase 2): re-checked on THIS device, for both counter layouts, under every collision pattern."""
    import ctypes
    h = (ctypes.c_int64 * 3)()
    ops._lib.call("ps_selftest_atom_order", h, None)
    checked, collided, bad = int(h[0]), int(h[1]), int(h[2])
    assert checked >= 10 ** 9 and collided > checked // 10
    assert ops._lib.lib().ps_rank_mode() == 0  # the only ranking mode this build has
    assert bad == 0, f"{bad} of {collided} colliding lane-instances were served out of lane order: the sorts are not stable here"


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_scatter_min_max_with_negative_zero(ops, dtype):
    """-0.0 must order with the negatives in the bit-pattern atomics of the column reductions (round 2: `v >= 0`
    sent it down the signed path, where it stuck as the minimum)."""
    rng = np.random.default_rng(8)
    n, S = 200_000, 300
    v = np.round(rng.standard_normal(n) * 2) / 2  # multiples of 0.5: many exact +0.0 / -0.0
    assert np.signbit(v[v == 0]).any() and (~np.signbit(v[v == 0])).any()
    idx = rng.integers(0, S, n).astype(I64)
    src = dev(v, dtype)
    for red, fn in (("min", np.minimum), ("max", np.maximum)):
        want = np.full(S, np.inf if red == "min" else -np.inf)
        fn.at(want, idx, v)
        got = host(ops.scatter(src, dev(idx), 0, None, S, red)).astype(np.float64)
        assert np.array_equal(got, want), red  # +0.0 == -0.0 compares equal; everything else must match exactly


def test_segment_csr_tile_kernel_variant(ops):
    """PS_SEG_TILE=1 (the entry-balanced tile kernel, measured and not the default) in a fresh process: same results as
    the default kernels on long rows, hubs, empty segments, a ptr that does not cover src, all reductions."""
    import os
    import subprocess
    import sys
    code = r"""
import os
import numpy as np, torch
from paddle_sparse_b200 import ops
rng = np.random.default_rng(5)
S, E = 3000, 1 << 20
lens = rng.poisson(300, S).astype(np.int64)
lens[rng.integers(0, S, 200)] = 0
lens[17] = 90000
lens[S - 1] = 5
lens = (lens * ((E - 4000) / lens.sum())).astype(np.int64)
ptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64) + 1500   # entries before ptr[0] / after ptr[S]: nobody's
v = (rng.integers(-8, 9, E) / 4.0).astype(np.float32)
for dt in (torch.float32, torch.bfloat16, torch.int32):
    src = torch.from_numpy(v if dt != torch.int32 else (v * 4).astype(np.int32)).cuda().to(dt)
    for red in ("sum", "mean", "min", "max"):
        got = ops.segment_csr(src, torch.from_numpy(ptr).cuda(), None, red)
        torch.save(got.cpu(), os.environ["OUT"] + f"_{str(dt)[6:]}_{red}.pt")
"""
    import tempfile
    outs = {}
    with tempfile.TemporaryDirectory() as d:
        for flag in ("0", "1"):
            env = dict(os.environ, PS_SEG_TILE=flag, OUT=os.path.join(d, "o" + flag))
            subprocess.run([sys.executable, "-c", code], check=True, env=env, cwd=os.path.dirname(os.path.dirname(__file__)),
                           capture_output=True)
        for name in sorted(os.listdir(d)):
            outs.setdefault(name[2:], {})[name[1]] = torch.load(os.path.join(d, name))
    assert len(outs) == 12
    for name, pair in outs.items():
        a, b = pair["0"].double(), pair["1"].double()
        if "bfloat16" in name and ("sum" in name or "mean" in name):
            assert torch.allclose(a, b, rtol=8e-3, atol=0.26), name
        else:
            assert torch.equal(a, b), name  # exactly representable values: order-independent


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.int64])
@pytest.mark.parametrize("width", [1, 3])
def test_segment_csr_long_segments(ops, dtype, width):
    """Hub rows: one segment holding 30 % of the entries, a few of several thousand, many short and empty ones -
    the long ones are handed to whole CTAs (segment_reduce_long_kernel), every result must still equal the oracle."""
    rng = np.random.default_rng(31 + width)
    S, E = 20_000, 600_000
    lens = rng.poisson(8, S).astype(I64)
    lens[rng.integers(0, S, S // 20)] = 0
    lens[123] = int(0.3 * E)
    lens[[7, 4000, S - 1]] = [5000, 129, 9000]
    lens = (lens * (E / lens.sum())).astype(I64)  # keep the proportions, hit roughly E entries
    ptr = np.concatenate([[0], np.cumsum(lens)]).astype(I64)
    n = int(ptr[-1])
    shape = (n,) if width == 1 else (n, width)
    if dtype == torch.int64:
        v = rng.integers(-50, 50, shape).astype(np.int64)
    else:
        v = (rng.integers(-8, 9, shape) / 4.0).astype(np.float32)  # exactly representable, exact sums
    src = dev(v, dtype)
    for red in REDUCES:
        got = host(ops.segment_csr(src, dev(ptr), None, red)).astype(np.float64)
        want = np_ref.segment_csr(v.astype(np.float64), ptr, red).astype(np.float64)
        if dtype == torch.int64 and red == "mean":
            want = np_ref.segment_csr(v, ptr, red).astype(np.float64)
        if red == "mean" and dtype != torch.int64:
            rtol = 8e-3 if dtype == torch.bfloat16 else 1e-6
            np.testing.assert_allclose(got, want, rtol=rtol, atol=1e-7, err_msg=red)
        elif dtype == torch.bfloat16 and red in ("sum", "add"):
            np.testing.assert_allclose(got, want, rtol=8e-3, atol=0.26, err_msg=red)  # one bf16 rounding of an exact sum
        else:
            assert np.array_equal(got, want), red
    if width == 1:  # a source that is not 16-byte aligned takes the element-wise loads: same results
        buf = torch.cat([src[:1], src])
        for red in ("sum", "max"):
            assert torch.equal(ops.segment_csr(buf[1:], dev(ptr), None, red), ops.segment_csr(src, dev(ptr), None, red))
