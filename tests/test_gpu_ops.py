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


def test_coalesce_empty_and_single(ops):
    e = dev(np.array([], I64))
    r, c, v = ops.coalesce(e, e, dev(np.array([], np.float32)), 5, 5)
    assert r.numel() == 0 and c.numel() == 0 and v.numel() == 0
    one = dev(np.array([3], I64))
    r, c, v = ops.coalesce(one, one, dev(np.array([2.5], np.float32)), 5, 5)
    assert host(r).tolist() == [3] and host(v).tolist() == [2.5]
