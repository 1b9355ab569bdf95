"""GPU parity tests: CSC structures, transpose, paddle_scatter replacements, broadcast mul,
select building blocks, row-range partitioning — operator layer vs the CPU oracle."""
import numpy as np
import pytest
import torch

from oracle import np_ref
from tests.golden import reference_vectors as G
from tests._util import DT, I64, _random_coo, assert_values, dev, host

pytestmark = pytest.mark.gpu


def _coalesced(rng, M, N, nnz):
    row, col = _random_coo(rng, M, N, nnz, 0.1)
    key = np.unique(row * N + col)
    return (key // N).astype(I64), (key % N).astype(I64)


# ----------------------------------------------------------------------------- csr2csc / colptr / csc2csr / t
def test_caching_golden(ops):
    g = G.CACHING
    row, col = dev(np.array(g["row"], I64)), dev(np.array(g["col"], I64))
    perm, colptr, colcount = ops.csr2csc(col, 2, True, True)
    assert host(perm).tolist() == g["csr2csc"] and host(colptr).tolist() == g["colptr"]
    assert host(colcount).tolist() == g["colcount"]
    assert host(ops.invert_perm(perm)).tolist() == g["csc2csr"]
    rowptr = ops.ind2ptr(row, 2)
    assert host(rowptr).tolist() == g["rowptr"] and host(ops.ptr2count(rowptr)).tolist() == g["rowcount"]


@pytest.mark.parametrize("M,N,nnz", [(1, 1, 1), (5, 300, 400), (1000, 1000, 30000), (10000, 10000, 100000),
                                     (200, 70000, 90000), (2500000, 2500000, 300000)])
def test_csc_structures_match_oracle(ops, M, N, nnz):
    rng = np.random.default_rng(N + nnz)
    row, col = _coalesced(rng, M, N, nnz)
    val = rng.random(row.size).astype(np.float32)
    A = np_ref.SparseTensor(row=row, col=col, value=val, sparse_sizes=(M, N), is_sorted=True, trust_data=True)
    st = A.storage
    perm, colptr, colcount = ops.csr2csc(dev(col), N, True, True)
    assert np.array_equal(host(perm), st.csr2csc())
    assert np.array_equal(host(colptr), st.colptr()) and np.array_equal(host(colcount), st.colcount())
    assert np.array_equal(host(ops.invert_perm(perm)), st.csc2csr())
    perm_only, none1, none2 = ops.csr2csc(dev(col), N, False, False)
    assert none1 is None and none2 is None and torch.equal(perm_only, perm)
    At = np_ref.t(A)
    for cp in (None, colptr):
        tr, tc, tv = ops.transpose_gather(perm, dev(row), dev(col), dev(val), cp, N)
        assert np.array_equal(host(tr), At.storage.row()) and np.array_equal(host(tc), At.storage.col())
        assert np.array_equal(host(tv), At.storage.value())
    wide = rng.random((row.size, 5)).astype(np.float32)
    _, _, tv = ops.transpose_gather(perm, dev(row), dev(col), dev(wide), None, N)
    assert np.array_equal(host(tv), wide[st.csr2csc()])
    tr, tc, tv = ops.transpose_gather(perm, dev(row), dev(col), None, colptr, N)
    assert tv is None and np.array_equal(host(tr), At.storage.row())


@pytest.mark.parametrize("M,N,nnz", [(1, 1, 1), (5, 300, 400), (1000, 1000, 30000), (200, 70000, 90000),
                                     (2500000, 2500000, 300000), (3000, 3000, 1500000)])
@pytest.mark.parametrize("vdt", ["float32", "int32", "bfloat16", "float64", None])
def test_csc_build_matches_oracle(ops, M, N, nnz, vdt):
    """ps_csc_build: csr2csc, colptr, colcount, csc2csr and the transposed arrays in one call (values carried
    through the sort when 4 bytes wide, gathered otherwise) == storage.py:386-447 + transpose.py:9-33."""
    rng = np.random.default_rng(N + nnz)
    row, col = _coalesced(rng, M, N, nnz)
    val = None if vdt is None else rng.integers(-50, 50, row.size).astype(np.float64)
    A = np_ref.SparseTensor(row=row, col=col, value=val, sparse_sizes=(M, N), is_sorted=True, trust_data=True)
    st = A.storage
    tval = None if vdt is None else dev(val).to(getattr(torch, vdt))
    r = ops.csc_build(dev(row), dev(col), tval, N, True, True)
    perm = st.csr2csc()
    assert np.array_equal(host(r["csr2csc"]), perm)
    assert np.array_equal(host(r["colptr"]), st.colptr())
    assert np.array_equal(host(r["colcount"]), st.colcount())
    assert np.array_equal(host(r["csc2csr"]), st.csc2csr())
    assert np.array_equal(host(r["t_row"]), col[perm])
    assert np.array_equal(host(r["t_col"]), row[perm])
    if vdt is None:
        assert r["t_value"] is None
    else:
        assert np.array_equal(host(r["t_value"].to(torch.float64)), val[perm])
    light = ops.csc_build(dev(row), dev(col), tval, N, False, False)
    assert light["csc2csr"] is None and light["colcount"] is None
    assert np.array_equal(host(light["csr2csc"]), perm) and np.array_equal(host(light["colptr"]), st.colptr())


def test_invert_perm_large(ops):
    """the bucketed inverse (radix partition on the high bits + windowed scatter) on a permutation beyond 2^20"""
    n = (1 << 21) + 12345
    perm = torch.randperm(n, device="cuda", generator=torch.Generator(device="cuda").manual_seed(5))
    inv = ops.invert_perm(perm)
    assert torch.equal(inv[perm], torch.arange(n, device="cuda"))


def test_csr2csc_on_uncoalesced_is_stable(ops):
    rng = np.random.default_rng(0)
    M = N = 50
    row, col = _random_coo(rng, M, N, 4000, 0.5)
    order = np.argsort(row * N + col, kind="stable")
    row, col = row[order], col[order]
    perm, _, _ = ops.csr2csc(dev(col), N, False, False)
    assert np.array_equal(host(perm), np.argsort(M * col + row, kind="stable"))


def test_csc_empty(ops):
    e = dev(np.array([], I64))
    perm, colptr, colcount = ops.csr2csc(e, 4, True, True)
    assert perm.numel() == 0 and host(colptr).tolist() == [0] * 5 and host(colcount).tolist() == [0] * 4
    assert ops.invert_perm(e).numel() == 0


def test_gather_rows(ops):
    rng = np.random.default_rng(2)
    idx = rng.integers(0, 1000, 5000).astype(I64)
    for shape, dt in [((1000,), np.float32), ((1000, 3), np.float16), ((1000, 4), np.float64), ((1000, 7), np.int8)]:
        src = (rng.random(shape) * 100).astype(dt)
        if dt == np.int8:
            continue  # uint8/int8 values are not a SparseTensor dtype
        assert np.array_equal(host(ops.gather_rows(dev(src), dev(idx))), src[idx])


# ----------------------------------------------------------------------------- paddle_scatter replacements
def _skewed_ptr(rng, S, E):
    cnt = rng.multinomial(E, rng.dirichlet(np.full(S, 0.3)))
    cnt[rng.integers(0, S, S // 5)] = 0  # empty segments
    ptr = np.concatenate([[0], np.cumsum(cnt)]).astype(I64)
    return ptr, int(ptr[-1])


@pytest.mark.parametrize("dtype", DT)
@pytest.mark.parametrize("reduce", ["sum", "mean", "min", "max"])
@pytest.mark.parametrize("S,E,width", [(1, 10, 1), (300, 20000, 1), (5000, 6000, 1), (64, 100000, 1), (200, 5000, 6)])
def test_segment_csr(ops, dtype, reduce, S, E, width):
    rng = np.random.default_rng(S + E)
    ptr, E = _skewed_ptr(rng, S, E)
    shape = (E,) if width == 1 else (E, width)
    if dtype in (torch.int32, torch.int64):
        src = rng.integers(-99, 99, shape)
    else:
        src = rng.random(shape).astype(np.float32)
    t = dev(src, dtype)
    seen = host(t)
    half = dtype in (torch.float16, torch.bfloat16)
    want = np_ref.segment_csr(seen.astype(np.float64) if half else seen, ptr, reduce)
    got = ops.segment_csr(t, dev(ptr), None, reduce)
    exact = reduce in ("min", "max") or dtype in (torch.int32, torch.int64)
    if dtype == torch.float32 and not exact:
        want = np_ref.segment_csr(seen.astype(np.float64), ptr, reduce)  # order-free reference for long rows
        np.testing.assert_allclose(host(got), want, rtol=2e-6)
    else:
        assert_values(host(got), want, dtype, exact)


@pytest.mark.parametrize("width", [1, 3])
def test_gather_csr_and_mul(ops, width):
    rng = np.random.default_rng(9)
    M, N = 700, 900
    row, col = _coalesced(rng, M, N, 40000)
    nnz = row.size
    rowptr = np_ref.SparseStorage(row=row, col=col, sparse_sizes=(M, N), is_sorted=True, trust_data=True).rowptr()
    for dtype in (torch.float32, torch.bfloat16, torch.float64, torch.int64):
        shape = (nnz,) if width == 1 else (nnz, width)
        val = dev((rng.random(shape) * 8).astype(np.float32), dtype)
        orow = dev((rng.random((M,) if width == 1 else (M, width)) * 8).astype(np.float32), dtype)
        ocol = dev((rng.random((N,) if width == 1 else (N, width)) * 8).astype(np.float32), dtype)
        assert np.array_equal(host(ops.gather_csr(orow, dev(rowptr), nnz)), np_ref.gather_csr(host(orow), rowptr))
        assert np.array_equal(host(ops.gather_csr(orow, dev(rowptr))), np_ref.gather_csr(host(orow), rowptr))
        want_r = (val.cpu() * orow.cpu()[torch.from_numpy(row)])
        want_c = (val.cpu() * ocol.cpu()[torch.from_numpy(col)])
        assert torch.equal(ops.mul_row(val, dev(rowptr), orow).cpu(), want_r)
        assert torch.equal(ops.mul_col(val, dev(col), ocol).cpu(), want_c)
        assert torch.equal(ops.mul_nnz(val, val).cpu(), val.cpu() * val.cpu())
        v2 = val.clone()
        ops.mul_row(v2, dev(rowptr), orow, out=v2)  # in place (mul_)
        assert torch.equal(v2.cpu(), want_r)


@pytest.mark.parametrize("dtype", DT)
@pytest.mark.parametrize("reduce", ["sum", "mean", "min", "max"])
def test_scatter_by_col(ops, dtype, reduce):
    rng = np.random.default_rng(21)
    N, nnz = 3000, 50000
    col = rng.integers(0, N - 100, nnz).astype(I64)  # the last 100 columns stay untouched -> 0
    for width in (1, 4):
        shape = (nnz,) if width == 1 else (nnz, width)
        if dtype in (torch.int32, torch.int64):
            src = rng.integers(-99, 99, shape)
        else:
            src = rng.standard_normal(shape).astype(np.float32)
        t = dev(src, dtype)
        seen = host(t)
        half = dtype in (torch.float16, torch.bfloat16)
        want = np_ref.scatter(seen.astype(np.float64) if half or dtype == torch.float32 else seen, col, N, reduce)
        got = host(ops.scatter(t, dev(col), 0, None, N, reduce))
        if reduce in ("min", "max") or dtype in (torch.int32, torch.int64):
            assert np.array_equal(got.astype(np.float64), want.astype(np.float64))
        else:  # signed values cancel: compare against the magnitude of the terms
            tol = {torch.float32: 2e-6, torch.float64: 1e-12, torch.float16: 2e-3, torch.bfloat16: 1e-2}[dtype]
            mag = np_ref.scatter(np.abs(seen).astype(np.float64), col, N, "sum" if reduce == "sum" else reduce)
            assert np.all(np.abs(got - want) <= tol * np.maximum(np.abs(mag), 1e-30) + 0)


@pytest.mark.parametrize("reduce", ["sum", "mean", "min", "max"])
def test_reduce_all(ops, reduce):
    rng = np.random.default_rng(4)
    x = rng.random(300001).astype(np.float32)
    want = {"sum": x.astype(np.float64).sum(), "mean": x.astype(np.float64).mean(), "min": x.min(), "max": x.max()}[reduce]
    np.testing.assert_allclose(float(ops.reduce_all(dev(x), reduce)), want, rtol=1e-6)
    xi = rng.integers(-9, 9, 1001).astype(I64)
    assert int(ops.reduce_all(dev(xi), "sum")) == xi.sum() and int(ops.reduce_all(dev(xi), "max")) == xi.max()


# ----------------------------------------------------------------------------- select building blocks
def test_mask_compaction(ops):
    rng = np.random.default_rng(6)
    mask = rng.random(100003) < 0.3
    assert np.array_equal(host(ops.mask_to_index(dev(mask))), np.nonzero(mask)[0])
    assert ops.mask_to_index(dev(np.zeros(10, bool))).numel() == 0
    assert ops.mask_to_index(dev(np.zeros(0, bool))).numel() == 0
    index = rng.integers(0, mask.size, 250000).astype(I64)
    assert np.array_equal(host(ops.mask_entries(dev(mask), dev(index), index.size)), np.nonzero(mask[index])[0])
    assert np.array_equal(host(ops.mask_entries(dev(mask), None, mask.size)), np.nonzero(mask)[0])


def test_index_select_fill(ops):
    rng = np.random.default_rng(8)
    M, N = 500, 400
    row, col = _coalesced(rng, M, N, 30000)
    val = rng.random(row.size).astype(np.float32)
    A = np_ref.SparseTensor(row=row, col=col, value=val, sparse_sizes=(M, N), is_sorted=True, trust_data=True)
    idx = rng.integers(0, M, 1200).astype(I64)  # duplicates allowed
    want = np_ref.index_select(A, 0, idx)
    rowptr = dev(A.storage.rowptr())
    rowcount = ops.ptr2count(rowptr)
    out_ptr, E = ops.count2ptr(rowcount, dev(idx), want_total=True)
    major, minor, v, perm = ops.index_select_fill(rowptr, dev(col), dev(val), dev(idx), out_ptr, E, want_perm=True)
    assert np.array_equal(host(out_ptr), want.storage.rowptr()) and E == want.nnz()
    assert np.array_equal(host(major), want.storage.row()) and np.array_equal(host(minor), want.storage.col())
    assert np.array_equal(host(v), want.storage.value())
    assert np.array_equal(val[host(perm)], want.storage.value())


# ----------------------------------------------------------------------------- row-range partitioning
@pytest.mark.parametrize("G_", [1, 2, 8])
def test_dist_histogram_and_partition(ops, G_):
    rng = np.random.default_rng(13)
    M, nnz = 100000, 300001
    row = rng.integers(0, M, nnz).astype(I64)
    col = rng.integers(0, M, nnz).astype(I64)
    val = rng.random(nnz).astype(np.float32)
    for shift, buckets in [(4, M // 16 + 1), (0, M)]:
        hist = host(ops.dist_row_histogram(dev(row), shift, buckets))
        assert np.array_equal(hist, np.bincount(row >> shift, minlength=buckets))
    splitters = np.concatenate([[0], np.sort(rng.integers(0, M, G_ - 1)), [M]]).astype(I64)
    r, c, v, counts = ops.dist_partition(dev(row), dev(col), dev(val), dev(splitters))
    dest = np.searchsorted(splitters[1:], row, side="right")
    order = np.argsort(dest, kind="stable")
    assert np.array_equal(host(counts), np.bincount(dest, minlength=G_))
    assert np.array_equal(host(r), row[order]) and np.array_equal(host(c), col[order])
    assert np.array_equal(host(v), val[order])
    # value-less tuples ride the same single pass; wide values take the permutation + gather route
    r, c, v, _ = ops.dist_partition(dev(row), dev(col), None, dev(splitters))
    assert v is None and np.array_equal(host(r), row[order]) and np.array_equal(host(c), col[order])
    wide = rng.random((nnz, 3))
    r, c, v, _ = ops.dist_partition(dev(row), dev(col), dev(wide), dev(splitters))
    assert np.array_equal(host(r), row[order]) and np.array_equal(host(v), wide[order])
    big = row + (1 << 31)  # indices above 2^31 survive the 32-bit payload words
    r, c, _, _ = ops.dist_partition(dev(big), dev(col), None, dev(splitters + (1 << 31)))
    assert np.array_equal(host(r), big[order])
