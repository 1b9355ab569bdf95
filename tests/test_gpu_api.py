"""GPU parity tests of the public API (SparseStorage / SparseTensor / free functions) against the
NumPy restatement of the reference's Python layer (oracle/np_ref.py) and the reference's golden
vectors.  These read like the reference's own test/test_storage.py, test_coalesce.py,
test_transpose.py, test_permute.py and test_sample.py."""
import numpy as np
import pytest
import torch

from oracle import np_ref
from tests.golden import reference_vectors as G
from tests._util import I64, _random_coo, dev, host

pytestmark = pytest.mark.gpu
DTYPES = [torch.float16, torch.float32, torch.float64, torch.int32, torch.int64, torch.bfloat16]  # testing.py:12-17


@pytest.fixture(scope="module")
def ps(ops):
    import paddle_sparse_b200
    return paddle_sparse_b200


def ten(x, dtype=torch.int64):
    return torch.tensor(x, dtype=dtype, device="cuda")


def _pair(rng, M, N, nnz, dtype=np.float32, width=None, coalesced=True):
    """The same matrix as an oracle SparseTensor and (constructor kwargs for) a GPU one."""
    row, col = _random_coo(rng, M, N, nnz, 0.15)
    if coalesced:
        key = np.unique(row * N + col)
        row, col = (key // N).astype(I64), (key % N).astype(I64)
    shape = (row.size,) if width is None else (row.size, width)
    val = rng.random(shape).astype(dtype)
    return row, col, val


def _same_storage(got, want, values_exact=True, check_ptr=True):
    assert got.sparse_sizes() == tuple(want.sparse_sizes())
    assert np.array_equal(host(got.storage.row()), want.storage.row())
    assert np.array_equal(host(got.storage.col()), want.storage.col())
    if check_ptr:
        assert np.array_equal(host(got.storage.rowptr()), want.storage.rowptr())
    if want.storage.value() is None:
        assert got.storage.value() is None
    elif values_exact:
        assert np.array_equal(host(got.storage.value()), want.storage.value())
    else:
        np.testing.assert_allclose(host(got.storage.value()), want.storage.value(), rtol=1e-6)
    for key in want.storage.cached_keys():  # every cache the reference carries over must agree
        mine = getattr(got.storage, "_" + key)
        if mine is not None:
            assert np.array_equal(host(mine), getattr(want.storage, "_" + key)), key


# ----------------------------------------------------------------------------- test_storage.py
@pytest.mark.parametrize("dtype", DTYPES)
def test_storage_sort_on_construct(ps, dtype):
    g = G.SORT_ON_CONSTRUCT
    st = ps.SparseStorage(row=ten(g["row"]), col=ten(g["col"]), value=ten(g["value"], dtype))
    assert st.row().tolist() == g["out_row"] and st.col().tolist() == g["out_col"]
    assert st.value().float().tolist() == g["out_value"]
    assert st.sparse_sizes() == g["sizes"]


def test_storage_caching(ps):
    g = G.CACHING
    st = ps.SparseStorage(row=ten(g["row"]), col=ten(g["col"]))
    assert st._rowcount is None and st._rowptr is None and st._colcount is None and st._colptr is None
    assert st._csr2csc is None and st.num_cached_keys() == 0
    st.fill_cache_()
    for k in ("rowcount", "rowptr", "colcount", "colptr", "csr2csc", "csc2csr"):
        assert getattr(st, "_" + k).tolist() == g[k], k
    assert st.num_cached_keys() == 5
    st2 = ps.SparseStorage(row=st._row, rowptr=st._rowptr, col=st._col, value=st._value,
                           sparse_sizes=st._sparse_sizes, rowcount=st._rowcount, colptr=st._colptr,
                           colcount=st._colcount, csr2csc=st._csr2csc, csc2csr=st._csc2csr)
    assert st2.num_cached_keys() == 5 and st2._csc2csr.tolist() == g["csc2csr"]
    st2.clear_cache_()
    assert st2._rowcount is None and st2._rowptr is not None and st2._colptr is None and st2.num_cached_keys() == 0
    # lazy one at a time: cached_keys evolves as in the reference
    st3 = ps.SparseStorage(row=ten(g["row"]), col=ten(g["col"]))
    st3.csr2csc()
    assert st3.cached_keys() == ["csr2csc"]
    assert st3.colptr().tolist() == g["colptr"] and st3.cached_keys() == ["colptr", "csr2csc"]
    st4 = ps.SparseStorage(row=ten(g["row"]), col=ten(g["col"]))
    assert st4.colptr().tolist() == g["colptr"] and st4.cached_keys() == ["colptr", "colcount"]


@pytest.mark.parametrize("dtype", DTYPES)
def test_storage_set_value_and_copy(ps, dtype):
    g = G.SET_VALUE
    st = ps.SparseStorage(row=ten(g["row"]), col=ten(g["col"]), value=ten(g["value"], dtype))
    st.set_value_(ten(g["value"], dtype), layout="csc")
    assert st.value().float().tolist() == [float(v) for v in g["csc_layout"]]
    st.set_value_(ten(g["value"], dtype), layout="coo")
    assert st.value().float().tolist() == [float(v) for v in g["coo_layout"]]
    out = st.set_value(ten(g["value"], dtype), layout="csc")
    assert st.value().float().tolist() == [float(v) for v in g["coo_layout"]]
    assert out.value().float().tolist() == [float(v) for v in g["csc_layout"]]
    assert out.col().data_ptr() == st.col().data_ptr()
    with pytest.warns(UserWarning):
        st.set_value_(ten(g["value"], dtype))
    # test/test_storage.py:145-154
    st = st.sparse_resize((3, 3))
    assert st.sparse_sizes() == (3, 3)
    new = st.copy()
    assert new != st and new is not st
    assert new.col().data_ptr() == st.col().data_ptr()
    new = st.clone()
    assert new != st
    assert new.col().data_ptr() != st.col().data_ptr()
    assert new.col().tolist() == st.col().tolist() and new.value().float().tolist() == st.value().float().tolist()


def test_sparse_resize_and_reshape_match_the_oracle(ps):
    """storage.py:281-368: golden vectors of test/test_storage.py:183-200, then random matrices with every cache
    built against the oracle's restatement (which caches survive, their contents, the shared row / col / value)."""
    g = G.SPARSE_RESHAPE
    st = ps.SparseStorage(row=ten(g["row"]), col=ten(g["col"]))
    for args, sizes, row, col in g["steps"]:
        st = st.sparse_reshape(*args)
        assert st.sparse_sizes() == sizes and st.row().tolist() == row and st.col().tolist() == col
    rng = np.random.default_rng(12)
    M, N = 700, 450
    key = np.unique(rng.integers(0, M * N, 20000))
    row, col = (key // N).astype(I64), (key % N).astype(I64)
    val = rng.random(key.size).astype(np.float32)
    O = np_ref.SparseStorage(row=row, col=col, value=val, sparse_sizes=(M, N), is_sorted=True)
    O.rowptr(); O.rowcount(); O.colptr(); O.colcount(); O.csr2csc(); O.csc2csr()
    A = ps.SparseStorage(row=dev(row), col=dev(col), value=dev(val), sparse_sizes=(M, N), is_sorted=True).fill_cache_()
    for sizes in ((M + 123, N + 77), (M, N + 5), (M + 1, N), (M, N)):
        got, want = A.sparse_resize(sizes), O.sparse_resize(sizes)
        assert got.sparse_sizes() == tuple(sizes) and got.cached_keys() == want.cached_keys()
        for k in ("rowptr", "rowcount", "colptr", "colcount", "csr2csc", "csc2csr"):
            assert np.array_equal(host(getattr(got, "_" + k)), getattr(want, "_" + k)), (sizes, k)
        assert got.col().data_ptr() == A.col().data_ptr() and got.value().data_ptr() == A.value().data_ptr()
        back = got.sparse_resize((M, N))  # cutting back restores the original pointers and counts
        for k in ("rowptr", "rowcount", "colptr", "colcount"):
            assert np.array_equal(host(getattr(back, "_" + k)), getattr(O, "_" + k)), (sizes, k)
    for shape in ((M * 2, N // 2), (N, M), (-1, 90), (350, -1), (1, M * N)):
        got, want = A.sparse_reshape(*shape), O.sparse_reshape(*shape)
        assert got.sparse_sizes() == want.sparse_sizes() and got.num_cached_keys() == 0
        assert np.array_equal(host(got.row()), want.row()) and np.array_equal(host(got.col()), want.col())
        assert np.array_equal(host(got.rowptr()), want.rowptr()) and got.value().data_ptr() == A.value().data_ptr()
    with pytest.raises(AssertionError):
        A.sparse_reshape(7, 11)


def test_storage_dtype_and_placement(ps):
    """storage.py:582-776: type / type_as / to_device / device_as / cuda / pin_memory / is_pinned / share_memory_ /
    is_shared.  Caches travel with the data; a host copy is an inert staging holder."""
    g = G.CACHING
    st = ps.SparseStorage(row=ten(g["row"]), col=ten(g["col"]), value=ten([1, 2, 3, 4], torch.float32)).fill_cache_()
    assert st.type(torch.float32) is st and st.type_as(torch.zeros(1)) is st
    h = st.type(torch.float16)
    assert h is not st and h.value().dtype == torch.float16 and h.col().data_ptr() == st.col().data_ptr()
    assert h.cached_keys() == st.cached_keys() and st.type_as(torch.zeros(1, dtype=torch.float64)).value().dtype == torch.float64
    novalue = ps.SparseStorage(row=ten(g["row"]), col=ten(g["col"]))
    assert novalue.type(torch.float16) is novalue
    assert st.cuda() is st and st.to_device("cuda") is st and st.to_device(st.col().device) is st
    assert st.device_as(st.col()) is st and not st.is_pinned()
    pinned = st.pin_memory()
    assert pinned.is_pinned() and not pinned.col().is_cuda and pinned.cached_keys() == st.cached_keys()
    assert pinned.col().tolist() == st.col().tolist() and pinned._csr2csc.tolist() == g["csr2csc"]
    with pytest.raises(RuntimeError, match="GPU-only"):  # a host copy holds data only: no operator runs on it
        pinned.clear_cache_().csr2csc()
    back = st.pin_memory().to_device("cuda", non_blocking=True)
    assert back.col().is_cuda and back.csc2csr().tolist() == g["csc2csr"] and back.num_cached_keys() == 5
    assert back.value().tolist() == [1, 2, 3, 4] and pinned.cuda().col().is_cuda
    assert back.device_as(torch.zeros(1)).col().device.type == "cpu"
    with pytest.warns(UserWarning):
        assert st.share_memory_() is None
    with pytest.warns(UserWarning):
        assert st.is_shared() is False


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two CUDA devices in one process")
def test_second_device_in_one_process(ps):
    """cudaFuncSetAttribute is per device: the > 48 KB shared-memory kernels must be configured on each device a
    process uses."""
    rng = np.random.default_rng(2)
    M, N, nnz = 3_000_000, 2_500_000, 300_000
    row, col = _random_coo(rng, M, N, nnz, 0.3)
    val = rng.random(nnz).astype(np.float32)
    want = np_ref.build_pipeline(row, col, val, M, N)
    A0 = ps.SparseTensor(row=dev(row), col=dev(col), value=dev(val), sparse_sizes=(M, N)).coalesce("add").fill_cache_()
    raw = ps.SparseTensor(row=dev(row).to("cuda:1"), col=dev(col).to("cuda:1"), value=dev(val).to("cuda:1"),
                          sparse_sizes=(M, N))
    A1 = raw.coalesce("add").fill_cache_()
    for A in (A0, A1, A0.to("cuda:1").clear_cache_().fill_cache_()):
        st = A.storage
        for k, t in dict(row=st.row(), col=st.col(), rowptr=st.rowptr(), colptr=st.colptr(), csr2csc=st.csr2csc(),
                         csc2csr=st.csc2csr()).items():
            assert np.array_equal(t.cpu().numpy(), want[k]), k
    assert A1.device() == torch.device("cuda:1")


@pytest.mark.parametrize("dtype", DTYPES)
def test_storage_coalesce(ps, dtype):
    g = G.STORAGE_COALESCE
    st = ps.SparseStorage(row=ten(g["row"]), col=ten(g["col"]), value=ten(g["value"], dtype))
    assert st.row().tolist() == g["row"] and st.col().tolist() == g["col"]
    assert not st.is_coalesced()
    st2 = st.coalesce()
    assert st2.is_coalesced() and st2.coalesce() is st2  # storage.py:459-460 returns self
    assert st2.row().tolist() == g["out_row"] and st2.col().tolist() == g["out_col"]
    assert st2.value().float().tolist() == [float(v) for v in g["out_value"]]
    assert st2.num_cached_keys() == 0


def test_storage_validation(ps):
    with pytest.raises(AssertionError):
        ps.SparseStorage(row=ten([0, 5]), col=ten([0, 1]), sparse_sizes=(3, 3))
    with pytest.raises(AssertionError):
        ps.SparseStorage(row=ten([0, 1]), col=ten([0, 9]), sparse_sizes=(3, 3))
    with pytest.raises(RuntimeError):
        ps.SparseStorage(row=torch.tensor([0]), col=torch.tensor([0]))  # CPU tensors: no fallback
    st = ps.SparseStorage(rowptr=ten([0, 1, 1, 3]), col=ten([2, 0, 1]))
    assert st.sparse_sizes() == (3, 3) and st.row().tolist() == [0, 2, 2]
    e = ps.SparseStorage.empty()
    assert e.sparse_sizes() == (0, 0) and e.rowptr().tolist() == [0]


def test_construction_inspect_validates_and_feeds_coalesce(ps, ops):
    """Large unsorted inputs: the constructor's validation pass is the bucket histogram of the coalesce that
    follows (ps_coo_inspect).  Out-of-range indices are still rejected at construction; the coalesce that reuses
    the inspection equals the one that does not; an in-place edit of row / col in between is noticed."""
    rng = np.random.default_rng(59)
    M, N, nnz = 3_000_000, 2_500_000, 400_000
    row, col = _random_coo(rng, M, N, nnz, 0.3)
    val = rng.random(nnz).astype(np.float32)
    for bad_row, bad_col in ((M, 0), (0, N), (M + (1 << 40), 0), (0, N + (1 << 45))):
        r, c = row.copy(), col.copy()
        r[1234], c[4321] = max(r[1234], bad_row), max(c[4321], bad_col)
        with pytest.raises(AssertionError):
            ps.SparseStorage(row=dev(r), col=dev(c), value=dev(val), sparse_sizes=(M, N))
    mx = ops.coo_inspect(dev(row), dev(col), M, N)
    assert mx[0] == row.max() and mx[1] == col.max() and mx[2] is not None
    assert ops.coo_inspect(dev(row[:1000]), dev(col[:1000]), M, N)[2] is None  # too small for the hybrid plan
    want = np_ref.SparseStorage(row=row, col=col, value=val.astype(np.float64), sparse_sizes=(M, N)).coalesce("add")
    st = ps.SparseStorage(row=dev(row), col=dev(col), value=dev(val), sparse_sizes=(M, N))
    assert st._inspection is not None
    fused = st.coalesce("add")
    plain = ops.coalesce(dev(row), dev(col), dev(val), M, N, "add", True)
    for got, ref_, w in zip((fused.row(), fused.col(), fused.rowptr()), (plain[0], plain[1], plain[3]),
                            (want.row(), want.col(), want.rowptr())):
        assert torch.equal(got, ref_) and np.array_equal(host(got), w)
    assert torch.equal(fused.value(), plain[2])
    np.testing.assert_allclose(host(fused.value()), want.value(), rtol=1e-6)
    # in-place edits between construction and first use
    d_row, d_col = dev(row), dev(col)
    st = ps.SparseStorage(row=d_row, col=d_col, value=dev(val), sparse_sizes=(M, N))
    d_col.add_(1).sub_(1)                      # same content, new version: the inspection is dropped, result unchanged
    assert torch.equal(st.coalesce("add").col(), fused.col())
    st = ps.SparseStorage(row=d_row, col=d_col, value=dev(val), sparse_sizes=(M, N))
    d_row[7] = M + 5
    with pytest.raises(AssertionError):
        st.coalesce("add")
    st = ps.SparseStorage(row=d_row, col=d_col, value=dev(val), sparse_sizes=(M, N), trust_data=True)
    assert st._inspection is None              # trust_data: nothing is read at construction


# ----------------------------------------------------------------------------- test_coalesce.py / test_transpose.py
def test_free_coalesce_golden(ps):
    g = G.COALESCE
    index = ten([g["row"], g["col"]])
    idx, val = ps.coalesce(index, None, m=g["m"], n=g["n"])
    assert idx.tolist() == g["index"] and val is None
    idx, val = ps.coalesce(index, ten(g["value"], torch.float32), m=g["m"], n=g["n"])
    assert idx.tolist() == g["index"] and val.tolist() == g["add"]
    idx, val = ps.coalesce(index, ten(g["value"], torch.float32), m=g["m"], n=g["n"], op="max")
    assert idx.tolist() == g["index"] and val.tolist() == g["max"]


@pytest.mark.parametrize("dtype", DTYPES)
def test_transpose_golden(ps, dtype):
    g = G.TRANSPOSE_MATRIX
    mat = ps.SparseTensor(row=ten(g["row"]), col=ten(g["col"]), value=ten(g["value"], dtype),
                          sparse_sizes=(g["m"], g["n"]))
    mt = mat.t()
    assert mt.sparse_sizes() == (g["n"], g["m"])
    r, c, v = mt.coo()
    assert [r.tolist(), c.tolist()] == g["index"] and v.float().tolist() == [float(x) for x in g["out_value"]]
    g = G.TRANSPOSE
    idx, val = ps.transpose(ten([g["row"], g["col"]]), ten(g["value"], dtype), m=g["m"], n=g["n"])
    assert idx.tolist() == g["index"] and val.float().tolist() == [[float(x) for x in r] for r in g["out_value"]]


def test_t_swaps_caches(ps):
    rng = np.random.default_rng(0)
    row, col, val = _pair(rng, 300, 200, 9000)
    A = ps.SparseTensor(row=dev(row), col=dev(col), value=dev(val), sparse_sizes=(300, 200)).fill_cache_()
    O = np_ref.SparseTensor(row=row, col=col, value=val, sparse_sizes=(300, 200))
    O.fill_cache_()
    At, Ot = A.t(), np_ref.t(O)
    _same_storage(At, Ot)
    assert sorted(At.storage.cached_keys()) == sorted(Ot.storage.cached_keys())
    _same_storage(At.t(), O)
    B = ps.SparseTensor(row=dev(row), col=dev(col), sparse_sizes=(300, 200))  # value-less, cold caches
    _same_storage(B.t(), np_ref.t(np_ref.SparseTensor(row=row, col=col, sparse_sizes=(300, 200))))


# ----------------------------------------------------------------------------- reduce.py / mul.py
@pytest.mark.parametrize("reduce", ["sum", "mean", "min", "max"])
@pytest.mark.parametrize("width", [None, 3])
def test_reductions_all_dims(ps, reduce, width):
    rng = np.random.default_rng(17)
    M, N = 400, 250
    row, col, val = _pair(rng, M, N, 12000, width=width)
    A = ps.SparseTensor(row=dev(row), col=dev(col), value=dev(val), sparse_sizes=(M, N))
    O = np_ref.SparseTensor(row=row, col=col, value=val.astype(np.float64), sparse_sizes=(M, N))
    for dim in ([None, 0, 1, -1] if width is None else [None, 0, 1, 2, -1]):
        got = getattr(A, reduce)(dim)
        want = np_ref.reduction(O, dim, reduce)
        np.testing.assert_allclose(host(got), want, rtol=0 if reduce in ("min", "max") else 2e-6)
    # the test_reduce.py:13-62 identity
    np.testing.assert_allclose(float(getattr(A, reduce)()), float(getattr(val.astype(np.float64), reduce)()), rtol=1e-6)
    # with the CSC caches filled, dim=0 runs over the parked value[csr2csc] (segment reduction over colptr), and
    # after t() took that away, with the per-column counts from the cached colcount: same answer each time
    A.fill_cache_()
    assert A.storage.has_colcount()
    for parked in (True, False):
        assert (A.storage.csc_ordered_value() is not None) == parked
        np.testing.assert_allclose(host(getattr(A, reduce)(0)), np_ref.reduction(O, 0, reduce),
                                   rtol=0 if reduce in ("min", "max") else 2e-6)
        A.t()
    V = ps.SparseTensor(row=dev(row), col=dev(col), sparse_sizes=(M, N))  # value-less
    OV = np_ref.SparseTensor(row=row, col=col, sparse_sizes=(M, N))
    for dim in (None, 0, 1):
        got, want = getattr(V, reduce)(dim), np_ref.reduction(OV, dim, reduce)
        assert got.dtype == torch.float32 and np.array_equal(host(got), want)
    with pytest.raises(ValueError):
        ps.reduction(A, 0, "prod")


def test_repeated_column_reductions_switch_to_contiguous_columns(ps):
    """reduce.py:40-42 asked repeatedly on one matrix: the first call is an atomic scatter, the second builds the
    carrying CSC sort privately and every later one is a segment reduction over colptr - same results, no cache key
    appears (cached_keys evolves as in the reference), and a new value tensor or clear_cache_ drops the helper."""
    rng = np.random.default_rng(23)
    M, N = 3000, 2000
    row, col, val = _pair(rng, M, N, 150_000)
    A = ps.SparseTensor(row=dev(row), col=dev(col), value=dev(val), sparse_sizes=(M, N))
    O = np_ref.SparseTensor(row=row, col=col, value=val.astype(np.float64), sparse_sizes=(M, N))
    for k, red in enumerate(("sum", "mean", "max", "min", "sum")):
        got = getattr(A, red)(0)
        np.testing.assert_allclose(host(got), np_ref.reduction(O, 0, red), rtol=0 if red in ("min", "max") else 2e-6)
        assert (getattr(A.storage, "_dim0_accel", None) is not None) == (k >= 1)
        assert A.storage.cached_keys() == []
    B = A.set_value(A.storage.value() * 2, layout="coo")
    assert getattr(B.storage, "_dim0_accel", None) is None
    np.testing.assert_allclose(host(B.sum(0)), 2 * np_ref.reduction(O, 0, "sum"), rtol=2e-6)
    A.storage.value().mul_(3.0)  # an in-place edit is noticed through the version counter
    np.testing.assert_allclose(host(A.sum(0)), 3 * np_ref.reduction(O, 0, "sum"), rtol=2e-6)
    A.clear_cache_()
    assert getattr(A.storage, "_dim0_accel", None) is None and A.storage._dim0_calls == 0


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float64])
def test_mul_broadcast(ps, dtype):
    rng = np.random.default_rng(23)
    M, N = 300, 500
    row, col, val = _pair(rng, M, N, 20000)
    tv = dev(val, dtype)
    seen = tv.float().cpu().numpy().astype(np.float64) if dtype != torch.float64 else host(tv)
    orow = dev(rng.random((M, 1)).astype(np.float32), dtype)
    ocol = dev(rng.random((1, N)).astype(np.float32), dtype)
    for other in (orow, ocol):
        A = ps.SparseTensor(row=dev(row), col=dev(col), value=tv.clone(), sparse_sizes=(M, N))
        O = np_ref.SparseTensor(row=row, col=col, value=seen.copy(), sparse_sizes=(M, N))
        want = np_ref.mul(O, other.float().cpu().numpy().astype(np.float64)).storage.value()
        rtol = {torch.float32: 1e-6, torch.bfloat16: 8e-3, torch.float64: 1e-12}[dtype]
        out = A.mul(other)
        np.testing.assert_allclose(host(out.storage.value()), want, rtol=rtol)
        assert torch.equal(A.storage.value(), tv)  # out of place
        assert torch.equal((A * other).storage.value(), out.storage.value())
        B = A.mul_(other)
        assert B is A and torch.equal(A.storage.value(), out.storage.value())
    V = ps.SparseTensor(row=dev(row), col=dev(col), sparse_sizes=(M, N))  # test_overload.py: value-less
    assert np.array_equal(host((V * orow).storage.value()), host(orow)[:, 0][row])
    assert np.array_equal(host((V * ocol).storage.value()), host(ocol)[0][col])
    with pytest.raises(ValueError):
        V.mul(dev(np.ones((3, 3), np.float32)))
    A = ps.SparseTensor(row=dev(row), col=dev(col), value=tv.clone(), sparse_sizes=(M, N))
    w = dev(rng.random(row.size).astype(np.float32), dtype)
    assert torch.equal(A.mul_nnz(w, layout="coo").storage.value(), tv * w)


def test_value_edits_after_fill_cache_reach_t_and_column_sums(ps):
    """fill_cache_ parks value[csr2csc] for t(); an in-place multiply (same tensor object), an in-place torch edit of
    storage.value() and set_value_ must each invalidate it."""
    rng = np.random.default_rng(53)
    M, N = 200, 300
    row, col, val = _pair(rng, M, N, 8000)
    other = rng.random((M, 1)).astype(np.float32)

    def fresh():
        A = ps.SparseTensor(row=dev(row), col=dev(col), value=dev(val), sparse_sizes=(M, N)).fill_cache_()
        assert A.storage.csc_ordered_value() is not None
        return A
    edits = {
        "mul_": (lambda A: A.mul_(dev(other)), val * other[row, 0]),
        "torch in-place": (lambda A: A.storage.value().mul_(3.0), val * np.float32(3.0)),
        "set_value_": (lambda A: A.set_value_(dev(val + 1), layout="coo"), val + 1),
    }
    for name, (edit, new_val) in edits.items():
        O = np_ref.SparseTensor(row=row, col=col, value=new_val.astype(np.float32), sparse_sizes=(M, N))
        A = fresh()
        edit(A)
        np.testing.assert_allclose(host(A.sum(0)), np_ref.reduction(O, 0, "sum"), rtol=2e-6, err_msg=name)
        A = fresh()
        edit(A)
        _, _, tv = A.t().coo()
        assert np.array_equal(host(tv), np_ref.t(O).coo()[2]), name


# ----------------------------------------------------------------------------- add / sparse mul / to_symmetric
@pytest.mark.parametrize("dtype", [torch.float32, torch.float64, torch.int64, torch.bfloat16, torch.float16])
def test_sparse_add_mul_golden(ps, dtype):
    """test/test_add.py:15-39, test/test_mul.py:15-61."""
    def T(r, c, v):
        return ps.SparseTensor(row=dev(np.array(r, I64)), col=dev(np.array(c, I64)),
                               value=dev(np.array(v, np.float32), dtype))
    g = G.SPARSE_ADD
    C = T(g["rowA"], g["colA"], g["valueA"]) + T(g["rowB"], g["colB"], g["valueB"])
    row, col, val = C.coo()
    assert row.tolist() == g["row"] and col.tolist() == g["col"] and host(val).tolist() == g["value"]
    assert val.dtype == dtype
    for g in (G.SPARSE_MUL, G.SPARSE_MUL_EMPTY):
        C = T(g["rowA"], g["colA"], g["valueA"]) * T(g["rowB"], g["colB"], g["valueB"])
        row, col, val = C.coo()
        assert row.tolist() == g["row"] and col.tolist() == g["col"] and host(val).tolist() == g["value"]
    dup = T([0, 0], [1, 1], [1, 2])
    with pytest.raises(ValueError):  # mul.py:38-41
        dup * dup


def test_to_symmetric_golden(ps):
    """test/test_tensor.py:71-88."""
    g = G.TO_SYMMETRIC
    A = ps.SparseTensor(row=dev(np.array(g["row"], I64)), col=dev(np.array(g["col"], I64)),
                        value=dev(np.array(g["value"], I64)))
    assert not A.is_symmetric()
    S = A.to_symmetric()
    assert S.is_symmetric()
    assert S.to_dense().tolist() == g["dense"]


@pytest.mark.parametrize("with_value", [True, False])
@pytest.mark.parametrize("sizes", [((400, 300), (350, 420)), ((5000, 5000), (5000, 5000))])
def test_sparse_add_matches_oracle(ps, sizes, with_value):
    rng = np.random.default_rng(31)
    (MA, NA), (MB, NB) = sizes
    rowA, colA = _random_coo(rng, MA, NA, 30000)
    rowB, colB = _random_coo(rng, MB, NB, 40000)
    vA = rng.integers(0, 1 << 20, rowA.size).astype(np.float64) if with_value else None
    vB = rng.integers(0, 1 << 20, rowB.size).astype(np.float64) if with_value else None
    A = ps.SparseTensor(row=dev(rowA), col=dev(colA), value=dev(vA) if with_value else None, sparse_sizes=(MA, NA))
    B = ps.SparseTensor(row=dev(rowB), col=dev(colB), value=dev(vB) if with_value else None, sparse_sizes=(MB, NB))
    OA = np_ref.SparseTensor(row=rowA, col=colA, value=vA, sparse_sizes=(MA, NA))
    OB = np_ref.SparseTensor(row=rowB, col=colB, value=vB, sparse_sizes=(MB, NB))
    C, OC = A + B, np_ref.add(OA, OB)
    assert tuple(C.sparse_sizes()) == tuple(OC.sparse_sizes())
    for got, want in zip(C.coo(), OC.coo()):
        assert (got is None) == (want is None)
        if got is not None:
            assert np.array_equal(host(got), want)  # integer-valued doubles: sums exact in any order
    assert np.array_equal(host(C.storage.rowptr()), OC.storage.rowptr())


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64, torch.bfloat16])
@pytest.mark.parametrize("width", [0, 3])
def test_sparse_mul_matches_oracle(ps, dtype, width):
    rng = np.random.default_rng(37)
    MA, NA, MB, NB = 600, 700, 650, 500
    rowA, colA = _random_coo(rng, MA, NA, 60000)
    rowB, colB = _random_coo(rng, MB, NB, 50000)
    shape = lambda n: (n,) if width == 0 else (n, width)  # noqa: E731
    vA = rng.integers(1, 16, shape(rowA.size)).astype(np.float64)  # products < 256: exact in bf16
    vB = rng.integers(1, 16, shape(rowB.size)).astype(np.float64)
    A = ps.SparseTensor(row=dev(rowA), col=dev(colA), value=dev(vA, dtype), sparse_sizes=(MA, NA)).coalesce("max")
    B = ps.SparseTensor(row=dev(rowB), col=dev(colB), value=dev(vB, dtype), sparse_sizes=(MB, NB)).coalesce("max")
    OA = np_ref.SparseTensor(row=rowA, col=colA, value=vA, sparse_sizes=(MA, NA)).coalesce("max")
    OB = np_ref.SparseTensor(row=rowB, col=colB, value=vB, sparse_sizes=(MB, NB)).coalesce("max")
    C, OC = A * B, np_ref.mul_sparse(OA, OB)
    assert C.nnz() == OC.nnz() > 0
    assert tuple(C.sparse_sizes()) == tuple(OC.sparse_sizes()) == (650, 700)
    for got, want in zip(C.coo(), OC.coo()):
        assert np.array_equal(host(got), want)
    assert C.is_coalesced()
    E = A * ps.SparseTensor(row=dev(np.array([MB - 1], I64)), col=dev(np.array([NB - 1], I64)),
                            value=dev(np.ones(shape(1)), dtype), sparse_sizes=(MB, NB))
    assert E.nnz() in (0, 1)


@pytest.mark.parametrize("reduce", ["sum", "mean", "min", "max"])
@pytest.mark.parametrize("with_value", [True, False])
def test_to_symmetric_matches_oracle(ps, reduce, with_value):
    rng = np.random.default_rng(41)
    M, N = 900, 1100
    row, col = _random_coo(rng, M, N, 50000)
    val = rng.integers(0, 1 << 16, row.size).astype(np.float64) if with_value else None
    A = ps.SparseTensor(row=dev(row), col=dev(col), value=dev(val) if with_value else None, sparse_sizes=(M, N))
    O = np_ref.SparseTensor(row=row, col=col, value=val, sparse_sizes=(M, N))
    S, OS = A.to_symmetric(reduce), np_ref.to_symmetric(O, reduce)
    assert tuple(S.sparse_sizes()) == tuple(OS.sparse_sizes()) == (1100, 1100)
    assert np.array_equal(host(S.storage.row()), OS.storage.row())
    assert np.array_equal(host(S.storage.col()), OS.storage.col())
    if with_value:
        np.testing.assert_allclose(host(S.storage.value()), OS.storage.value(), rtol=1e-12)
    else:
        assert S.storage.value() is None
    assert S.is_symmetric()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float64])
def test_add_broadcast(ps, dtype):
    """add.py:12-28, 53-100 against the oracle's restatement."""
    rng = np.random.default_rng(29)
    M, N = 300, 500
    row, col, val = _pair(rng, M, N, 20000)
    tv = dev(val, dtype)
    seen = tv.float().cpu().numpy().astype(np.float64) if dtype != torch.float64 else host(tv)
    orow = dev(rng.random((M, 1)).astype(np.float32), dtype)
    ocol = dev(rng.random((1, N)).astype(np.float32), dtype)
    for other in (orow, ocol):
        A = ps.SparseTensor(row=dev(row), col=dev(col), value=tv.clone(), sparse_sizes=(M, N))
        O = np_ref.SparseTensor(row=row, col=col, value=seen.copy(), sparse_sizes=(M, N))
        want = np_ref.add(O, other.float().cpu().numpy().astype(np.float64)).storage.value()
        rtol = {torch.float32: 1e-6, torch.bfloat16: 8e-3, torch.float64: 1e-12}[dtype]
        out = A.add(other)
        np.testing.assert_allclose(host(out.storage.value()), want, rtol=rtol)
        assert torch.equal(A.storage.value(), tv)  # out of place
        assert torch.equal((A + other).storage.value(), out.storage.value())
        B = A.add_(other)
        assert B is A and torch.equal(A.storage.value(), out.storage.value())
    V = ps.SparseTensor(row=dev(row), col=dev(col), sparse_sizes=(M, N))  # value-less: entries count as 1
    srow, scol = host(V.storage.row()), host(V.storage.col())
    np.testing.assert_allclose(host((V + orow).storage.value()), host(orow)[:, 0][srow] + 1, rtol=8e-3)
    np.testing.assert_allclose(host((V + ocol).storage.value()), host(ocol)[0][scol] + 1, rtol=8e-3)
    with pytest.raises(ValueError):
        V.add(dev(np.ones((3, 3), np.float32)))
    A = ps.SparseTensor(row=dev(row), col=dev(col), value=tv.clone(), sparse_sizes=(M, N))
    w = dev(rng.random(row.size).astype(np.float32), dtype)
    assert torch.equal(A.add_nnz(w, layout="coo").storage.value(), tv + w)
    wide = dev(rng.random((row.size, 3)).astype(np.float32), dtype)
    Aw = ps.SparseTensor(row=dev(row), col=dev(col), value=wide.clone(), sparse_sizes=(M, N))
    o3 = dev(rng.random((M, 1, 3)).astype(np.float32), dtype)
    got = Aw.add(o3).storage.value()
    want = Aw.storage.value().double() + o3[:, 0][Aw.storage.row()].double()
    np.testing.assert_allclose(got.double().cpu().numpy(), want.cpu().numpy(), rtol=8e-3)


# ----------------------------------------------------------------------------- test_eye.py / test_overload.py
@pytest.mark.parametrize("dtype", [torch.float32, torch.float64, torch.int64, torch.bfloat16, torch.float16])
def test_eye_golden(ps, dtype):
    """test/test_eye.py:14-66."""
    g = G.EYE["square"]
    mat = ps.SparseTensor.eye(3, dtype=dtype, device="cuda")
    st = mat.storage
    assert mat.device().type == "cuda" and st.sparse_sizes() == (3, 3)
    assert st.row().tolist() == g["row"] and st.rowptr().tolist() == g["rowptr"] and st.col().tolist() == g["col"]
    assert host(st.value()).tolist() == g["value"] and st.value().dtype == dtype
    assert st.num_cached_keys() == 0
    st = ps.SparseTensor.eye(3, has_value=False, device="cuda").storage
    assert st.row().tolist() == g["row"] and st.rowptr().tolist() == g["rowptr"] and st.col().tolist() == g["col"]
    assert st.value() is None and st.num_cached_keys() == 0
    for key in ("wide", "tall"):
        g = G.EYE[key]
        st = ps.SparseTensor.eye(g["M"], g["N"], fill_cache=True, device="cuda").storage
        assert st.sparse_sizes() == (g["M"], g["N"]) and st.num_cached_keys() == 5
        for name in ("row", "rowptr", "col", "rowcount", "colptr", "colcount", "csr2csc", "csc2csr"):
            assert getattr(st, name)().tolist() == g[name], (key, name)


def test_overload_golden(ps):
    """test/test_overload.py:6-22: a dense [M,1] / [1,N] operand on either side of + and * of a value-less matrix;
    the stored entries count as 1 for + and take the operand's value for *."""
    g = G.OVERLOAD
    row, col = np.array(g["row"], I64), np.array(g["col"], I64)
    mat = ps.SparseTensor(row=dev(row), col=dev(col))
    for shape, picked in (((3, 1), row), ((1, 3), col)):
        other = dev(np.array(g["other"], I64)).view(shape)
        want = np.array(g["other"])[picked]
        for out in (other + mat, mat + other):
            assert out.storage.value().tolist() == (want + 1).tolist()
        for out in (other * mat, mat * other):
            assert out.storage.value().tolist() == want.tolist()
            assert out.storage.row().tolist() == g["row"] and out.storage.col().tolist() == g["col"]


def test_equal_and_to(ps):
    """test/test_tensor.py:91-124: `==` compares sizes, indices and values; `to(dtype)` converts the values.  The
    reference's `to(CPUPlace)` has no counterpart here (GPU-only) and says so."""
    row, col = dev(np.array([0, 0, 0, 1, 1], I64)), dev(np.array([0, 1, 2, 0, 2], I64))
    value = torch.arange(1, 6, device="cuda")
    matA = ps.SparseTensor(row=row, col=col, value=value)
    matB = ps.SparseTensor(row=row, col=col, value=value)
    matC = ps.SparseTensor(row=row, col=dev(np.array([0, 1, 2, 0, 1], I64)), value=value)
    assert matA is not matB and matA == matB
    assert matA is not matC and matA != matC
    assert value.dtype == torch.int64
    mat = matA.to(torch.float32)
    assert mat.storage.value().dtype == torch.float32 and matA.storage.value().dtype == torch.int64
    assert matA.to("cuda", torch.float64).storage.value().dtype == torch.float64
    assert matA.to(torch.zeros(1, dtype=torch.float16, device="cuda")).storage.value().dtype == torch.float16
    assert matA.to("cuda:0").storage.row().device == row.device
    with pytest.raises(RuntimeError):
        matA.to("cpu", torch.float32)
    with pytest.raises(TypeError):
        matA.to()


# ----------------------------------------------------------------------------- converters
@pytest.mark.parametrize("fmt", ["coo", "csr", "csc"])
@pytest.mark.parametrize("has_value", [True, False])
def test_scipy_round_trip(ps, fmt, has_value):
    """tensor.py:796-856: from_scipy keeps rowptr (and colptr for CSC input) and skips the sort; to_scipy in every
    layout gives back the same matrix."""
    import scipy.sparse as sp
    rng = np.random.default_rng(47)
    M, N = 120, 90
    row, col, val = _pair(rng, M, N, 2000)
    mat = getattr(sp.coo_matrix((val, (row, col)), (M, N)), "to" + fmt)()
    A = ps.SparseTensor.from_scipy(mat, has_value=has_value)
    assert A.storage.has_rowptr() and A.storage.has_colptr() == (fmt == "csc")
    assert tuple(A.sparse_sizes()) == (M, N) and A.has_value() == has_value
    O = np_ref.SparseTensor(row=row, col=col, value=val if has_value else None, sparse_sizes=(M, N))
    for got, want in zip(A.coo(), O.coo()):
        assert (got is None) == (want is None)
        if got is not None:
            assert np.array_equal(host(got), want)
    assert np.array_equal(host(A.storage.rowptr()), O.storage.rowptr())
    assert np.array_equal(host(A.storage.colptr()), O.storage.colptr())
    ref = mat if has_value else (mat != 0).astype(np.float32)
    for layout in ("coo", "csr", "csc"):
        back = A.to_scipy(layout=layout, dtype=torch.float32)
        assert back.format == layout and back.shape == (M, N)
        assert (abs(back - ref)).nnz == 0


def test_torch_sparse_round_trip(ps):
    """tensor.py:110-155, 564-591 with torch.sparse standing in for paddle.sparse; test/test_convert.py:22-30."""
    g = G.CONVERT
    index, value, N = dev(np.array(g["index"], I64)), dev(np.array(g["value"], np.float32)), g["N"]
    A = ps.SparseTensor.from_torch_sparse_coo_tensor(torch.sparse_coo_tensor(index, value, (N, N)))
    assert torch.equal(torch.stack(A.coo()[:2]), index) and torch.equal(A.coo()[2], value)
    coo = A.to_torch_sparse_coo_tensor()
    assert torch.equal(coo.to_dense(), A.to_dense())
    csr = A.to_paddle_sparse_csr_tensor()
    assert torch.equal(csr.to_dense(), A.to_dense())
    B = ps.SparseTensor.from_torch_sparse_csr_tensor(csr)
    assert B.storage.has_rowptr() and torch.equal(B.to_dense(), A.to_dense())
    V = ps.SparseTensor(row=index[0], col=index[1], sparse_sizes=(N, N))
    assert torch.equal(V.to_torch_sparse_coo_tensor(torch.float32).to_dense(), (A.to_dense() != 0).float())
    with pytest.raises(NotImplementedError):
        A.to_paddle_sparse_csc_tensor()


# ----------------------------------------------------------------------------- cat
def test_cat_golden(ps):
    """test/test_cat.py:11-66."""
    g = G.CAT
    m1 = ps.SparseTensor(row=dev(np.array(g["row1"], I64)), col=dev(np.array(g["col1"], I64))).fill_cache_()
    m2 = ps.SparseTensor(row=dev(np.array(g["row2"], I64)), col=dev(np.array(g["col2"], I64))).fill_cache_()
    out = ps.cat([m1, m2], dim=0)
    assert out.to_dense().tolist() == g["dim0"]
    assert out.storage.has_row() and out.storage.has_rowptr() and out.storage.has_rowcount()
    assert out.storage.num_cached_keys() == g["dim0_cached"]
    out = ps.cat([m1, m2], dim=1)
    assert out.to_dense().tolist() == g["dim1"]
    assert out.storage.has_row() and not out.storage.has_rowptr()
    assert out.storage.num_cached_keys() == g["dim1_cached"]
    out = ps.cat([m1, m2], dim=(0, 1))
    assert out.to_dense().tolist() == g["diag"]
    assert out.storage.has_row() and out.storage.has_rowptr()
    assert out.storage.num_cached_keys() == g["diag_cached"]
    m1 = m1.set_value_(torch.randn((m1.nnz(), g["value_width"]), device="cuda"), layout="coo")
    out = ps.cat([m1, m1], dim=-1)
    assert list(out.storage.value().shape) == [m1.nnz(), g["value_cat_width"]]
    assert out.storage.has_row() and out.storage.has_rowptr()
    assert out.storage.num_cached_keys() == g["value_cat_cached"]
    with pytest.raises(IndexError):
        ps.cat([m1, m1], dim=3)


@pytest.mark.parametrize("dim", [0, 1, (0, 1)])
@pytest.mark.parametrize("cached", ["all", "first-only", "none"])
@pytest.mark.parametrize("with_value", [True, False])
def test_cat_matches_oracle(ps, dim, cached, with_value):
    """Indices, values and every stitched structure against the oracle, for 3 operands (40 when dim=0: more than one
    launch of the concat kernel), one of them empty; a structure survives only if every operand carries it."""
    rng = np.random.default_rng(43)
    shapes = [(300, 200, 5000), (50, 400, 0), (700, 100, 9000)]
    if dim == 0 and cached == "all":
        shapes = shapes + [(int(rng.integers(1, 40)), int(rng.integers(1, 90)), int(rng.integers(0, 300)))
                           for _ in range(37)]
    gs, os_ = [], []
    for k, (M, N, nnz) in enumerate(shapes):
        row, col, val = _pair(rng, M, N, max(nnz, 1), width=2)
        if nnz == 0:
            row, col, val = row[:0], col[:0], val[:0]
        A = ps.SparseTensor(row=dev(row), col=dev(col), value=dev(val) if with_value else None, sparse_sizes=(M, N))
        O = np_ref.SparseTensor(row=row, col=col, value=val if with_value else None, sparse_sizes=(M, N))
        if cached == "all" or (cached == "first-only" and k == 0):
            A.fill_cache_()
            O.fill_cache_()
        gs.append(A)
        os_.append(O)
    got, want = ps.cat(gs, dim), np_ref.cat(os_, dim)
    assert tuple(got.sparse_sizes()) == tuple(want.sparse_sizes())
    assert sorted(got.storage.cached_keys()) == sorted(want.storage.cached_keys())
    assert got.storage.has_rowptr() == (want.storage._rowptr is not None)
    for g_, w_ in zip(got.coo(), want.coo()):
        assert (g_ is None) == (w_ is None)
        if g_ is not None:
            assert np.array_equal(host(g_), w_)
    for key in ("_rowptr", "_rowcount", "_colptr", "_colcount", "_csr2csc", "_csc2csr"):
        w_ = getattr(want.storage, key)
        g_ = getattr(got.storage, key)
        assert (g_ is None) == (w_ is None), key
        if w_ is not None:
            assert np.array_equal(host(g_), w_), key
    # the stitched structures are the ones a fresh build would compute
    fresh = np_ref.SparseTensor(row=want.coo()[0], col=want.coo()[1], sparse_sizes=want.sparse_sizes()).fill_cache_()
    assert np.array_equal(host(got.storage.rowptr()), fresh.storage.rowptr())
    assert np.array_equal(host(got.storage.colptr()), fresh.storage.colptr())
    assert np.array_equal(host(got.storage.csr2csc()), fresh.storage.csr2csc())


# ----------------------------------------------------------------------------- narrow / __narrow_diag__ / select
_STRUCTS = ("_row", "_rowptr", "_rowcount", "_colptr", "_colcount", "_csr2csc", "_csc2csr")


def _same_structs(got, want):
    assert tuple(got.sparse_sizes()) == tuple(want.sparse_sizes())
    assert sorted(got.storage.cached_keys()) == sorted(want.storage.cached_keys())
    for g_, w_ in zip(got.coo(), want.coo()):
        assert (g_ is None) == (w_ is None)
        if g_ is not None:
            assert np.array_equal(host(g_), w_)
    for key in _STRUCTS:
        w_, g_ = getattr(want.storage, key), getattr(got.storage, key)
        if key in ("_row", "_rowptr"):  # either may be derived lazily on our side; when held it must agree
            if g_ is not None and w_ is not None:
                assert np.array_equal(host(g_), w_), key
            continue
        assert (g_ is None) == (w_ is None), key
        if w_ is not None:
            assert np.array_equal(host(g_), w_), key


@pytest.mark.parametrize("cached", [True, False])
@pytest.mark.parametrize("with_value", [True, False])
def test_narrow_matches_oracle(ps, cached, with_value):
    """narrow.py:11-100 against its restatement: indices, values, sizes and which caches survive, for interior,
    full, single, empty and negative-start ranges on both sparse dimensions, and a value dimension."""
    rng = np.random.default_rng(77)
    M, N = 4000, 3100
    row, col, val = _pair(rng, M, N, 90000, width=3)
    A = ps.SparseTensor(row=dev(row), col=dev(col), value=dev(val) if with_value else None, sparse_sizes=(M, N))
    O = np_ref.SparseTensor(row=row, col=col, value=val if with_value else None, sparse_sizes=(M, N))
    if cached:
        A.fill_cache_()
        O.fill_cache_()
    for dim, start, length in ((0, 700, 2100), (0, 0, M), (0, M - 1, 1), (0, 1234, 0), (0, -50, 20),
                               (1, 500, 1700), (1, 0, N), (1, N - 1, 1), (1, 77, 0), (-2 if not with_value else -3, 5, 9)):
        got, want = A.narrow(dim, start, length), np_ref.narrow(O, dim, start, length)
        _same_structs(got, want)
        # what was carried over is what a fresh build computes
        fresh = np_ref.SparseTensor(row=want.coo()[0], col=want.coo()[1], sparse_sizes=want.sparse_sizes()).fill_cache_()
        assert np.array_equal(host(got.storage.rowptr()), fresh.storage.rowptr())
        assert np.array_equal(host(got.storage.colptr()), fresh.storage.colptr())
        assert np.array_equal(host(got.storage.csr2csc()), fresh.storage.csr2csc())
    if with_value:
        got, want = A.narrow(2, 1, 2), np_ref.narrow(O, 2, 1, 2)
        assert np.array_equal(host(got.storage.value()), want.storage.value())
        got = ps.select(A, 0, 17)
        _same_structs(got, np_ref.select(O, 0, 17))
        # slicing through __getitem__ goes the same way (tensor.py:704-786)
        _same_structs(A[100:300, 40:900], np_ref.narrow(np_ref.narrow(O, 0, 100, 200), 1, 40, 860))
    else:
        with pytest.raises(ValueError):
            A.narrow(2, 0, 1)
    with pytest.raises(ValueError):
        A.narrow(0, M - 3, 10)


def test_narrow_owns_its_memory(ps):
    """The reference's `paddle.slice` results are copies: an in-place multiply on a slice must not reach the source."""
    rng = np.random.default_rng(3)
    row, col, val = _pair(rng, 500, 400, 6000)
    A = ps.SparseTensor(row=dev(row), col=dev(col), value=dev(val), sparse_sizes=(500, 400)).fill_cache_()
    before = host(A.storage.value()).copy()
    for B in (A.narrow(0, 100, 200), A.narrow(1, 30, 300), A.__narrow_diag__((0, 0), (500, 400))):
        B.storage.value().mul_(0)
        B.storage.col().add_(1)
        B.storage.rowptr().add_(1)
    assert np.array_equal(host(A.storage.value()), before)
    assert np.array_equal(host(A.storage.col()), col)
    assert int(A.storage.rowptr()[0]) == 0


@pytest.mark.parametrize("cached", ["all", "first-only", "none"])
def test_narrow_diag_inverts_cat_diag(ps, cached):
    """narrow.py:103-168: blocks of a diagonal stack come back with every structure the stack carried, equal to the
    operands' own (golden operands of test/test_cat.py first, then four random blocks, one of them empty)."""
    g = G.CAT
    m1 = ps.SparseTensor(row=dev(np.array(g["row1"], I64)), col=dev(np.array(g["col1"], I64))).fill_cache_()
    m2 = ps.SparseTensor(row=dev(np.array(g["row2"], I64)), col=dev(np.array(g["col2"], I64))).fill_cache_()
    both = ps.cat([m1, m2], dim=(0, 1))
    for m, start in ((m1, (0, 0)), (m2, tuple(m1.sparse_sizes()))):
        back = both.__narrow_diag__(start, tuple(m.sparse_sizes()))
        assert back.to_dense().tolist() == m.to_dense().tolist()
        assert back.storage.num_cached_keys() == 5
        for key in _STRUCTS:
            assert torch.equal(getattr(back.storage, key), getattr(m.storage, key)), key

    rng = np.random.default_rng(91)
    shapes = [(900, 700, 20000), (40, 60, 0), (1500, 1100, 50000), (3, 5, 7)]
    gs, os_ = [], []
    for k, (M, N, nnz) in enumerate(shapes):
        row, col, val = _pair(rng, M, N, max(nnz, 1), width=2)
        if nnz == 0:
            row, col, val = row[:0], col[:0], val[:0]
        A = ps.SparseTensor(row=dev(row), col=dev(col), value=dev(val), sparse_sizes=(M, N))
        O = np_ref.SparseTensor(row=row, col=col, value=val, sparse_sizes=(M, N))
        if cached == "all" or (cached == "first-only" and k == 0):
            A.fill_cache_()
            O.fill_cache_()
        gs.append(A)
        os_.append(O)
    stack, ostack = ps.cat(gs, (0, 1)), np_ref.cat(os_, (0, 1))
    r0 = c0 = 0
    for A, O in zip(gs, os_):
        size = tuple(O.sparse_sizes())
        got, want = stack.__narrow_diag__((r0, c0), size), np_ref.narrow_diag(ostack, (r0, c0), size)
        _same_structs(got, want)
        assert np.array_equal(host(got.storage.value()), O.storage.value())
        assert np.array_equal(host(got.storage.row()), O.storage.row())
        assert np.array_equal(host(got.storage.col()), O.storage.col())
        r0, c0 = r0 + size[0], c0 + size[1]


def test_narrow_cols_full_shape(ps):
    """configs[1]-sized column range: entries kept in order, shifted, counts add up across a partition of the
    columns (size-independent properties; the oracle would take minutes here)."""
    M = N = 2449029
    row, col, val = ps.ops.synth_coo(0x5eed0001, 0, 20_000_000, M, N, "cuda")
    A = ps.SparseTensor(row=row, col=col, value=val, sparse_sizes=(M, N)).coalesce()
    total, mass = 0, 0.0
    for start, length in ((0, 1_000_000), (1_000_000, 449_029), (1_449_029, 1_000_000)):
        B = A.narrow(1, start, length)
        r, c, v = B.coo()
        total += B.nnz()
        mass += float(v.double().sum())
        assert int(c.min()) >= 0 and int(c.max()) < length
        key = r * length + c
        assert bool((key[1:] > key[:-1]).all())
    assert total == A.nnz()
    assert abs(mass - float(A.storage.value().double().sum())) <= 1e-9 * mass


# ----------------------------------------------------------------------------- index_select / masked_select
def test_permute_golden(ps):
    g = G.PERMUTE
    A = ps.SparseTensor(row=ten(g["row"]), col=ten(g["col"]), value=ten(g["value"], torch.float32))
    perm = ten(g["perm"])
    out = A.index_select(0, perm).index_select(1, perm)  # permute.py:8-13
    r, c, v = out.coo()
    assert r.tolist() == g["out_row"] and c.tolist() == g["out_col"] and v.tolist() == g["out_value"]
    r, c, v = A.permute(perm).coo()
    assert r.tolist() == g["out_row"] and c.tolist() == g["out_col"] and v.tolist() == g["out_value"]
    r, c, v = ps.permute(A, perm).coo()
    assert r.tolist() == g["out_row"] and c.tolist() == g["out_col"] and v.tolist() == g["out_value"]


def test_device_dtype_and_grad_helpers(ps):
    """tensor.py:232-242, 472-489, 597-603, 692-693 and the dtype shorthands: same names and meaning; every CPU
    target is refused (no CPU representation)."""
    A = ps.SparseTensor(row=ten([0, 1, 1]), col=ten([1, 0, 2]), sparse_sizes=(2, 3))
    assert not A.requires_grad() and not A.has_value()
    A.requires_grad_()  # fills ones first, like the reference
    assert A.has_value() and A.requires_grad() and A.storage.value().tolist() == [1.0, 1.0, 1.0]
    A.requires_grad_(False)
    assert not A.requires_grad()
    assert A.device_as(A.storage.col()) is A and A.to_device("cuda") is A
    for name, dtype in (("bool", torch.bool), ("byte", torch.uint8), ("char", torch.int8), ("short", torch.int16)):
        assert getattr(A, name)().dtype() == dtype
    assert not A.is_pinned() and not A.is_shared() and A.share_memory_() is A
    for refuse in (A.cpu, A.pin_memory, lambda: A.to_device("cpu")):
        with pytest.raises(RuntimeError):
            refuse()


@pytest.mark.parametrize("with_value", [True, False])
@pytest.mark.parametrize("coalesced", [True, False])
def test_index_and_masked_select(ps, with_value, coalesced):
    rng = np.random.default_rng(31)
    M, N = 350, 270
    row, col, val = _pair(rng, M, N, 15000, coalesced=coalesced)
    if not coalesced:
        o = np.argsort(row * N + col, kind="stable")
        row, col, val = row[o], col[o], val[o]
    kw = dict(sparse_sizes=(M, N), is_sorted=True)
    A = ps.SparseTensor(row=dev(row), col=dev(col), value=dev(val) if with_value else None, **kw)
    O = np_ref.SparseTensor(row=row, col=col, value=val if with_value else None, **kw)
    for dim, size in ((0, M), (1, N)):
        idx = rng.integers(0, size, 500).astype(I64)  # duplicates allowed
        got, want = A.index_select(dim, dev(idx)), np_ref.index_select(O, dim, idx)
        _same_storage(got, want)
        assert sorted(got.storage.cached_keys()) == sorted(want.storage.cached_keys())
        mask = rng.random(size) < 0.4
        got, want = A.masked_select(dim, dev(mask)), np_ref.masked_select(O, dim, mask)
        _same_storage(got, want)
        assert sorted(got.storage.cached_keys()) == sorted(want.storage.cached_keys())
        assert A[dev(idx)].sparse_sizes()[0] == 500 if dim == 0 else True
    eidx = np.sort(rng.choice(row.size, 700, replace=False)).astype(I64)
    emask = rng.random(row.size) < 0.3
    for layout in ("coo", "csc"):
        # layout="csc" leaves the entries in CSC order while the reference still declares them sorted
        # (index_select.py:113-120): rows are compared, the (undefined) rowptr of that result is not
        ok = layout == "coo"
        _same_storage(A.index_select_nnz(dev(eidx), layout), np_ref.index_select_nnz(O, eidx, layout), check_ptr=ok)
        _same_storage(A.masked_select_nnz(dev(emask), layout), np_ref.masked_select_nnz(O, emask, layout),
                      check_ptr=ok)


def test_getitem_and_to_dense(ps):
    rng = np.random.default_rng(2)
    mat = (rng.random((50, 40)) < 0.2) * rng.random((50, 40))
    A = ps.SparseTensor.from_dense(dev(mat.astype(np.float32)))
    assert np.array_equal(host(A.to_dense()), mat.astype(np.float32))
    idx = dev(np.array([3, 1, 3, 7], I64))
    assert np.array_equal(host(A[idx].to_dense()), mat.astype(np.float32)[[3, 1, 3, 7]])
    assert np.array_equal(host(A[:, idx].to_dense()), mat.astype(np.float32)[:, [3, 1, 3, 7]])
    assert np.array_equal(host(A[5:17, -9:].to_dense()), mat.astype(np.float32)[5:17, -9:])
    assert np.array_equal(host(A[3].to_dense()), mat.astype(np.float32)[3:4])
    m = dev(rng.random(50) < 0.5)
    assert np.array_equal(host(A[m].to_dense()), mat.astype(np.float32)[host(m)])
    assert A == A.clone() and not (A == A.t())
    assert A.nnz() == int((mat != 0).sum()) and A.sizes() == [50, 40] and A.dtype() == torch.float32


# ----------------------------------------------------------------------------- sample.py
def test_sample_adj_api(ps):
    g = G.SAMPLE_ADJ
    A = ps.SparseTensor(row=ten(g["row"]), col=ten(g["col"]), value=torch.arange(14, device="cuda"),
                        sparse_sizes=g["sizes"])
    out, n_id = A.sample_adj(ten(g["subset"]), num_neighbors=-1)
    assert n_id.tolist() == g["all_n_id"]
    r, c, v = out.coo()
    assert r.tolist() == g["all_row"] and c.tolist() == g["all_col"] and v.tolist() == g["all_val"]
    out, n_id = A.sample_adj(ten(g["subset"]), num_neighbors=2, replace=True)
    assert out.nnz() == g["nnz_replace_k2"]
    out, n_id = A.sample_adj(ten(g["subset"]), num_neighbors=2, replace=False)
    assert out.nnz() == g["nnz_noreplace_k2"]
    s = A.sample(3)
    assert s.shape == (6, 3) and int(s.min()) >= 0 and int(s.max()) < 6  # test_sample.py:8-14


# ----------------------------------------------------------------------------- the headline pipeline
def _pipeline(ps, row, col, val, M, N):
    A = ps.SparseTensor(row=row, col=col, value=val, sparse_sizes=(M, N)).coalesce("add").fill_cache_()
    return A, A.t()


def test_pipeline_config1_full_compare(ps):
    """BASELINE.json configs[0]: N=10k, nnz=100k (20 % duplicates) fp32: coalesce + rowptr/colptr + t() + sum(1)."""
    rng = np.random.default_rng(1)
    M = N = 10000
    row, col = _random_coo(rng, M, N, 100000, 0.2)
    val = rng.random(100000).astype(np.float32)
    want = np_ref.build_pipeline(row, col, val, M, N)
    A, At = _pipeline(ps, dev(row), dev(col), dev(val), M, N)
    st = A.storage
    for k, t_ in dict(row=st.row(), col=st.col(), rowptr=st.rowptr(), csr2csc=st.csr2csc(), colptr=st.colptr(),
                      csc2csr=st.csc2csr(), t_row=At.storage.row(), t_col=At.storage.col()).items():
        assert np.array_equal(host(t_), want[k]), k
    np.testing.assert_allclose(host(st.value()), want["value"], rtol=1e-6)
    np.testing.assert_allclose(host(At.storage.value()), want["t_value"], rtol=1e-6)
    O = np_ref.SparseTensor(row=want["row"], col=want["col"], value=want["value"].astype(np.float64), sparse_sizes=(M, N))
    np.testing.assert_allclose(host(A.sum(1)), np_ref.reduction(O, 1, "sum"), rtol=2e-6)


def test_pipeline_config2_full_size_properties(ps, ops):
    """BASELINE.json configs[1] at full size (2,449,029 nodes, 61,859,140 nnz): size-independent
    properties — sortedness, strictness, checksum conservation, permutation validity, t(t(A)) == A."""
    M = N = 2449029
    nnz = 61859140
    row, col, val = ops.synth_coo(0x5EED0002, 0, nnz, M, N, "cuda")
    A, At = _pipeline(ps, row, col, val, M, N)
    st = A.storage
    assert ops.coo_order(st.row(), st.col()) == (True, True)
    assert ops.coo_order(At.storage.row(), At.storage.col()) == (True, True)
    nz = A.nnz()
    assert 0 < nz <= nnz and int(st.rowptr()[-1]) == nz and int(st.colptr()[-1]) == nz
    # duplicates only merge: total mass and key checksums are conserved
    np.testing.assert_allclose(float(val.double().sum()), float(st.value().double().sum()), rtol=1e-9)
    assert int(st.rowcount().sum()) == nz and int(st.colcount().sum()) == nz
    iota = torch.arange(nz, device="cuda")
    assert torch.equal(st.csr2csc()[st.csc2csr()], iota)
    assert torch.equal(st.col()[st.csr2csc()], At.storage.row()) and torch.equal(st.row()[st.csr2csc()], At.storage.col())
    assert torch.equal(st.value()[st.csr2csc()], At.storage.value())
    assert torch.equal(ops.ptr2ind(st.rowptr(), nz), st.row())
    Att = At.t()
    assert torch.equal(Att.storage.row(), st.row()) and torch.equal(Att.storage.col(), st.col())
    assert torch.equal(Att.storage.value(), st.value())
    # distinct-key count agrees with an independent computation on the packed keys
    keys = torch.unique(row * N + col)
    assert keys.numel() == nz
    assert torch.equal(keys, st.row() * N + st.col())


@pytest.mark.parametrize("rmat", [False, True])
def test_synthetic_generators_match_the_oracle_restatement(ops, rmat):
    """host (oracle/synth.py) and device (csrc/synth.cu) produce the same stream, bit for bit, for
    the uniform and the R-MAT variant, at an offset and with non-power-of-two sizes."""
    from oracle import synth
    M, N, first, n = 2449029, 1234567, 987654321, 200000
    row, col, val = ops.synth_coo(0x5EED0002, first, n, M, N, "cuda", rmat=rmat)
    h_row, h_col, h_val = (synth.synth_rmat if rmat else synth.synth_coo)(0x5EED0002, first, n, M, N)
    assert np.array_equal(host(row), h_row) and np.array_equal(host(col), h_col) and np.array_equal(host(val), h_val)
    if rmat:  # skew: the heaviest row holds far more than its uniform share
        assert np.bincount(h_row).max() > 50 * n / M


def test_pipeline_rmat_small_full_compare(ps, ops):
    """The skewed variant through the whole pipeline against the oracle: hub rows overflow the shared-memory buckets
    of the hybrid coalesce, so this covers its fall-back on realistic skew."""
    from oracle import synth
    M = N = 200000
    nnz = 1500000
    row, col, val = synth.synth_rmat(0x5EED0007, 0, nnz, M, N)
    want = np_ref.build_pipeline(row, col, val, M, N)
    d_row, d_col, d_val = ops.synth_coo(0x5EED0007, 0, nnz, M, N, "cuda", rmat=True)
    A, At = _pipeline(ps, d_row, d_col, d_val, M, N)
    st = A.storage
    for k, t_ in dict(row=st.row(), col=st.col(), rowptr=st.rowptr(), csr2csc=st.csr2csc(), colptr=st.colptr(),
                      csc2csr=st.csc2csr(), t_row=At.storage.row(), t_col=At.storage.col()).items():
        assert np.array_equal(host(t_), want[k]), k
    np.testing.assert_allclose(host(st.value()), want["value"], rtol=2e-6)
    np.testing.assert_allclose(host(At.storage.value()), want["t_value"], rtol=2e-6)


def test_pipeline_config2_rmat_full_size_properties(ps, ops):
    """configs[1] sizes with the R-MAT stream (SURVEY.md 8d's skewed variant): the same size-independent
    properties as the uniform case."""
    M = N = 2449029
    nnz = 61859140
    row, col, val = ops.synth_coo(0x5EED0002, 0, nnz, M, N, "cuda", rmat=True)
    A, At = _pipeline(ps, row, col, val, M, N)
    st = A.storage
    assert ops.coo_order(st.row(), st.col()) == (True, True)
    assert ops.coo_order(At.storage.row(), At.storage.col()) == (True, True)
    nz = A.nnz()
    assert 0 < nz < nnz and int(st.rowptr()[-1]) == nz and int(st.colptr()[-1]) == nz
    np.testing.assert_allclose(float(val.double().sum()), float(st.value().double().sum()), rtol=1e-9)
    iota = torch.arange(nz, device="cuda")
    assert torch.equal(st.csr2csc()[st.csc2csr()], iota)
    assert torch.equal(st.col()[st.csr2csc()], At.storage.row()) and torch.equal(st.row()[st.csr2csc()], At.storage.col())
    assert torch.equal(ops.ptr2ind(st.rowptr(), nz), st.row())
    keys = torch.unique(row * N + col)
    assert keys.numel() == nz and torch.equal(keys, st.row() * N + st.col())
