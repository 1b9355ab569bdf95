"""Gradients of `value` through the public API: the reference gets them from Paddle's autograd over
`value[csr2csc]` (transpose.py:15), `segment_csr` (reduce.py:51, storage.py:471), `scatter` (reduce.py:42) and
`gather_csr(...) * value` (mul.py:79-90).  Here every value-producing kernel is wrapped (paddle_sparse_b200/autograd.py);
each test builds the same scalar loss once through this package and once from plain torch index ops on the raw
(row, col, value) triple, and compares the gradients.  float64 values, so the comparison is tight."""
import numpy as np
import pytest
import torch

from tests._util import _random_coo

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ps(ops):
    import paddle_sparse_b200
    return paddle_sparse_b200


def _problem(seed, M=300, N=257, nnz=4000, dup=0.25, dtype=torch.float64):
    rng = np.random.default_rng(seed)
    row, col = _random_coo(rng, M, N, nnz, dup)
    row, col = torch.from_numpy(row).cuda(), torch.from_numpy(col).cuda()
    val = torch.from_numpy(rng.standard_normal(nnz)).to(dtype).cuda()
    return M, N, row, col, val, rng


def _w(rng, n, dtype=torch.float64):
    return torch.from_numpy(rng.standard_normal(n)).to(dtype).cuda()


def test_coalesce_reductions_and_transpose(ps):
    M, N, row, col, val, rng = _problem(1)
    wr, wc = _w(rng, M), _w(rng, N)
    v = val.clone().requires_grad_()
    A = ps.SparseTensor(row=row, col=col, value=v, sparse_sizes=(M, N)).coalesce("add")
    assert A.storage.value().grad_fn is not None
    loss = (A.sum(dim=1) * wr).sum() + (A.sum(dim=0) * wc).sum() + (A.t().sum(dim=1) * wc).sum() + 0.5 * A.sum()
    loss.backward()
    r = val.clone().requires_grad_()
    ref = (torch.zeros(M, dtype=r.dtype, device="cuda").index_add(0, row, r) * wr).sum() \
        + 2 * (torch.zeros(N, dtype=r.dtype, device="cuda").index_add(0, col, r) * wc).sum() + 0.5 * r.sum()
    ref.backward()
    torch.testing.assert_close(loss.detach(), ref.detach())
    torch.testing.assert_close(v.grad, r.grad)


def test_cached_transpose_and_set_value_csc(ps):
    M, N, row, col, val, rng = _problem(2, dup=0.0)
    key = torch.unique(row * N + col)
    row, col = key // N, key % N
    val = val[: row.numel()]
    wc = _w(rng, N)
    v = val.clone().requires_grad_()
    A = ps.SparseTensor(row=row, col=col, value=v, sparse_sizes=(M, N), is_sorted=True).fill_cache_()
    At = A.t()  # takes the values parked by the CSC build: gathered differentiably when grad is needed
    B = A.set_value(At.storage.value() * 3.0, layout="csc")  # back to CSR order through csc2csr
    loss = (At.sum(dim=1) * wc).sum() + (B.sum(dim=0) * wc).sum()
    loss.backward()
    r = val.clone().requires_grad_()
    ref = 4 * (torch.zeros(N, dtype=r.dtype, device="cuda").index_add(0, col, r) * wc).sum()
    ref.backward()
    torch.testing.assert_close(v.grad, r.grad)
    assert torch.equal(B.storage.value().detach(), 3.0 * val)


@pytest.mark.parametrize("reduce", ["mean", "min", "max"])
def test_reductions_with_ties(ps, reduce):
    M, N, row, col, val, rng = _problem(3, dup=0.0)
    key = torch.unique(row * N + col)
    row, col = key // N, key % N
    # many ties; never exactly 0: torch's reference counts its own zero-initialised output as a tie partner there
    val = torch.round(val[: row.numel()] * 2) / 2 + 0.25
    wr, wc = _w(rng, M), _w(rng, N)
    v = val.clone().requires_grad_()
    A = ps.SparseTensor(row=row, col=col, value=v, sparse_sizes=(M, N), is_sorted=True)
    loss = (getattr(A, reduce)(dim=1) * wr).sum() + (getattr(A, reduce)(dim=0) * wc).sum()
    loss.backward()
    r = val.clone().requires_grad_()
    tr = {"mean": "mean", "min": "amin", "max": "amax"}[reduce]
    y1 = torch.zeros(M, dtype=r.dtype, device="cuda").scatter_reduce(0, row, r, tr, include_self=False)
    y0 = torch.zeros(N, dtype=r.dtype, device="cuda").scatter_reduce(0, col, r, tr, include_self=False)
    ref = (y1 * wr).sum() + (y0 * wc).sum()
    ref.backward()
    torch.testing.assert_close(loss.detach(), ref.detach())
    torch.testing.assert_close(v.grad, r.grad)


def test_broadcast_mul_add_grads_to_both_operands(ps):
    M, N, row, col, val, rng = _problem(4, dup=0.0)
    key = torch.unique(row * N + col)
    row, col = key // N, key % N
    val = val[: row.numel()]
    wr, wc = _w(rng, M), _w(rng, N)
    a0, b0 = _w(rng, M), _w(rng, N)
    v, a, b = val.clone().requires_grad_(), a0.clone().requires_grad_(), b0.clone().requires_grad_()
    A = ps.SparseTensor(row=row, col=col, value=v, sparse_sizes=(M, N), is_sorted=True)
    B = A.mul(a.view(M, 1)).add(b.view(1, N)).mul(b.view(1, N)).add(a.view(M, 1))
    loss = (B.sum(dim=1) * wr).sum() + (B.sum(dim=0) * wc).sum()
    loss.backward()
    rv, ra, rb = val.clone().requires_grad_(), a0.clone().requires_grad_(), b0.clone().requires_grad_()
    e = (rv * ra[row] + rb[col]) * rb[col] + ra[row]
    ref = (torch.zeros(M, dtype=e.dtype, device="cuda").index_add(0, row, e) * wr).sum() \
        + (torch.zeros(N, dtype=e.dtype, device="cuda").index_add(0, col, e) * wc).sum()
    ref.backward()
    torch.testing.assert_close(v.grad, rv.grad)
    torch.testing.assert_close(a.grad, ra.grad)
    torch.testing.assert_close(b.grad, rb.grad)
    with pytest.raises(RuntimeError, match="in-place"):
        A.mul_(a.view(M, 1))


def test_select_narrow_and_sparse_ops(ps):
    M, N, row, col, val, rng = _problem(5, dup=0.0)
    key = torch.unique(row * N + col)
    row, col = key // N, key % N
    val = val[: row.numel()]
    idx = torch.from_numpy(rng.integers(0, M, 120)).cuda()
    mask = torch.from_numpy(rng.random(N) < 0.4).cuda()
    v = val.clone().requires_grad_()
    A = ps.SparseTensor(row=row, col=col, value=v, sparse_sizes=(M, N), is_sorted=True)
    S = A.index_select(0, idx)
    Mk = A.masked_select(1, mask)
    Nr = A.narrow(1, 20, 100)
    sym = A.sparse_resize((max(M, N), max(M, N))).to_symmetric("sum")
    loss = S.sum() * 1.5 + (Mk.sum(dim=1) ** 2).sum() + Nr.sum() * 0.25 + (sym.sum(dim=1) ** 2).sum()
    loss.backward()
    r = val.clone().requires_grad_()
    cnt = torch.bincount(idx, minlength=M).to(r.dtype)
    keep = mask[col].to(r.dtype)
    K = max(M, N)
    srow = torch.zeros(K, dtype=r.dtype, device="cuda").index_add(0, row, r).index_add(0, col, r)
    ref = 1.5 * (r * cnt[row]).sum() + (torch.zeros(M, dtype=r.dtype, device="cuda").index_add(0, row, r * keep) ** 2).sum() \
        + 0.25 * (r * ((col >= 20) & (col < 120)).to(r.dtype)).sum() + (srow ** 2).sum()
    ref.backward()
    torch.testing.assert_close(loss.detach(), ref.detach())
    torch.testing.assert_close(v.grad, r.grad)


def test_no_grad_path_is_the_plain_kernel_path(ps):
    M, N, row, col, val, _ = _problem(6)
    A = ps.SparseTensor(row=row, col=col, value=val.float(), sparse_sizes=(M, N)).coalesce("add")
    assert A.storage.value().grad_fn is None and not A.requires_grad()
    v = val.float().requires_grad_()
    with torch.no_grad():
        B = ps.SparseTensor(row=row, col=col, value=v, sparse_sizes=(M, N)).coalesce("add")
    assert B.storage.value().grad_fn is None
    torch.testing.assert_close(B.storage.value(), A.storage.value())
