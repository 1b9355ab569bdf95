"""GPU sample_adj vs the oracle: bit-exact against the plain-C port (same Philox mapping, same
insertion-order relabelling) and, on canonicalised per-row sets, against the reference's own
sample_cpu.cpp compiled unmodified (oracle/_ref) — SURVEY.md 8c comparison rule."""
import numpy as np
import pytest

from oracle import native
from tests.golden import reference_vectors as G
from tests._util import I64, dev, host

pytestmark = pytest.mark.gpu


def _graph(rng, M, avg_deg, multi=False):
    deg = rng.poisson(avg_deg, M)
    deg[rng.integers(0, M, max(M // 10, 1))] = 0
    deg[rng.integers(0, M, 3)] = avg_deg * 40  # hubs
    rowptr = np.concatenate([[0], np.cumsum(deg)]).astype(I64)
    col = np.empty(rowptr[-1], I64)
    for r in range(M):
        d = deg[r]
        if multi:
            col[rowptr[r]:rowptr[r + 1]] = np.sort(rng.integers(0, M, d))
        else:
            col[rowptr[r]:rowptr[r + 1]] = np.sort(rng.choice(M, size=min(d, M), replace=False))[:d] if d <= M else 0
    return rowptr, col


def _run_gpu(ops, rowptr, col, idx, k, replace, seed):
    out = ops.sample_adj(dev(rowptr), dev(col), dev(idx), k, replace, seed)
    return [host(t) for t in out]


def _check_invariants(rowptr, col, idx, out, k, replace):
    out_rowptr, out_col, n_id, e_id = out
    S = idx.size
    assert out_rowptr[0] == 0 and out_rowptr[-1] == out_col.size == e_id.size
    assert np.array_equal(n_id[:S], idx)
    assert len(set(n_id[S:].tolist())) == n_id.size - S  # new nodes are distinct
    assert np.array_equal(n_id[out_col], col[e_id])  # relabelling is consistent
    for i in range(S):
        lo, hi = out_rowptr[i], out_rowptr[i + 1]
        assert np.all(np.diff(out_col[lo:hi]) >= 0)  # sample_cpu.cpp:133-140
        r0, r1 = rowptr[idx[i]], rowptr[idx[i] + 1]
        assert np.all((e_id[lo:hi] >= r0) & (e_id[lo:hi] < r1))
        if not replace:
            assert len(set(e_id[lo:hi].tolist())) == hi - lo
            assert hi - lo == (r1 - r0 if k < 0 else min(k, r1 - r0))


def test_sample_adj_golden(ops):
    g = G.SAMPLE_ADJ
    row, col = np.array(g["row"], I64), np.array(g["col"], I64)
    rowptr = native.ind2ptr(row, 6)
    idx = np.array(g["subset"], I64)
    out_rowptr, out_col, n_id, e_id = _run_gpu(ops, rowptr, col, idx, -1, False, 0)
    assert n_id.tolist() == g["all_n_id"] and out_col.tolist() == g["all_col"]
    assert native.ptr2ind(out_rowptr, out_col.size).tolist() == g["all_row"]
    assert e_id.tolist() == g["all_val"]  # value = arange(14) in test_sample.py:20
    assert _run_gpu(ops, rowptr, col, idx, 2, True, 7)[1].size == g["nnz_replace_k2"]
    assert _run_gpu(ops, rowptr, col, idx, 2, False, 7)[1].size == g["nnz_noreplace_k2"]


@pytest.mark.parametrize("k,replace", [(-1, False), (-1, True), (1, False), (5, False), (5, True), (25, False),
                                       (32, False), (33, False), (70, False), (70, True), (0, False)])
@pytest.mark.parametrize("multi", [False, True])
def test_sample_adj_matches_port_bit_exact(ops, k, replace, multi):
    rng = np.random.default_rng(100 + k + 7 * multi)
    M = 3000
    rowptr, col = _graph(rng, M, 12, multi)
    idx = rng.choice(M, 400, replace=False).astype(I64)
    idx[5] = idx[17]  # duplicate seeds: last position wins (sample_cpu.cpp:35-40)
    for seed in (0, 42, 2**63 + 5):
        got = _run_gpu(ops, rowptr, col, idx, k, replace, seed)
        want = native.sample_adj_port(rowptr, col, idx, k, replace, seed)
        for a, b, name in zip(got, want, ("rowptr", "col", "n_id", "e_id")):
            assert np.array_equal(a, b), name
    if not (idx[5] == idx[17]):
        _check_invariants(rowptr, col, idx, got, k, replace)


@pytest.mark.skipif(not native.have_ref(), reason="oracle/_ref not built")
@pytest.mark.parametrize("k,replace", [(-1, False), (3, False), (10, False), (25, False), (10, True), (40, False)])
def test_sample_adj_matches_reference_sets(ops, k, replace):
    rng = np.random.default_rng(55 + k)
    M = 2000
    rowptr, col = _graph(rng, M, 15)
    idx = rng.choice(M, 300, replace=False).astype(I64)
    got = _run_gpu(ops, rowptr, col, idx, k, replace, 42)
    want = native.sample_adj_ref(rowptr, col, idx, k, replace, 42)
    assert np.array_equal(got[0], want[0])
    assert np.array_equal(got[2][: idx.size], want[2][: idx.size])
    assert set(got[2].tolist()) == set(want[2].tolist())
    assert native.canonical_sample(*got) == native.canonical_sample(*want)
    if k < 0:  # CSR order: n_id and local ids are compared bit-exactly (test_sample.py:26-31)
        for a, b in zip(got, want):
            assert np.array_equal(a, b)
    _check_invariants(rowptr, col, idx, got, k, replace)


def test_sample_adj_edge_cases(ops):
    rowptr = np.array([0, 0, 2, 2], I64)
    col = np.array([0, 2], I64)
    e = np.array([], I64)
    out = _run_gpu(ops, rowptr, col, e, 5, False, 1)
    assert out[0].tolist() == [0] and all(t.size == 0 for t in out[1:])
    out = _run_gpu(ops, rowptr, col, np.array([0, 2], I64), 5, True, 1)  # only empty rows
    assert out[0].tolist() == [0, 0, 0] and out[2].tolist() == [0, 2] and out[1].size == 0
    out = _run_gpu(ops, rowptr, col, np.array([1], I64), -1, False, 1)
    assert out[0].tolist() == [0, 2] and out[2].tolist() == [1, 0, 2] and out[1].tolist() == [1, 2]


def test_sample_adj_two_hop_products_like(ops):
    """BASELINE.json configs[3] at reduced scale: 1024 seeds, fanout 25 then 10, replace=False, seed 42."""
    rng = np.random.default_rng(1)
    M = 50000
    rowptr, col = _graph(rng, M, 25, multi=True)
    seeds = rng.choice(M, 1024, replace=False).astype(I64)
    hop1 = _run_gpu(ops, rowptr, col, seeds, 25, False, 42)
    want1 = native.sample_adj_port(rowptr, col, seeds, 25, False, 42)
    assert all(np.array_equal(a, b) for a, b in zip(hop1, want1))
    hop2 = _run_gpu(ops, rowptr, col, hop1[2], 10, False, 42)
    want2 = native.sample_adj_port(rowptr, col, want1[2], 10, False, 42)
    assert all(np.array_equal(a, b) for a, b in zip(hop2, want2))


@pytest.mark.parametrize("k", [1, 4, 25])
def test_sample_uniform_matches_the_oracle(ops, k):
    """sample.py:10-28 as one kernel: bit-exact against the oracle's restatement under the same Philox mapping, with
    and without `subset`, rows without entries included; every draw is a neighbour of its row."""
    import torch

    from oracle import np_ref
    rng = np.random.default_rng(k)
    M = 300
    row = np.sort(rng.integers(0, M, 5000)).astype(np.int64)
    row = row[(row % 7) != 3]  # some rows stay empty
    col = rng.integers(0, M, row.size).astype(np.int64)
    A = np_ref.SparseTensor(row=row, col=col, sparse_sizes=(M, M))
    rowptr, cc, _ = A.csr()
    d_rowptr, d_col = torch.from_numpy(rowptr).cuda(), torch.from_numpy(cc).cuda()
    sub = rng.integers(0, M, 64).astype(np.int64)
    for subset in (None, sub):
        want = np_ref.sample(A, k, subset, seed=42)
        got = ops.sample_uniform(d_rowptr, d_col, None if subset is None else torch.from_numpy(subset).cuda(), k, 42)
        assert np.array_equal(got.cpu().numpy(), want)
    import paddle_sparse_b200 as ps
    T = ps.SparseTensor(row=torch.from_numpy(row).cuda(), col=torch.from_numpy(col).cuda(), sparse_sizes=(M, M))
    out = ps.sample(T, k, torch.from_numpy(sub).cuda(), seed=42)
    assert np.array_equal(out.cpu().numpy(), np_ref.sample(A, k, sub, seed=42))
    assert T.sample(k).shape == (M, k)
