"""Host logic of the row-range-partitioned build on CPU: splitter selection, and the rank/exchange
plumbing under a world_size-2 gloo group with NumPy stand-ins (from the oracle side) for the two
device kernels.  The product path always uses the CUDA operators; `_ops` exists for this test."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import np_ref
from paddle_sparse_b200 import dist as psd


def test_bucket_shift():
    assert psd.bucket_shift(10) == (0, 10)
    assert psd.bucket_shift(32768) == (0, 32768)
    assert psd.bucket_shift(32769) == (1, 16385)
    s, b = psd.bucket_shift(111059956)
    assert b <= psd.MAX_BUCKETS and ((111059956 + (1 << s) - 1) >> s) == b and s == 12


@pytest.mark.parametrize("G", [1, 2, 4, 8])
def test_choose_splitters_balanced_and_row_aligned(G):
    rng = np.random.default_rng(G)
    M, shift = 100000, 3
    buckets = (M + 7) >> 3
    w = rng.pareto(1.5, buckets) + 0.01  # skewed degrees
    hist = torch.from_numpy(rng.multinomial(2_000_000, w / w.sum()))
    sp = psd.choose_splitters(hist, G, shift, M)
    assert sp[0] == 0 and sp[-1] == M and sp.numel() == G + 1
    assert torch.all(sp[1:] >= sp[:-1])
    assert all(int(r) % (1 << shift) == 0 or int(r) == M for r in sp)
    cum = np.concatenate([[0], np.cumsum(hist.numpy())])
    per = [cum[min((int(sp[g + 1]) + 7) >> 3, buckets)] - cum[min((int(sp[g]) + 7) >> 3, buckets)] for g in range(G)]
    assert sum(per) == 2_000_000
    assert max(per) <= 2_000_000 / G + hist.max().item()  # off by at most one bucket


def test_route_tuples_refuses_ids_beyond_32_bits():
    """The exchange carries row / col as 32-bit words (csrc/dist.cu): larger sparse sizes are refused up front."""
    e = torch.zeros(0, dtype=torch.int64)
    for sizes in ((2 ** 32 + 1, 8), (8, 2 ** 40)):
        with pytest.raises(ValueError, match=r"2\^32"):
            psd.route_tuples(e, e, None, sizes)


def test_choose_splitters_degenerate():
    sp = psd.choose_splitters(torch.zeros(4, dtype=torch.int64), 4, 0, 4)
    assert sp.tolist()[0] == 0 and sp.tolist()[-1] == 4 and all(a <= b for a, b in zip(sp.tolist(), sp.tolist()[1:]))
    one = torch.tensor([0, 1000, 0, 0])  # everything in one bucket: one rank gets it all
    sp = psd.choose_splitters(one, 2, 0, 4)
    assert sp.tolist() == [0, 2, 4]


class _NumpyOps:
    """Stand-ins with the contracts of ps_dist_row_histogram / ps_dist_partition (include/ps_abi.h)."""

    @staticmethod
    def dist_row_histogram(row, shift, buckets):
        return torch.from_numpy(np.bincount(row.numpy() >> shift, minlength=buckets).astype(np.int64))

    @staticmethod
    def dist_partition(row, col, value, splitters):
        dest = np.searchsorted(splitters.numpy()[1:], row.numpy(), side="right")
        order = torch.from_numpy(np.argsort(dest, kind="stable"))
        counts = torch.from_numpy(np.bincount(dest, minlength=splitters.numel() - 1).astype(np.int64))
        return row[order], col[order], None if value is None else value[order], counts


def _worker(rank, world, port, M, per_rank, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(1000 + rank)
        row = torch.from_numpy((rng.pareto(1.2, per_rank) * 50).astype(np.int64) % M)
        col = torch.from_numpy(rng.integers(0, M, per_rank))
        val = torch.from_numpy(rng.random(per_rank).astype(np.float32))
        r, c, v, sp, sent, recv = psd.route_tuples(row, col, val, (M, M), _ops=_NumpyOps)
        out[rank] = dict(row=r.numpy(), col=c.numpy(), val=v.numpy(), sp=sp.numpy(), sent=sent, recv=recv,
                         src=(row.numpy(), col.numpy(), val.numpy()))
    finally:
        dist.destroy_process_group()


def test_route_tuples_gloo_world2():
    world, M, per_rank = 2, 5000, 20000
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = mp.Manager().dict()
    mp.spawn(_worker, args=(world, port, M, per_rank, out), nprocs=world, join=True)
    res = [out[r] for r in range(world)]
    assert np.array_equal(res[0]["sp"], res[1]["sp"])  # identical splitters on every rank
    sp = res[0]["sp"]
    all_src = [np.concatenate([res[r]["src"][k] for r in range(world)]) for k in range(3)]
    for g in range(world):
        r = res[g]
        assert r["recv"] == r["row"].size
        assert np.all((r["row"] >= sp[g]) & (r["row"] < sp[g + 1]))  # ownership
        own = (all_src[0] >= sp[g]) & (all_src[0] < sp[g + 1])
        # arrival order = source rank order, source order kept (stable partition + all-to-all)
        assert np.array_equal(r["row"], all_src[0][own]) and np.array_equal(r["col"], all_src[1][own])
        assert np.array_equal(r["val"], all_src[2][own])
    assert sum(r["recv"] for r in res) == world * per_rank
    assert abs(res[0]["recv"] - res[1]["recv"]) < 0.2 * per_rank  # nnz-balanced despite the skew


def _col_worker(rank, world, port, N, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(2000 + rank)
        res = {}
        for dt in (np.float64, np.int64):
            col = rng.integers(0, N, 400)
            col = col[col % 5 != rank]           # columns only one rank touches, and (c % 5 in {0,1} on both missing) none
            col = col[col % 7 != 3]              # columns nobody touches
            val = (rng.integers(-50, 50, (col.size, 2))).astype(dt)
            res[dt.__name__] = dict(col=col, val=val)
            for red in ("sum", "mean", "min", "max"):
                local = "sum" if red == "mean" else red
                partial = torch.from_numpy(np_ref.scatter(val, col, N, local))
                count = torch.from_numpy(np.bincount(col, minlength=N).astype(np.int64))
                res[dt.__name__][red] = psd.combine_column_partials(partial, count, red).numpy()
        out[rank] = res
    finally:
        dist.destroy_process_group()


def test_combine_column_partials_gloo_world2():
    """dim=0 reductions across row shards == the reduction over the union (reduce.py:40-42 semantics)."""
    world, N = 2, 64
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = mp.Manager().dict()
    mp.spawn(_col_worker, args=(world, port, N, out), nprocs=world, join=True)
    for dt in ("float64", "int64"):
        col = np.concatenate([out[r][dt]["col"] for r in range(world)])
        val = np.concatenate([out[r][dt]["val"] for r in range(world)])
        assert np.setdiff1d(np.arange(N), col).size > 0  # some columns are empty everywhere
        for red in ("sum", "mean", "min", "max"):
            want = np_ref.scatter(val, col, N, red)
            for r in range(world):
                got = out[r][dt][red]
                assert got.dtype == want.dtype
                if dt == "int64":
                    assert np.array_equal(got, want), (dt, red)
                else:
                    np.testing.assert_allclose(got, want, rtol=1e-12, err_msg=f"{dt} {red}")
