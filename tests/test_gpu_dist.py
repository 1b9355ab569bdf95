"""Multi-GPU parity of the row-range-partitioned build (needs >= 2 GPUs; run under `gpurun --gpus 2`):
the concatenation of the shards must equal the single-GPU build of the union, bit for bit on the
indices, and each shard must carry a self-contained rowptr on local row ids."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, M, N, per_rank, out, exchange="peer"):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), PS_DIST_EXCHANGE=exchange)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        import paddle_sparse_b200 as ps
        from paddle_sparse_b200 import dist as psd
        row, col, val = ps.ops.synth_coo(77, rank * per_rank, per_rank, M, N, f"cuda:{rank}")
        row = (row * row) % M  # skew the rows so that the splitters matter
        shard = psd.build_partitioned(row, col, val, (M, N), fill_cache=True, transpose=True)
        st = shard.tensor.storage
        out[rank] = dict(peer=bool(psd._peer_buffers) and not psd._peer_unavailable,
                         r0=shard.row_start, r1=shard.row_end, row=shard.global_row().cpu().numpy(),
                         col=st.col().cpu().numpy(), val=st.value().cpu().numpy(), rowptr=st.rowptr().cpu().numpy(),
                         colptr=st.colptr().cpu().numpy(), sizes=shard.tensor.sparse_sizes(),
                         t_sizes=shard.transposed.sparse_sizes(), src=(row.cpu().numpy(), col.cpu().numpy(),
                                                                        val.cpu().numpy()))
    finally:
        dist.destroy_process_group()


# fused = the partition pass stores into the peers' buffers (default); peer = partition, then device-to-device copies
# into them; nccl = partition, then all_to_all_single (the fall-back)
@pytest.mark.parametrize("exchange", ["fused", "peer", "nccl"])
@pytest.mark.parametrize("world", [2])
def test_partitioned_build_matches_single_gpu(cuda, world, exchange):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    from oracle import np_ref
    M, N, per_rank = 50000, 40000, 300000
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = mp.Manager().dict()
    mp.spawn(_worker, args=(world, port, M, N, per_rank, out, exchange), nprocs=world, join=True)
    res = [out[r] for r in range(world)]
    assert all(r["peer"] == (exchange != "nccl") for r in res)  # the path that was asked for is the one that ran
    src = [np.concatenate([res[r]["src"][k] for r in range(world)]) for k in range(3)]
    idx, val = np_ref.coalesce(np.stack([src[0], src[1]]), src[2].astype(np.float64), M, N, "add")
    assert res[0]["r0"] == 0 and res[-1]["r1"] == M
    assert all(res[g]["r1"] == res[g + 1]["r0"] for g in range(world - 1))
    assert np.array_equal(np.concatenate([r["row"] for r in res]), idx[0])
    assert np.array_equal(np.concatenate([r["col"] for r in res]), idx[1])
    np.testing.assert_allclose(np.concatenate([r["val"] for r in res]), val, rtol=1e-6)
    for r in res:
        rows = r["r1"] - r["r0"]
        assert r["sizes"] == (rows, N) and r["t_sizes"] == (N, rows)
        assert r["rowptr"].size == rows + 1 and r["rowptr"][-1] == r["row"].size
        assert np.array_equal(r["rowptr"], np.searchsorted(r["row"] - r["r0"], np.arange(rows + 1)))
        assert np.array_equal(r["colptr"], np.concatenate([[0], np.cumsum(np.bincount(r["col"], minlength=N))]))
    n = [r["row"].size for r in res]
    assert max(n) < 1.25 * sum(n) / world  # nnz-balanced shards


def _second_exchange_worker(rank, world, port, M, N, per_rank, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    os.environ.pop("PS_DIST_EXCHANGE", None)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        import paddle_sparse_b200 as ps
        from paddle_sparse_b200 import dist as psd
        row, col, val = ps.ops.synth_coo(91, rank * per_rank, per_rank, M, N, f"cuda:{rank}")
        col = (col * 7) % (N - N // 8)  # the last eighth of the columns stays empty everywhere
        val = torch.floor(val * 1024) - 300  # integer-valued fp32 of both signs: sums exact in any order
        shard = psd.build_partitioned(row, col, val, (M, N), fill_cache=True)
        res = dict(src=(row.cpu().numpy(), col.cpu().numpy(), val.cpu().numpy()))
        for red in ("sum", "mean", "min", "max"):
            res[red] = psd.reduce_columns(shard, red).cpu().numpy()
        T = psd.transpose_partitioned(shard)
        st = T.tensor.storage
        res["t"] = dict(r0=T.row_start, r1=T.row_end, sizes=T.sparse_sizes, row=T.global_row().cpu().numpy(),
                        col=st.col().cpu().numpy(), val=st.value().cpu().numpy(), rowptr=st.rowptr().cpu().numpy())
        out[rank] = res
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2])
def test_column_reductions_and_global_transpose_across_shards(cuda, world):
    """`reduction(dim=0)` (reduce.py:40-42) and `t()` (transpose.py:9-33) of a matrix held as
    row shards equal the single-matrix oracle results."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    from oracle import np_ref
    M, N, per_rank = 30000, 20000, 200000
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = mp.Manager().dict()
    mp.spawn(_second_exchange_worker, args=(world, port, M, N, per_rank, out), nprocs=world, join=True)
    res = [out[r] for r in range(world)]
    src = [np.concatenate([res[r]["src"][k] for r in range(world)]) for k in range(3)]
    A = np_ref.SparseTensor(row=src[0], col=src[1], value=src[2].astype(np.float64), sparse_sizes=(M, N))
    A = A.coalesce("sum")
    for red in ("sum", "mean", "min", "max"):
        want = np_ref.reduction(A, 0, red)
        assert np.all(want[N - N // 8:] == 0)
        for r in res:
            np.testing.assert_allclose(r[red], want, rtol=1e-6 if red == "mean" else 0, err_msg=red)
    At = np_ref.t(A)
    t_row, t_col, t_val = At.coo()
    ts = [r["t"] for r in res]
    assert ts[0]["r0"] == 0 and ts[-1]["r1"] == N and all(t["sizes"] == (N, M) for t in ts)
    assert all(ts[g]["r1"] == ts[g + 1]["r0"] for g in range(world - 1))
    assert np.array_equal(np.concatenate([t["row"] for t in ts]), t_row)
    assert np.array_equal(np.concatenate([t["col"] for t in ts]), t_col)
    assert np.array_equal(np.concatenate([t["val"] for t in ts]), t_val)
    for t in ts:
        rows = t["r1"] - t["r0"]
        assert np.array_equal(t["rowptr"], np.searchsorted(t["row"] - t["r0"], np.arange(rows + 1)))
