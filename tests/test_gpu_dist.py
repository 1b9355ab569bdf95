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


def _worker(rank, world, port, M, N, per_rank, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        import paddle_sparse_b200 as ps
        from paddle_sparse_b200 import dist as psd
        row, col, val = ps.ops.synth_coo(77, rank * per_rank, per_rank, M, N, f"cuda:{rank}")
        row = (row * row) % M  # skew the rows so that the splitters matter
        shard = psd.build_partitioned(row, col, val, (M, N), fill_cache=True, transpose=True)
        st = shard.tensor.storage
        out[rank] = dict(r0=shard.row_start, r1=shard.row_end, row=shard.global_row().cpu().numpy(),
                         col=st.col().cpu().numpy(), val=st.value().cpu().numpy(), rowptr=st.rowptr().cpu().numpy(),
                         colptr=st.colptr().cpu().numpy(), sizes=shard.tensor.sparse_sizes(),
                         t_sizes=shard.transposed.sparse_sizes(), src=(row.cpu().numpy(), col.cpu().numpy(),
                                                                        val.cpu().numpy()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2])
def test_partitioned_build_matches_single_gpu(cuda, world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    from oracle import np_ref
    M, N, per_rank = 50000, 40000, 300000
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = mp.Manager().dict()
    mp.spawn(_worker, args=(world, port, M, N, per_rank, out), nprocs=world, join=True)
    res = [out[r] for r in range(world)]
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
