"""Per-stage device time of the row-partitioned build (run under torchrun on the GPU box):
   python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/diststage.py"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import paddle_sparse_b200 as ps  # noqa: E402
from paddle_sparse_b200 import dist as psd, ops  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
M = N = 111059956
npg = 201960734
row, col, val = ops.synth_coo(0x5EED0005, rank * npg, npg, M, N, "cuda")


def ev():
    e = torch.cuda.Event(enable_timing=True)
    e.record()
    return e


def step():
    t = [ev()]
    shift, buckets = psd.bucket_shift(M)
    hist = ops.dist_row_histogram(row, shift, buckets)
    dist.all_reduce(hist)
    splitters = psd.choose_splitters(hist, world, shift, M)
    t.append(ev())
    p_row, p_col, p_val, counts = ops.dist_partition(row, col, val, splitters.to(row.device))
    t.append(ev())
    r_row, r_col, r_val, sent, n_recv, _ = psd._exchange_auto(p_row, p_col, p_val, counts, None)
    t.append(ev())
    r0, r1 = int(splitters[rank]), int(splitters[rank + 1])
    local_row = ops.add_scalar(r_row, -r0) if r0 else r_row
    t.append(ev())
    c = ops.coalesce(local_row, r_col, r_val, max(r1 - r0, 0), N, "add", True)
    t.append(ev())
    A = ps.SparseTensor(row=c[0], rowptr=c[3], col=c[1], value=c[2], sparse_sizes=(max(r1 - r0, 0), N), is_sorted=True,
                        trust_data=True)
    A.fill_cache_()
    t.append(ev())
    return t


names = ["histogram + all-reduce + splitters", "partition by destination", "exchange (counts, row, col, value)",
         "local row ids", "local coalesce", "fill_cache_ (CSC side)"]
for _ in range(3):
    step()
tot = [0.0] * len(names)
R = 4
for _ in range(R):
    dist.barrier()
    t = step()
    torch.cuda.synchronize()
    for i in range(len(names)):
        tot[i] += t[i].elapsed_time(t[i + 1]) / R
if rank == 0:
    for n, v in zip(names, tot):
        print(f"{n:40s} {v:8.3f} ms")
    print(f"{'sum':40s} {sum(tot):8.3f} ms")
dist.destroy_process_group()
