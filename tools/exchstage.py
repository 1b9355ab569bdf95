"""Where the fused exchange spends its time (run under torchrun on 2+ GPUs): device time between the stages and the
host's wall clock per stage.
   python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/exchstage.py"""
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from paddle_sparse_b200 import dist as psd, ops  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
M = N = 111059956
npg = 201960734
row, col, val = ops.synth_coo(0x5EED0005, rank * npg, npg, M, N, "cuda")


def step(rec):
    def mark(name):
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        rec.append((name, e, time.perf_counter()))
    mark("start")
    shift, buckets = psd.bucket_shift(M)
    hist = ops.dist_row_histogram(row, shift, buckets)
    mark("row histogram (enqueue)")
    dist.all_reduce(hist)
    mark("all-reduce (enqueue)")
    splitters = psd.choose_splitters(hist, world, shift, M)
    mark("splitters on the host (sync)")
    sp = splitters.to(row.device)
    dest, counts = ops.dist_dest(row, sp)
    mark("dest + counts (enqueue)")
    c, buf, vb = psd._peer_plan(counts, val, row.device, None)
    mark("all-gather of the counts + plan (sync)")
    recv_base = torch.cumsum(c, 0) - c
    bases = recv_base[rank].to(row.device)
    buf.h_row.barrier()
    mark("barrier 1")
    ops.dist_scatter_peer(dest, row, col, val, bases, buf.tab_row, buf.tab_col, buf.tab_val, narrow=True)
    mark("scatter pass (enqueue)")
    buf.h_row.barrier()
    mark("barrier 2")
    torch.cuda.synchronize()
    mark("synchronize")


for _ in range(3):
    step([])
acc = {}
R = 5
for _ in range(R):
    dist.barrier()
    torch.cuda.synchronize()
    rec = []
    step(rec)
    for (n0, e0, t0), (n1, e1, t1) in zip(rec[:-1], rec[1:]):
        a = acc.setdefault(n1, [0.0, 0.0])
        a[0] += e0.elapsed_time(e1) / R
        a[1] += (t1 - t0) * 1e3 / R
if rank == 0:
    for n, (dev_ms, host_ms) in acc.items():
        print(f"{n:45s} device {dev_ms:7.3f} ms   host {host_ms:7.3f} ms")
    print(f"{'sum':45s} device {sum(v[0] for v in acc.values()):7.3f} ms   host {sum(v[1] for v in acc.values()):7.3f} ms")
dist.destroy_process_group()
