"""ONE GPU building ONE row shard of the papers100M shape (BASELINE.json configs[4] at 8 ranks: 13.9 M local rows,
111 M columns, 202 M entries), no exchange: the shard-local part of the partitioned build, stage by stage.
   python tools/shardbench.py [ranks=8] [reps=3]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import paddle_sparse_b200 as ps  # noqa: E402
from paddle_sparse_b200 import ops  # noqa: E402

G = int(sys.argv[1]) if len(sys.argv) > 1 else 8
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
N = 111059956
M = (N + G - 1) // G
n = 201960734
row, col, val = ops.synth_coo(0x5EED0005, 0, n, M, N, "cuda")


def step(ev, carry):
    ev[0].record()
    r, c, v, ptr = ops.coalesce(row, col, val, M, N, "add", True)
    ev[1].record()
    A = ps.SparseTensor(row=r, rowptr=ptr, col=c, value=v, sparse_sizes=(M, N), is_sorted=True, trust_data=True)
    if carry:
        A.fill_cache_()
    else:
        A.storage.fill_cache_(carry=False)
    ev[2].record()
    return A


for carry in (True, False):
    for _ in range(2):
        step([torch.cuda.Event(enable_timing=True) for _ in range(3)], carry)
    tot = [0.0, 0.0]
    for _ in range(reps):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        A = step(ev, carry)
        torch.cuda.synchronize()
        for i in range(2):
            tot[i] += ev[i].elapsed_time(ev[i + 1]) / reps
    print(f"G={G} rows {M} nnz {n} -> {A.nnz()} | coalesce {tot[0]:.3f} ms | fill_cache_(carry={carry}) {tot[1]:.3f} ms | "
          f"sum {sum(tot):.3f} ms", flush=True)
    del A

# the same coalesce fed the way the fused exchange feeds it: 32-bit global row ids + offset, 32-bit columns
off = 3 * M
row32 = (row + off).to(torch.int32)  # bits of the uint32 value
col32 = col.to(torch.int32)
for name, fn in (("int64 local rows", lambda: ops.coalesce(row, col, val, M, N, "add", True)),
                 ("uint32 global rows + offset", lambda: ops.coalesce_u32(row32, col32, off, val, M, N, "add", True))):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    print(f"coalesce from {name}: {e0.elapsed_time(e1) / reps:.3f} ms -> {out[0].numel()} entries", flush=True)
