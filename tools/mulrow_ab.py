"""A/B of the two width-1 mul_row kernels on the configs[2] shape: the 128-bit kernel (aligned value / out) against the
element-wise one (the same data at an odd element offset), both on the sorted-table expansion.
   python tools/mulrow_ab.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import paddle_sparse_b200 as ps  # noqa: E402
from paddle_sparse_b200 import ops  # noqa: E402

M = N = 232965
row, col, val = ops.synth_coo(0x5EED0003, 0, 114615892, M, N, "cuda")
A = ps.SparseTensor(row=row, col=col, value=val, sparse_sizes=(M, N)).coalesce("add")
rowptr, n = A.storage.rowptr(), A.nnz()
for dt in (torch.float32, torch.bfloat16):
    buf = torch.empty(n + 8, dtype=dt, device="cuda")
    buf[:n] = A.storage.value().to(dt)
    buf[n:] = 1
    other = torch.rand(M, device="cuda").to(dt)
    for name, off in (("128-bit", 0), ("element-wise", 1)):
        v = buf[off:off + n]
        out = torch.empty(n + 8, dtype=dt, device="cuda")[off:off + n]
        for _ in range(3):
            ops.mul_row(v, rowptr, other, out=out)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        for _ in range(20):
            ops.mul_row(v, rowptr, other, out=out)
        b.record()
        torch.cuda.synchronize()
        print(dt, name, round(a.elapsed_time(b) / 20, 4), "ms")
