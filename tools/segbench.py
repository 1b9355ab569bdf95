"""Row reductions on the configs[2] shape, device time and host enqueue time per call (is the loop host-bound?):
   [PS_B200_LIB=other.so] python tools/segbench.py [--rmat]"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import paddle_sparse_b200 as ps  # noqa: E402
from paddle_sparse_b200 import ops  # noqa: E402

M = N = 232965
row, col, val = ops.synth_coo(0x5EED0003, 0, 114615892, M, N, "cuda", rmat="--rmat" in sys.argv)
A = ps.SparseTensor(row=row, col=col, value=val, sparse_sizes=(M, N)).coalesce("add")
rowptr = A.storage.rowptr()
n = A.storage.col().numel()
for dt in (torch.float32, torch.bfloat16):
    v = A.storage.value().to(dt)
    for red in ("sum", "max"):
        for _ in range(5):
            ops.segment_csr(v, rowptr, None, red)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(50):
            ops.segment_csr(v, rowptr, None, red)
        e1.record()
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        print(f"{str(dt):16s} {red:4s} device {e0.elapsed_time(e1) / 50 * 1e3:7.1f} us/call   host enqueue {(t1 - t0) / 50 * 1e6:6.1f} us/call   nnz {n}")
