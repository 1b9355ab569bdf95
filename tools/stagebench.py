"""Per-API-call device time of the bench step on the configs[1] shape (run on the GPU box):
   python tools/stagebench.py [--rmat]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import paddle_sparse_b200 as ps  # noqa: E402
from paddle_sparse_b200 import ops  # noqa: E402

M = N = 2449029
NNZ = 61859140
row, col, val = ops.synth_coo(0x5EED0002, 0, NNZ, M, N, "cuda", rmat="--rmat" in sys.argv)


def step(ev):
    ev[0].record()
    A = ps.SparseTensor(row=row, col=col, value=val, sparse_sizes=(M, N))
    ev[1].record()
    A = A.coalesce("add")
    ev[2].record()
    A.fill_cache_()
    ev[3].record()
    At = A.t()
    ev[4].record()
    return At


for _ in range(3):
    step([torch.cuda.Event(enable_timing=True) for _ in range(5)])
tot = [0.0] * 4
R = 10
for _ in range(R):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
    step(ev)
    torch.cuda.synchronize()
    for i in range(4):
        tot[i] += ev[i].elapsed_time(ev[i + 1]) / R
print("constructor %.3f ms | coalesce %.3f ms | fill_cache_ %.3f ms | t() %.3f ms | sum %.3f ms" % (*tot, sum(tot)))
