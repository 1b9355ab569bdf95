"""Host-side cost of one bench step: the same API sequence on a tiny input (GPU time ~ 0), wall-clocked, plus a
cProfile of it.  What it prints bounds the idle time the GPU sees after each of the step's host syncs.
Run on the GPU box:   python tools/hostcost.py"""
import cProfile
import os
import pstats
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import paddle_sparse_b200 as ps  # noqa: E402
from paddle_sparse_b200 import ops  # noqa: E402

M = N = 2449029
row, col, val = ops.synth_coo(0x5EED0002, 0, 100000, M, N, "cuda")


def step():
    A = ps.SparseTensor(row=row, col=col, value=val, sparse_sizes=(M, N)).coalesce("add")
    A.fill_cache_()
    return A.t()


for _ in range(20):
    step()
torch.cuda.synchronize()
t0 = time.perf_counter()
R = 200
for _ in range(R):
    step()
torch.cuda.synchronize()
print(f"host cost of one step on a 100 K-entry input: {(time.perf_counter() - t0) / R * 1e3:.3f} ms")
pr = cProfile.Profile()
pr.enable()
for _ in range(R):
    step()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
