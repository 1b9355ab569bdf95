"""A few coalesce calls on the configs[1] shape (for ncu captures of the coalesce-side kernels):
   python tools/prof_coalesce.py [reps]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import paddle_sparse_b200 as ps  # noqa: E402
from paddle_sparse_b200 import ops  # noqa: E402

M = N = 2449029
NNZ = 61859140
row, col, val = ops.synth_coo(0x5EED0002, 0, NNZ, M, N, "cuda")
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    r, c, v, p = ops.coalesce(row, col, val, M, N, "add", True)
torch.cuda.synchronize()
print("ok", r.numel(), p[-1].item())
