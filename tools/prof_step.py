"""The bench step (configs[1]) a few times, nothing else: the command line ncu captures are taken from.
   python tools/prof_step.py [steps]        (16 kernel launches per step)"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import paddle_sparse_b200 as ps  # noqa: E402
from paddle_sparse_b200 import ops  # noqa: E402

M = N = 2449029
NNZ = 61859140
row, col, val = ops.synth_coo(0x5EED0002, 0, NNZ, M, N, "cuda")
lib = ps._lib.lib()
torch.cuda.synchronize()
n0 = int(lib.ps_launch_count())
for i in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    A = ps.SparseTensor(row=row, col=col, value=val, sparse_sizes=(M, N)).coalesce("add").fill_cache_()
    At = A.t()
    torch.cuda.synchronize()
    if i == 0:
        print("launches per step:", int(lib.ps_launch_count()) - n0)
print("ok", A.nnz(), At.storage.row()[:2].tolist())
