"""coalesce at the shard shape from int64 and from uint32 sources, once each after a warm-up (for an ncu launch list)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from paddle_sparse_b200 import ops  # noqa: E402

N = 111059956
M = (N + 7) // 8
n = 201960734
row, col, val = ops.synth_coo(0x5EED0005, 0, n, M, N, "cuda")
off = 3 * M
row32 = (row + off).to(torch.int32)
col32 = col.to(torch.int32)
for _ in range(2):
    ops.coalesce(row, col, val, M, N, "add", True)
    torch.cuda.synchronize()
    ops.coalesce_u32(row32, col32, off, val, M, N, "add", True)
    torch.cuda.synchronize()
print("ok")
