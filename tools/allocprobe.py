"""Why the first timed steps after a short warm-up run slower than steady state (run on the GPU box):
step time of successive groups of steps, with the caching allocators' counters beside them."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import paddle_sparse_b200 as ps  # noqa: E402
from paddle_sparse_b200 import ops  # noqa: E402

M = N = 2449029
row, col, val = ops.synth_coo(0x5EED0002, 0, 61859140, M, N, "cuda")


def step():
    A = ps.SparseTensor(row=row, col=col, value=val, sparse_sizes=(M, N)).coalesce("add")
    A.fill_cache_()
    return A, A.t()


def timed(n):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = None
    for _ in range(n):
        out = step()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n, out


mode = sys.argv[1] if len(sys.argv) > 1 else "drop"
out = None
for _ in range(3):
    if mode == "keep":
        out = step()
    else:
        step()
for g in range(6):
    st0 = torch.cuda.memory_stats()
    ms, out = timed(5)
    st1 = torch.cuda.memory_stats()
    print(f"{mode}: steps {3 + 5 * g}-{7 + 5 * g}: {ms:.3f} ms/step, torch cudaMalloc calls +"
          f"{st1['num_device_alloc'] - st0['num_device_alloc']}, frees +{st1['num_device_free'] - st0['num_device_free']}, "
          f"reserved {st1['reserved_bytes.all.current'] / 2**30:.2f} GiB")
