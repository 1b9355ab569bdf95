"""Device timeline of bench steps on the configs[1] shape: every kernel / memcpy / memset of the process with
its start, duration and the idle gap before it (CUPTI through torch.profiler; no nsys on the box).
Run on the GPU box:   python tools/timeline.py [steps]
A number taken under the profiler is not a bench value - this is for finding gaps."""
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import paddle_sparse_b200 as ps  # noqa: E402
from paddle_sparse_b200 import ops  # noqa: E402

M = N = 2449029
NNZ = 61859140
row, col, val = ops.synth_coo(0x5EED0002, 0, NNZ, M, N, "cuda")


def step():
    A = ps.SparseTensor(row=row, col=col, value=val, sparse_sizes=(M, N)).coalesce("add")
    A.fill_cache_()
    return A.t()


def main():
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(steps):
            step()
        torch.cuda.synchronize()
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    evs.sort(key=lambda e: e.time_range.start)
    t0 = evs[0].time_range.start
    prev_end = t0
    busy = gaps = 0.0
    print(f"{'start_us':>10} {'gap_us':>8} {'dur_us':>9}  name")
    for e in evs:
        st, en = e.time_range.start, e.time_range.end
        gap = st - prev_end
        print(f"{st - t0:10.1f} {gap:8.1f} {en - st:9.1f}  {e.name[:90]}")
        busy += en - st
        if gap > 0:
            gaps += gap
        prev_end = max(prev_end, en)
    print(f"# {steps} steps: span {(prev_end - t0) / steps / 1e3:.3f} ms/step, busy {busy / steps / 1e3:.3f}, "
          f"idle {gaps / steps / 1e3:.3f}, {len(evs) / steps:.0f} device activities/step")


if __name__ == "__main__":
    main()
