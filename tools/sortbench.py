"""Micro-benchmark of the building blocks on the configs[1] shape (run on the GPU box):
   python tools/sortbench.py [what ...]     what in {sort, coalesce, csc, all}
Prints the live per-kernel timings of include/ps_abi.h: ps_prof_read."""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import paddle_sparse_b200 as ps  # noqa: E402
from paddle_sparse_b200 import ops  # noqa: E402

lib = ps._lib.lib()
M = N = 2449029
NNZ = 61859140


def prof():
    out = {}
    for i in range(lib.ps_prof_kinds()):
        ms, n = ctypes.c_double(), ctypes.c_int64()
        lib.ps_prof_read(i, ctypes.byref(ms), ctypes.byref(n))
        if n.value:
            out[lib.ps_prof_name(i).decode()] = (round(ms.value / n.value, 4), n.value)
    return out


def timeit(name, fn, reps=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    prof()
    lib.ps_prof_enable(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    lib.ps_prof_enable(0)
    print(f"{name}: {e0.elapsed_time(e1) / reps:.3f} ms  per-launch avg ms (launches): {prof()}", flush=True)


def main():
    what = set(sys.argv[1:]) or {"all"}
    row, col, val = ops.synth_coo(0x5EED0002, 0, NNZ, M, N, "cuda")
    print("variant", os.environ.get("PS_RS_VARIANT", "0"))
    if what & {"sort", "all"}:
        timeline("(reset)")
        timeit("sort_coo", lambda: ops.sort_coo(row, col, val, M, N))
        timeline("sort_coo")
    if what & {"coalesce", "all"}:
        timeit("coalesce", lambda: ops.coalesce(row, col, val, M, N))
    if what & {"csc", "all"}:
        r, c, v = ops.coalesce(row, col, val, M, N)
        timeit("csr2csc", lambda: ops.csr2csc(c, N, True, False))
        perm, colptr, _ = ops.csr2csc(c, N, True, False)
        timeit("invert_perm", lambda: ops.invert_perm(perm))
        timeit("transpose_gather", lambda: ops.transpose_gather(perm, r, c, v, colptr, N))
        timeit("ind2ptr", lambda: ops.ind2ptr(r, M))
        timeit("csc_build(all)", lambda: ops.csc_build(r, c, v, N, True, True))
        timeit("csc_build(light)", lambda: ops.csc_build(r, c, v, N, False, False))


def timeline(label):
    """debug builds (make TIMELINE=1): share of CTA time per phase of the radix pass"""
    buf = (ctypes.c_ulonglong * 16)()
    if lib.ps_debug_timeline(buf) != 0:
        return
    names = ["zero counters", "rank (waits for keys)", "B1", "column scans+publish", "B3", "positions+staging+request",
             "look-back", "B5", "scatter", "(payload staging)", "(positions + key staging)"]
    tot = float(sum(buf[:11])) or 1.0
    print(label, "timeline:", ", ".join(f"{n} {100 * buf[i] / tot:.1f}%" for i, n in enumerate(names)),
          f"| total {tot / 1e6:.1f} Mcycles", flush=True)


if __name__ == "__main__":
    main()
