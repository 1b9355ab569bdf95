"""Consumers of the built storage on the shapes of BASELINE.json configs[2] (Reddit shape: broadcast mul, reductions)
and configs[3] (sample_adj on the ogbn-products shape), timed on the device against the compulsory bytes of SURVEY.md
8(d) (figure A).  Run on the GPU box:   python tools/opbench.py [--small] [--rmat]
Prints one JSON line per operation: ms, achieved GB/s on the figure-A bytes, fraction of the measured HBM peak."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import paddle_sparse_b200 as ps  # noqa: E402
from paddle_sparse_b200 import ops  # noqa: E402


def peak():
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    return float(json.load(open(p))["hbm_gbs"]) if os.path.exists(p) else 6650.0


def timeit(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def report(name, ms, nbytes, extra=None):
    gbs = nbytes / ms / 1e6
    line = {"op": name, "ms": round(ms, 4), "figure_A_bytes": int(nbytes), "achieved_GBps": round(gbs, 1),
            "frac_of_measured_hbm": round(gbs / peak(), 3)}
    if extra:
        line.update(extra)
    print(json.dumps(line), flush=True)


def main():
    small = "--small" in sys.argv
    rmat = "--rmat" in sys.argv  # the skewed (R-MAT) generator of SURVEY 8d instead of the uniform one
    # ---- configs[2]: Reddit shape, already coalesced + CSR (build not timed)
    M = N = 232965
    nnz = 114615892 if not small else 8000000
    row, col, val = ops.synth_coo(0x5EED0003, 0, nnz, M, N, "cuda", rmat=rmat)
    A = ps.SparseTensor(row=row, col=col, value=val, sparse_sizes=(M, N)).coalesce("add")
    st = A.storage
    n = st.col().numel()
    rowptr, colv = st.rowptr(), st.col()
    colcount = ops.scatter_add_ones(colv, N)
    del row, col, val
    for dt, bv in ((torch.float32, 4), (torch.bfloat16, 2)):
        v = st.value().to(dt)
        o_row = torch.rand(M, device="cuda").to(dt)
        o_col = torch.rand(N, device="cuda").to(dt)
        tag = "f32" if bv == 4 else "bf16"
        report(f"mul_row[{tag}] (mul.py:76-93, [M,1] operand)", timeit(lambda: ops.mul_row(v, rowptr, o_row)),
               8 * (M + 1) + bv * M + 2 * bv * n)
        report(f"mul_col[{tag}] (mul.py:76-93, [1,N] operand)", timeit(lambda: ops.mul_col(v, colv, o_col)),
               8 * n + bv * N + 2 * bv * n)
        for red in ("sum", "mean", "max"):
            report(f"reduce dim=1 {red}[{tag}] (reduce.py:51 segment_csr)", timeit(lambda: ops.segment_csr(v, rowptr, None, red)),
                   8 * (M + 1) + bv * n + bv * M)
            report(f"reduce dim=0 {red}[{tag}] (reduce.py:42 scatter)",
                   timeit(lambda: ops.scatter(v, colv, 0, None, N, red)), (8 + bv) * n + bv * N)
            if red != "sum":
                report(f"reduce dim=0 {red}[{tag}] with the cached colcount",
                       timeit(lambda: ops.scatter(v, colv, 0, None, N, red, colcount)), (8 + bv) * n + (8 + bv) * N)
    A.fill_cache_()  # parks value[csr2csc]: reduction(dim=0) becomes a segment reduction over colptr
    for red in ("sum", "mean", "max"):
        report(f"reduce dim=0 {red}[f32] through the public API after fill_cache_ (parked CSC-ordered values)",
               timeit(lambda: ps.reduction(A, 0, red)), 8 * (N + 1) + 4 * n + 4 * N)
    A.clear_cache_()
    report("ptr2ind / row() (csrc/convert.cpp:46-76)", timeit(lambda: ops.ptr2ind(rowptr, n)), 8 * (M + 1) + 8 * n)
    report("ind2ptr / rowptr() (csrc/convert.cpp:13-43)", timeit(lambda: ops.ind2ptr(st.row(), M)), 8 * n + 8 * (M + 1))
    del A, st, rowptr, colv
    torch.cuda.empty_cache()

    # ---- configs[3]: sample_adj on the products shape, 1024 seeds, fanout 25 then 10, replace=False, seed 42
    M = N = 2449029
    nnz = 61859140 if not small else 6000000
    row, col, val = ops.synth_coo(0x5EED0002, 0, nnz, M, N, "cuda", rmat=rmat)
    A = ps.SparseTensor(row=row, col=col, value=val, sparse_sizes=(M, N)).coalesce("add")
    rowptr, colv = A.storage.rowptr(), A.storage.col()
    g = torch.Generator(device="cuda").manual_seed(1)
    seeds = torch.randperm(M, device="cuda", generator=g)[:1024].contiguous()
    out1 = ops.sample_adj(rowptr, colv, seeds, 25, False, 42)
    ms1 = timeit(lambda: ops.sample_adj(rowptr, colv, seeds, 25, False, 42), 20)
    n_id = out1[2]
    out2 = ops.sample_adj(rowptr, colv, n_id, 10, False, 42)
    ms2 = timeit(lambda: ops.sample_adj(rowptr, colv, n_id, 10, False, 42), 20)
    e1, e2 = out1[1].numel(), out2[1].numel()
    report("sample_adj hop 1 (csrc/sample.cpp:8-40; 1024 seeds, 25 neighbours)", ms1,
           16 * 1024 + 32 * e1 + 16 * e1 + 8 * n_id.numel() + 8 * 1025,
           {"sampled_edges": e1, "edges_per_s": round(e1 / ms1 * 1e3)})
    report("sample_adj hop 2 (on hop 1's n_id, 10 neighbours)", ms2,
           16 * n_id.numel() + 32 * e2 + 16 * e2 + 8 * out2[2].numel() + 8 * (n_id.numel() + 1),
           {"sampled_edges": e2, "edges_per_s": round(e2 / ms2 * 1e3)})


if __name__ == "__main__":
    main()
