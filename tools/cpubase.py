"""the reference's CPU path (oracle/: its compiled csrc/cpu sources + the NumPy restatement of its
Python op chains) and the CUDA path side by side on every config of BASELINE.json, run on the GPU box so that the CPU
numbers come from that box's host cores.  One JSON line per config.
   python tools/cpubase.py [--skip-c5]           (about 4 minutes of CPU work; 1 core: the reference's CPU code is
                                                 single-threaded, csrc/ holds no `#pragma omp`)
This is a measurement tool (it imports oracle/ like bench.py's cpu_baseline leg does); the product never does."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import paddle_sparse_b200 as ps  # noqa: E402
from oracle import native, np_ref, synth  # noqa: E402
from paddle_sparse_b200 import ops  # noqa: E402


def gpu_ms(fn, reps=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def cpu_s(fn, reps=1):
    t0 = time.perf_counter()
    for _ in range(reps):
        out = fn()
    return (time.perf_counter() - t0) / reps, out


def emit(**kw):
    kw["host_cores_present"] = os.cpu_count()
    kw["cpu_cores_used"] = 1
    print(json.dumps(kw), flush=True)


def build_gpu(row, col, val, M, N):
    A = ps.SparseTensor(row=row, col=col, value=val, sparse_sizes=(M, N)).coalesce("add").fill_cache_()
    return A, A.t()


def c1():
    """configs[0]: random COO N=10k, nnz=100k (20 % duplicates) fp32: coalesce + rowptr/colptr + t() + sum(dim=1)."""
    rng = np.random.default_rng(1)
    M = N = 10000
    base = 80000
    row = rng.integers(0, M, base)
    col = rng.integers(0, N, base)
    dup = rng.integers(0, base, 20000)
    row, col = np.concatenate([row, row[dup]]).astype(np.int64), np.concatenate([col, col[dup]]).astype(np.int64)
    val = rng.random(row.size).astype(np.float32)

    def cpu():
        A = np_ref.SparseTensor(row=row, col=col, value=val, sparse_sizes=(M, N)).coalesce("sum")
        A.storage.rowptr(), A.storage.colptr()
        return np_ref.reduction(np_ref.t(A), 1, "sum")

    d = [torch.from_numpy(x).cuda() for x in (row, col, val)]

    def gpu():
        A = ps.SparseTensor(row=d[0], col=d[1], value=d[2], sparse_sizes=(M, N)).coalesce("add")
        A.storage.rowptr(), A.storage.colptr()
        return ps.reduction(A.t(), 1, "sum")

    s, want = cpu_s(cpu, 5)
    ms = gpu_ms(gpu, 20, 5)
    got = gpu().cpu().numpy()
    assert np.allclose(got, want, rtol=1e-5)
    emit(config="C1", cpu_s=round(s, 5), gpu_ms=round(ms, 4), nnz=int(row.size), gpu_nnz_per_s=round(row.size / ms * 1e3),
         note="launch-latency bound at this size (a few dozen launches and three host syncs)")


def c2():
    """configs[1]: products shape, full pipeline.  CPU: oracle on the full 61.86 M-entry stream (what bench.py times)."""
    M = N = 2449029
    nnz = 61859140
    row, col, val = synth.synth_coo(0x5EED0002, 0, nnz, M, N)
    s, out = cpu_s(lambda: np_ref.build_pipeline(row, col, val, M, N))
    d = ops.synth_coo(0x5EED0002, 0, nnz, M, N, "cuda")
    ms = gpu_ms(lambda: build_gpu(d[0], d[1], d[2], M, N), 10, 3)
    fig_a = 128.6 * nnz
    emit(config="C2", cpu_s=round(s, 2), gpu_ms=round(ms, 3), nnz=nnz, gpu_nnz_per_s=round(nnz / ms * 1e3),
         cpu_nnz_per_s=round(nnz / s), frac_of_8TBps_figure_A=round(fig_a / (ms * 1e-3) / 8e12, 3))
    del out, d
    return row, col, val


def c3():
    """configs[2]: Reddit shape, broadcast mul and reductions, fp32 (CPU: NumPy restatement of mul.py / reduce.py)."""
    M = N = 232965
    nnz = 114615892
    d = ops.synth_coo(0x5EED0003, 0, nnz, M, N, "cuda")
    A = ps.SparseTensor(row=d[0], col=d[1], value=d[2], sparse_sizes=(M, N)).coalesce("add")
    st = A.storage
    rowptr, col, val = st.rowptr().cpu().numpy(), st.col().cpu().numpy(), st.value().cpu().numpy()
    ref = np_ref.SparseTensor(rowptr=rowptr, col=col, value=val, sparse_sizes=(M, N), is_sorted=True)
    o_row = np.random.default_rng(2).random((M, 1)).astype(np.float32)
    o_col = np.ascontiguousarray(o_row.reshape(1, N))
    g_row, g_col = torch.from_numpy(o_row).cuda(), torch.from_numpy(o_col).cuda()
    rows = {}
    for name, cpu, gpu in (
            ("mul [M,1]", lambda: np_ref.mul(ref, o_row), lambda: ps.mul(A, g_row)),
            ("mul [1,N]", lambda: np_ref.mul(ref, o_col), lambda: ps.mul(A, g_col)),
            ("sum dim=1", lambda: np_ref.reduction(ref, 1, "sum"), lambda: ps.reduction(A, 1, "sum")),
            ("mean dim=1", lambda: np_ref.reduction(ref, 1, "mean"), lambda: ps.reduction(A, 1, "mean")),
            ("max dim=1", lambda: np_ref.reduction(ref, 1, "max"), lambda: ps.reduction(A, 1, "max")),
            ("sum dim=0", lambda: np_ref.reduction(ref, 0, "sum"), lambda: ps.reduction(A, 0, "sum")),
            ("max dim=0", lambda: np_ref.reduction(ref, 0, "max"), lambda: ps.reduction(A, 0, "max"))):
        s, _ = cpu_s(cpu)
        rows[name] = {"cpu_s": round(s, 3), "gpu_ms": round(gpu_ms(gpu, 10, 3), 4)}
    emit(config="C3", nnz=int(col.size), ops=rows)


def c4(rowptr_col=None):
    """configs[3]: sample_adj on the products graph, 1024 seeds, 25 then 10, replace=False, seed 42.
    CPU: the reference's own sample_cpu.cpp (compiled unmodified) and the plain-C port."""
    M = N = 2449029
    nnz = 61859140
    d = ops.synth_coo(0x5EED0002, 0, nnz, M, N, "cuda")
    A = ps.SparseTensor(row=d[0], col=d[1], value=d[2], sparse_sizes=(M, N)).coalesce("add")
    rowptr, col = A.storage.rowptr(), A.storage.col()
    g = torch.Generator(device="cuda").manual_seed(1)
    seeds = torch.randperm(M, device="cuda", generator=g)[:1024].contiguous()
    h_rowptr, h_col, h_seeds = rowptr.cpu().numpy(), col.cpu().numpy(), seeds.cpu().numpy()

    def gpu():
        o1 = ops.sample_adj(rowptr, col, seeds, 25, False, 42)
        return o1, ops.sample_adj(rowptr, col, o1[2], 10, False, 42)

    def cpu(fn):
        o1 = fn(h_rowptr, h_col, h_seeds, 25, False, 42)
        return o1, fn(h_rowptr, h_col, np.ascontiguousarray(o1[2]), 10, False, 42)

    ms = gpu_ms(gpu, 20, 3)
    o1, o2 = gpu()
    edges = int(o1[1].numel() + o2[1].numel())
    s_port, (p1, p2) = cpu_s(lambda: cpu(native.sample_adj_port), 3)
    assert np.array_equal(p2[1], o2[1].cpu().numpy()) and np.array_equal(p2[2], o2[2].cpu().numpy())
    line = dict(config="C4", gpu_ms=round(ms, 4), sampled_edges=edges, gpu_edges_per_s=round(edges / ms * 1e3),
                cpu_port_s=round(s_port, 5), cpu_port_edges_per_s=round(edges / s_port))
    if native.have_ref():
        s_ref, _ = cpu_s(lambda: cpu(native.sample_adj_ref), 3)
        line.update(cpu_reference_s=round(s_ref, 5), cpu_reference_edges_per_s=round(edges / s_ref))
    emit(**line)


def c5():
    """configs[4]: one papers100M-shaped row shard of the 8-GPU split (13.88 M rows x 111 M cols, 201.96 M entries):
    the CPU builds that shard; the GPU column is bench.py's single_gpu_same_shape_ms (no exchange)."""
    M, N = 13882495, 111059956
    nnz = 201960734
    row, col, val = synth.synth_coo(0x5EED0005, 0, nnz, M, N)
    s, out = cpu_s(lambda: np_ref.build_pipeline(row, col, val, M, N))
    del out
    d = [torch.from_numpy(x).cuda() for x in (row, col, val)]

    def gpu():
        A = ps.SparseTensor(row=d[0], col=d[1], value=d[2], sparse_sizes=(M, N)).coalesce("add")
        A.storage.fill_cache_(carry=False)
        return A

    ms = gpu_ms(gpu, 5, 2)
    emit(config="C5 (one of eight row shards)", cpu_s=round(s, 2), gpu_ms=round(ms, 3), nnz=nnz,
         cpu_nnz_per_s=round(nnz / s), gpu_nnz_per_s=round(nnz / ms * 1e3))


if __name__ == "__main__":
    c1()
    c4()
    c3()
    c2()
    if "--skip-c5" not in sys.argv:
        c5()
