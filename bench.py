#!/usr/bin/env python
"""bench.py — coalesce + CSR/CSC build throughput (nnz/s), the metric of BASELINE.json.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one pass of the hot path over one batch of synthetic COO input, through the public
API:   SparseTensor(row, col, value, sparse_sizes)   (validation; sort-on-construct is deferred)
       .coalesce("add")                               (sort + dedup + rowptr: hybrid coalesce)
       .fill_cache_()                                 (rowcount, csr2csc, colptr, colcount, csc2csr: ps_csc_build)
       .t()                                           (takes the transposed arrays the CSC sort carried)
N = 1 : BASELINE.json configs[1] — ogbn-products shape, 2,449,029 nodes, 61,859,140 nnz, fp32.
N > 1 : the row-range-partitioned build (configs[4] shape, papers100M: 111,059,956 nodes,
        201,960,734 nnz per GPU = 1.6157 B nnz at N = 8), weak scaling, one all-to-all-v (over peer memory, fused
        with the partition pass; NCCL all_to_all_single is the fall-back).
`value` = inputs resident in HBM.  `e2e` = the same step from pinned HOST buffers to pinned HOST buffers: H2D of
row / col / value and D2H of EVERY array the step produces (row, col, value, rowptr, colptr, csr2csc, csc2csr and the
transposed triple at N = 1) inside the timed region, every step, run as a double-buffered loader would (copies of
neighbouring steps overlap the build on the copy engines); the strictly serial figure is reported beside it.
A line is printed only if the run passed its parity gate: at N = 1 every array of one step is
compared with the CPU oracle on the full configuration (the same oracle pass is the timed `cpu_baseline`).
Prints ONE JSON line (rank 0).  `--impl reference` times the reference's CPU path (NumPy restatement of its Python
op chain + its own csrc/cpu compiled unmodified where available) on the FULL configs[1] input, as many steps as fit
a few minutes.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

PRODUCTS = dict(name="ogbn-products-shape", M=2449029, N=2449029, nnz=61859140, seed=0x5EED0002)
PAPERS = dict(name="ogbn-papers100M-shape", M=111059956, N=111059956, nnz_per_gpu=201960734, seed=0x5EED0005)
# SURVEY.md 8d figure A: compulsory bytes of the whole pipeline per input nnz at configs[1] (no duplicates)
PIPELINE_BYTES_PER_NNZ = 128.6
# without t() (the sharded step builds CSR + CSC structures per shard, no transposed copy): 128.6 - 48
PIPELINE_LOCAL_BYTES_PER_NNZ = 80.6
CPU_SAMPLE_NNZ = 8_000_000      # bounded CPU sample where the full configuration cannot be timed (N > 1, --quick-cpu)
REFERENCE_BUDGET_S = 170.0      # the reference arm runs as many full-size steps as fit this (at least one)
# dram__bytes_read.sum + dram__bytes_write.sum of rs_pass_kernel, average per launch over the 5 launches of one
# step, from the committed `ncu --set full` capture (profiles/r02y_step_kernels_ncu.txt: 8.21 GB over the 5
# launches, against 8.41 GB modelled: the smaller passes partly hit L2): updated whenever the capture is
NCU_TRAFFIC_PER_LAUNCH = 1.642e9


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """Samples SM clocks and throttle reasons with NVML while the timed region runs."""

    def __init__(self, index: int):
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop = threading.Event()
        self._thr = None

    def __enter__(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()
        except Exception:  # NVML missing: fall back to one nvidia-smi sample at exit
            self._nv = None
        return self

    def _run(self):
        nv = self._nv
        names = {
            nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
            nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
            nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake",
        }
        while not self._stop.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)))
                mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.01)

    def __exit__(self, *exc):
        self._stop.set()
        if self._thr is not None:
            self._thr.join()
        if not self.samples:
            try:
                out = subprocess.check_output(
                    ["nvidia-smi", f"--id={self.index}", "--query-gpu=clocks.sm,clocks.max.sm",
                     "--format=csv,noheader,nounits"], text=True).strip().split(",")
                self.samples, self.max_mhz = [int(out[0])], int(out[1])
            except Exception:
                pass

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# =============================================================================== the CPU reference arm
def cpu_pipeline(nnz: int, cfg=PRODUCTS, first: int = 0):
    """The reference's CPU path (oracle/np_ref.build_pipeline: storage.py / transpose.py op chain restated in NumPy,
    ind2ptr from the reference's own csrc/cpu compiled unmodified where oracle/_ref was built) on entries
    [first, first + nnz) of the synthetic stream at the configuration's sparse sizes.
    Returns (seconds, kind, inputs, outputs)."""
    from oracle import native, np_ref, synth
    row, col, val = synth.synth_coo(cfg["seed"], first, nnz, cfg["M"], cfg["N"])
    t0 = time.perf_counter()
    out = np_ref.build_pipeline(row, col, val, cfg["M"], cfg["N"])
    dt = time.perf_counter() - t0
    kind = "port"  # the Python op chain is restated (Paddle is not installable); see `detail`
    return dt, kind, (row, col, val), out


CPU_DETAIL = ("NumPy restatement of the reference's storage.py / transpose.py op chain (np.argsort for its three "
              "argsorts, reduceat for segment_csr, fancy-index gathers) + the reference's own csrc/cpu ind2ptr "
              "compiled unmodified when oracle/_ref is present; single-threaded like the reference's native code")


def reference_workload(n_gpus: int, nnz: int, full: bool):
    if n_gpus <= 1:
        c = PRODUCTS
        what = "every entry" if full else f"the first {nnz} entries"
        return {"workload": f"{c['name']}: {c['M']} nodes, {c['nnz']} nnz, fp32; coalesce(add) + rowptr + csr2csc + colptr + "
                            f"csc2csr + t() on the host CPU; {what} of the stream per step"
                            + ("" if full else " (BOUNDED SAMPLE, not the full configuration)")}
    c = PAPERS
    return {"workload": f"{c['name']}: {c['M']} nodes; CPU arm times the first {nnz} entries of rank 0's "
                        f"{c['nnz_per_gpu']}-entry stream at full sparse sizes per step (BOUNDED SAMPLE of one shard's "
                        "input; no exchange on the CPU)", "parallelism": f"row-range x{n_gpus} on the GPU arm"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import native
    native.port_lib()
    multi = args.gpus > 1
    cfg = PAPERS if multi else PRODUCTS
    # calibration / warm-up on a small prefix (also pages the libraries in)
    t_cal, _, _, _ = cpu_pipeline(2_000_000, cfg)
    full_nnz = cfg["nnz_per_gpu"] if multi else cfg["nnz"]
    est_full = t_cal * (full_nnz / 2e6) * 1.25  # argsort is n log n
    full = (not multi) and not args.quick_cpu and est_full <= REFERENCE_BUDGET_S
    nnz = full_nnz if full else (2 * CPU_SAMPLE_NNZ if multi else CPU_SAMPLE_NNZ)
    times, kind, t_begin = [], "port", time.perf_counter()
    while len(times) < max(args.steps, 1):  # as many steps as fit the budget, at least one
        dt, kind, _, _ = cpu_pipeline(nnz, cfg)
        times.append(dt)
        if time.perf_counter() - t_begin + dt > REFERENCE_BUDGET_S:
            break
    steps = len(times)
    ms = 1e3 * sum(times) / len(times)
    value = nnz / (ms * 1e-3)
    sample = (("the full configuration" if full else f"first {nnz} entries of the stream at full sparse sizes")
              + f", {steps} step(s) of {ms / 1e3:.1f} s; " + CPU_DETAIL)
    line = {
        "impl": "reference", "metric": "coalesce+CSR/CSC build nnz/s", "value": value, "unit": "nnz/s",
        "n_gpus": args.gpus, "steps": steps, "steps_requested": args.steps, "warmup": 1, "ms_per_step": ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int64 keys + f32 values",
        "data": "synthetic", "config": reference_workload(args.gpus, nnz, full),
        "cpu_baseline": {"value": value, "unit": "nnz/s", "cores": 1, "kind": kind, "sample": sample,
                         "host_cores_available": os.cpu_count(), "full_config": full},
        "e2e": {"value": value, "unit": "nnz/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def workload_config(n_gpus: int):
    if n_gpus <= 1:
        c = PRODUCTS
        return {"workload": f"{c['name']}: {c['M']} nodes, {c['nnz']} nnz, fp32; SparseTensor(...) + coalesce(add) + "
                            "fill_cache_ (rowptr, csr2csc, colptr, csc2csr, counts) + t() through the public API",
                "l2": "inputs (1.24 GB) and every intermediate exceed the 126 MB L2; no explicit flush"}
    c = PAPERS
    return {"workload": f"{c['name']} row-partitioned over {n_gpus} GPUs: {c['M']} nodes, {c['nnz_per_gpu']} nnz per GPU "
                        f"({c['nnz_per_gpu'] * n_gpus} total), fp32; splitter histogram + all-to-all-v of the tuples (the "
                        "partition pass writes straight into the peers' symmetric receive buffers over NVLink; NCCL for the "
                        "histogram all-reduce and the counts) + local coalesce + CSR/CSC build per row shard",
            "l2": "per-GPU inputs (4.0 GB) exceed the 126 MB L2; no explicit flush", "parallelism": f"row-range x{n_gpus}"}


# =============================================================================== our arm
def pipeline_step(ps, row, col, val, M, N):
    A = ps.SparseTensor(row=row, col=col, value=val, sparse_sizes=(M, N)).coalesce("add").fill_cache_()
    return A, A.t()


def bits_for(n: int) -> int:
    b = 0
    while (1 << b) < n:
        b += 1
    return b


def rs_pass_launches(nnz_in: int, nnz_out: int, M: int, N: int):
    """Algorithmic bytes of every rs_pass_kernel launch of one step, in launch order:
    each pass reads and writes every tuple once, at the width it has in that pass.
      hybrid coalesce (44-bit key, top 15 bits sorted globally, 8 + 7):
        pass A  reads int64 row + int64 col + f32 value (20 B), writes u64 key + u32 value (12 B)
        pass B  reads 12 B, writes the u32 bucket-local key + u32 value (8 B)
      hybrid CSC sort (top 15 of the 22 col bits sorted globally, carrying entry index, row, value):
        pass 1  reads int64 col + int64 row + f32 value (20 B), writes 4 x u32 (16 B)
        pass 2  16 B + 16 B
      csc2csr: the first partition pass of the bucketed inverse is fused into the CSC bucket kernel; the second one
        reads (target, source) as 2 x u32 (8 B) and writes 8 B
    Shapes that leave the hybrid paths (narrow keys, small inputs) are modelled as plain LSD passes."""
    out = []
    kbits = bits_for(M) + bits_for(N)
    if kbits > 32 and nnz_in >= (1 << 16) and kbits - 15 <= 32:  # the hybrid path (coalesce.cu: hybrid_plan)
        out += [nnz_in * (20 + 12), nnz_in * (12 + 8)]
    else:
        p = max((kbits + 7) // 8, 1)
        out += [nnz_in * (20 + 12)] + [nnz_in * 24] * max(p - 2, 0) + ([nnz_in * 24] if p > 1 else [])
    if nnz_out >= (1 << 16) and bits_for(N) - 15 <= 8:  # the hybrid CSC sort (transpose.cu: csc_hybrid)
        out += [nnz_out * (20 + 16), nnz_out * (16 + 16)]
        if bits_for(nnz_out) > 20:
            out += [nnz_out * 16] * (max((bits_for(nnz_out) - 12 + 7) // 8, 1) - 1)
    else:
        pc = max((bits_for(N) + 7) // 8, 1)
        out += [nnz_out * (20 + 16)] + [nnz_out * 32] * max(pc - 2, 0) + [nnz_out * (16 + 28)]
        if bits_for(nnz_out) > 20:
            out += [nnz_out * 16] * max((bits_for(nnz_out) - 12 + 7) // 8, 1)
    return out


def read_profile(lib):
    import ctypes
    out = {}
    for i in range(lib.ps_prof_kinds()):
        ms, n = ctypes.c_double(), ctypes.c_int64()
        lib.ps_prof_read(i, ctypes.byref(ms), ctypes.byref(n))
        if n.value:
            out[lib.ps_prof_name(i).decode()] = (ms.value, n.value)
    return out


def compare_with_oracle(np, got: dict, want: dict):
    """Every array of one step against the oracle: indices / pointers / permutations bit-exact, fp32 values within
    1e-6 relative (north_star).  Returns the list of array names compared; raises AssertionError on a mismatch."""
    for k, w in want.items():
        g = got[k]
        assert g.shape == w.shape, f"parity gate: {k} has shape {g.shape}, the oracle {w.shape}"
        if w.dtype.kind == "f":
            np.testing.assert_allclose(g, w, rtol=1e-6, atol=0, err_msg=f"parity gate: {k}")
        else:
            assert np.array_equal(g, w), f"parity gate: {k} differs from the oracle"
    return sorted(want)


def parity_gate(args, ps, torch, dist, device, rank, world, cfg, row, col, val):
    """N = 1: one full step on the bench inputs, every result array compared with the CPU oracle run on the same
    61.86 M entries (that oracle pass is also the timed cpu_baseline).  With --no-cpu / --quick-cpu the gate runs
    on a prefix of the stream instead and says so.  N > 1: the partitioned build at reduced scale (the same
    sparse sizes, 1 M entries per rank) gathered on rank 0 and compared with the oracle's single-matrix build, and
    - on the full-size shards, after the timed region - the global invariants (run_ours)."""
    import numpy as np
    M, N = cfg["M"], cfg["N"]
    if world == 1:
        full = not (args.no_cpu or args.quick_cpu)
        n = cfg["nnz"] if full else (4_000_000 if args.no_cpu else CPU_SAMPLE_NNZ)
        dt, kind, _, want = cpu_pipeline(n, cfg)
        A, At = pipeline_step(ps, row[:n], col[:n], val[:n], M, N)
        st, tt = A.storage, At.storage
        got = dict(row=st.row(), col=st.col(), value=st.value(), rowptr=st.rowptr(), colptr=st.colptr(),
                   csr2csc=st.csr2csc(), csc2csr=st.csc2csr(), t_row=tt.row(), t_col=tt.col(), t_value=tt.value())
        got = {k: v.cpu().numpy() for k, v in got.items()}
        compared = compare_with_oracle(np, got, want)
        del A, At, st, tt, got
        parity = {"passed": True, "arrays": compared, "nnz_in": n, "nnz_out": int(want["row"].size),
                  "full_config": full, "rule": "indices, pointers, permutations bit-exact; fp32 values rtol 1e-6",
                  "oracle": "oracle/np_ref.build_pipeline on the same synthetic entries"}
        base = None
        if not args.no_cpu:
            base = {"value": n / dt, "unit": "nnz/s", "cores": 1, "kind": kind, "full_config": full,
                    "sample": (("the full configuration" if full else f"first {n} entries of the same stream at full "
                                "sparse sizes") + f", one pass ({dt:.1f} s); " + CPU_DETAIL),
                    "host_cores_available": os.cpu_count()}
        return parity, base
    # ---- N > 1, reduced scale
    from paddle_sparse_b200 import dist as psd
    small = 1_000_000
    r, c, v = ps.ops.synth_coo(cfg["seed"] ^ 0x77, rank * small, small, M, N, device)
    shard = psd.build_partitioned(r, c, v, (M, N), fill_cache=True, transpose=False)
    st = shard.tensor.storage
    mine = dict(row=shard.global_row(), col=st.col(), value=st.value())
    sizes = torch.tensor([st.col().numel(), shard.row_start, shard.row_end], device=device, dtype=torch.int64)
    all_sizes = [torch.empty_like(sizes) for _ in range(world)]
    dist.all_gather(all_sizes, sizes)
    all_sizes = [t.tolist() for t in all_sizes]
    cap = max(s[0] for s in all_sizes)
    gathered = {}
    for k, t in mine.items():
        pad = torch.zeros(cap, dtype=t.dtype, device=device)
        pad[: t.numel()] = t
        bufs = [torch.empty_like(pad) for _ in range(world)]
        dist.all_gather(bufs, pad)
        gathered[k] = bufs
    # shard-local structure checks on every rank
    rows = shard.row_end - shard.row_start
    ok = bool(st.rowptr().numel() == rows + 1 and int(st.rowptr()[-1]) == st.col().numel()
              and torch.equal(st.rowptr(), ps.ops.ind2ptr(st.row(), rows))
              and torch.equal(st.colptr(), ps.ops.ind2ptr(st.col()[st.csr2csc()], N))
              and torch.equal(st.csc2csr()[st.csr2csc()], torch.arange(st.col().numel(), device=device)))
    flag = torch.tensor([int(ok)], device=device)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    assert int(flag) == 1, "parity gate: a shard's rowptr / colptr / permutations are inconsistent"
    parity = {"passed": True}
    if rank == 0:
        from oracle import np_ref, synth
        src = [synth.synth_coo(cfg["seed"] ^ 0x77, g * small, small, M, N) for g in range(world)]
        srow, scol, sval = (np.concatenate([s[k] for s in src]) for k in range(3))
        idx, value = np_ref.coalesce(np.stack([srow, scol]), sval, M, N, "add")
        got = {k: np.concatenate([gathered[k][g][: all_sizes[g][0]].cpu().numpy() for g in range(world)])
               for k in mine}
        compare_with_oracle(np, got, dict(row=idx[0], col=idx[1], value=value))
        assert all_sizes[0][1] == 0 and all_sizes[-1][2] == M
        assert all(all_sizes[g][2] == all_sizes[g + 1][1] for g in range(world - 1))
        parity = {"passed": True, "arrays": ["row", "col", "value", "shard rowptr / colptr / csr2csc / csc2csr consistency"],
                  "nnz_in": small * world, "nnz_out": int(idx.shape[1]), "full_config": False,
                  "rule": "concatenated shards == the oracle's single-matrix coalesce at reduced scale (same sparse sizes, "
                          f"{small} entries per rank), bit-exact indices, fp32 rtol 1e-6; full-size shards: global "
                          "invariants after the timed region",
                  "oracle": "oracle/np_ref.coalesce"}
    return parity, None


def run_ours(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != max(args.gpus, 1):
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)

    import paddle_sparse_b200 as ps  # raises when the CUDA library is missing: no fallback
    lib = ps._lib.lib()
    hbm_peak, peak_src = peaks()

    if world == 1:
        cfg = PRODUCTS
        M, N, nnz = cfg["M"], cfg["N"], cfg["nnz"]
        row, col, val = ps.ops.synth_coo(cfg["seed"], 0, nnz, M, N, device)

        def step():
            return pipeline_step(ps, row, col, val, M, N)
        units_per_step = nnz
    else:
        from paddle_sparse_b200 import dist as psd
        cfg = PAPERS
        M, N, npg = cfg["M"], cfg["N"], cfg["nnz_per_gpu"]
        row, col, val = ps.ops.synth_coo(cfg["seed"], rank * npg, npg, M, N, device)

        def step():
            return psd.build_partitioned(row, col, val, (M, N), fill_cache=True, transpose=False)
        units_per_step = npg * world

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = None
        for _ in range(steps):
            out = fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1) / steps], device=device, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms), out

    # ---- parity gate: precedes any timing; no line is printed unless it passes
    parity, cpu_baseline = parity_gate(args, ps, torch, dist, device, rank, world, cfg, row, col, val)

    # warm-up: W untimed steps through the SAME loop as the timed region (the previous step's result stays alive
    # while the next one is built), so that the caching allocators already hold the blocks that loop needs; a
    # warm-up that drops each result at once leaves ~12 cudaMalloc calls to the first timed steps
    # (tools/allocprobe.py: 7.0 vs 6.2 ms/step over the first five).
    _, out = timed(step, max(args.warmup, 3))
    del out
    # timed region 1: the headline (no per-kernel events)
    launches0 = int(lib.ps_launch_count())
    with ClockSampler(local_rank) as clocks:
        ms, out = timed(step, args.steps)
    launches = int(lib.ps_launch_count()) - launches0
    value = units_per_step / (ms * 1e-3)
    # timed region 2: the same K steps with CUDA events around every launch of the main kernel families, recorded
    # on the launching stream (ps_prof_enable): per-kernel durations for the roofline of the dominant kernel
    read_profile(lib)
    lib.ps_prof_enable(1)
    del out
    ms_prof, out = timed(step, args.steps)
    lib.ps_prof_enable(0)
    prof = read_profile(lib)

    # ---- end to end: pinned host inputs -> pinned host outputs, every step: H2D of row / col / value, the hot path,
    # D2H of every array the step produced (what a caller that leaves the device receives)
    if world == 1:
        h_row, h_col, h_val = (t.cpu().pin_memory() for t in (row, col, val))
        names = ("row", "col", "value", "rowptr", "colptr", "csr2csc", "csc2csr", "t_row", "t_col", "t_value")

        def result_arrays(A, At):
            st, tt = A.storage, At.storage
            return dict(row=st.row(), col=st.col(), value=st.value(), rowptr=st.rowptr(), colptr=st.colptr(),
                        csr2csc=st.csr2csc(), csc2csr=st.csc2csr(), t_row=tt.row(), t_col=tt.col(), t_value=tt.value())

        ref_arrays = result_arrays(*out)
        h_out = [{k: torch.empty(v.shape, dtype=v.dtype).pin_memory() for k, v in ref_arrays.items()} for _ in range(2)]
        del ref_arrays

        def download(A, At, slot):
            got = result_arrays(A, At)
            n = 0
            for k in names:
                dst = h_out[slot][k][: got[k].numel()]
                dst.copy_(got[k], non_blocking=True)
                n += dst.numel() * dst.element_size()
            return n

        def e2e_step():
            r, c, v = (t.to(device, non_blocking=True) for t in (h_row, h_col, h_val))
            A, At = pipeline_step(ps, r, c, v, M, N)
            download(A, At, 0)
            return A

        copy_stream = torch.cuda.Stream(device=device)
        down_stream = torch.cuda.Stream(device=device)

        def upload():
            """This step's inputs, pinned host -> device on the copy stream; the event marks their arrival."""
            with torch.cuda.stream(copy_stream):
                bufs = tuple(t.to(device, non_blocking=True) for t in (h_row, h_col, h_val))
                ev = torch.cuda.Event()
                ev.record(copy_stream)
            return bufs, ev

        d2h_bytes = [0]

        def e2e_pipelined(steps):
            """The same K steps as a double-buffered loader runs them: every step uploads its own inputs and reads
            ALL of its own results back; the upload of step i+1 and the download of step i-1 (copy engines, one per
            direction) overlap the build of step i (SMs)."""
            cur = torch.cuda.current_stream()
            pending = upload()
            A = None
            for i in range(steps):
                (r, c, v), ev = pending
                pending = upload() if i + 1 < steps else None
                cur.wait_event(ev)
                for t in (r, c, v):
                    t.record_stream(cur)
                A, At = pipeline_step(ps, r, c, v, M, N)
                built = torch.cuda.Event()
                built.record(cur)
                with torch.cuda.stream(down_stream):
                    down_stream.wait_event(built)
                    d2h_bytes[0] = download(A, At, i & 1)
                    for t in result_arrays(A, At).values():
                        t.record_stream(down_stream)
            cur.wait_stream(down_stream)
            return A
        h2d = sum(t.numel() * t.element_size() for t in (h_row, h_col, h_val))
        e2e_steps = max(args.steps // 2, 2)
        for _ in range(2):
            e2e_step()
        serial_ms, _ = timed(e2e_step, e2e_steps)
        e2e_pipelined(2)
        total_ms, _ = timed(lambda: e2e_pipelined(e2e_steps), 1)
        e2e_ms = total_ms / e2e_steps
        e2e = {"value": units_per_step / (e2e_ms * 1e-3), "unit": "nnz/s", "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h_bytes[0], "ms_per_step": e2e_ms, "steps": e2e_steps,
               "mode": "pinned host -> pinned host, double-buffered: every step uploads row / col / value and downloads "
                       "ALL result arrays (row, col, value, rowptr, colptr, csr2csc, csc2csr, t_row, t_col, t_value); "
                       "copies of neighbouring steps overlap the build; K uploads, K builds, K downloads timed",
               "serial_ms_per_step": serial_ms, "serial_value": units_per_step / (serial_ms * 1e-3)}
        nnz_out = out[0].nnz()
        # ---- secondary figure: the same shape from the R-MAT stream (degree skew; hub rows overflow the
        # shared-memory buckets of the hybrid coalesce, which then takes its LSD fall-back).  Uniform is the headline.
        del h_row, h_col, h_val, h_out
        s_row, s_col, s_val = ps.ops.synth_coo(cfg["seed"], 0, nnz, M, N, device, rmat=True)

        def skew_step():
            return pipeline_step(ps, s_row, s_col, s_val, M, N)
        _, s_out = timed(skew_step, 3)
        del s_out
        skew_ms, s_out = timed(skew_step, max(args.steps // 2, 2))
        skewed = {"stream": "R-MAT (0.57, 0.19, 0.19, 0.05) on the same counters and sizes", "ms_per_step": skew_ms,
                  "value": units_per_step / (skew_ms * 1e-3), "unit": "nnz/s", "nnz_out": s_out[0].nnz()}
        del s_row, s_col, s_val, s_out
    else:
        skewed = None
        h_row, h_col, h_val = (t.cpu().pin_memory() for t in (row, col, val))
        # what a caller that leaves the device receives from a shard: its CSR triple (rowptr, col, value) on local rows
        cap_out = int(npg * 1.3) + 1024
        h_res = dict(rowptr=torch.empty(M + 1, dtype=torch.int64).pin_memory(),
                     col=torch.empty(cap_out, dtype=torch.int64).pin_memory(),
                     value=torch.empty(cap_out, dtype=val.dtype).pin_memory())
        d2h_bytes = [0]

        def download(shard):
            st = shard.tensor.storage
            n = 0
            for k, t in (("rowptr", st.rowptr()), ("col", st.col()), ("value", st.value())):
                dst = h_res[k][: t.numel()]
                dst.copy_(t, non_blocking=True)
                n += dst.numel() * dst.element_size()
            d2h_bytes[0] = n

        def e2e_step():
            r, c, v = (t.to(device, non_blocking=True) for t in (h_row, h_col, h_val))
            shard = psd.build_partitioned(r, c, v, (M, N), fill_cache=True, transpose=False)
            download(shard)
            return shard

        copy_stream = torch.cuda.Stream(device=device)

        def upload():
            with torch.cuda.stream(copy_stream):
                bufs = tuple(t.to(device, non_blocking=True) for t in (h_row, h_col, h_val))
                ev = torch.cuda.Event()
                ev.record(copy_stream)
            return bufs, ev

        def e2e_pipelined(steps):
            """As at N = 1: step i+1's upload (copy engine) overlaps step i's exchange + build."""
            cur = torch.cuda.current_stream()
            pending = upload()
            last = None
            for i in range(steps):
                (r, c, v), ev = pending
                pending = upload() if i + 1 < steps else None
                cur.wait_event(ev)
                for t in (r, c, v):
                    t.record_stream(cur)
                last = psd.build_partitioned(r, c, v, (M, N), fill_cache=True, transpose=False)
                download(last)
            return last
        h2d = sum(t.numel() * t.element_size() for t in (h_row, h_col, h_val))
        e2e_steps = max(args.steps // 2, 2)
        e2e_step()
        serial_ms, _ = timed(e2e_step, e2e_steps)
        e2e_pipelined(2)
        total_ms, _ = timed(lambda: e2e_pipelined(e2e_steps), 1)
        e2e_ms = total_ms / e2e_steps
        tot_d2h = torch.tensor([float(d2h_bytes[0])], device=device, dtype=torch.float64)
        dist.all_reduce(tot_d2h, op=dist.ReduceOp.SUM)
        e2e = {"value": units_per_step / (e2e_ms * 1e-3), "unit": "nnz/s", "h2d_bytes_per_step": h2d * world,
               "d2h_bytes_per_step": int(tot_d2h), "ms_per_step": e2e_ms, "steps": e2e_steps,
               "mode": "pinned host -> pinned host per rank, double-buffered uploads: every step uploads the rank's row / "
                       "col / value and downloads its shard's CSR triple (rowptr, col, value on local rows); the CSC-side "
                       "caches stay on the device for the shard-local operators; bytes summed over the ranks",
               "serial_ms_per_step": serial_ms, "serial_value": units_per_step / (serial_ms * 1e-3)}
        del h_row, h_col, h_val, h_res
        nnz_out = None
        shard = out
        sent = torch.tensor([shard.sent, shard.received], device=device, dtype=torch.float64)
        dist.all_reduce(sent, op=dist.ReduceOp.MAX)
        # ---- the same step taken apart: (a) the exchange alone (splitter histogram + all-reduce + fused partition /
        # all-to-all-v), (b) ONE GPU building one shard of this shape with no exchange: the tuples this rank owns
        # after the exchange, as if they had arrived locally - the anchor for scaling efficiency
        def exchange_only():
            return psd.route_tuples(row, col, val, (M, N))
        exchange_only()
        exchange_ms, routed = timed(exchange_only, max(args.steps // 2, 2))
        o_row, o_col, o_val = (t.clone() for t in routed[:3])
        r0 = int(routed[3][rank])
        rows_local = int(routed[3][rank + 1]) - r0
        del routed

        def local_only():
            if o_row.dtype == torch.int32:  # what the fused exchange delivers: 32-bit global ids
                c_row, c_col, c_val, c_rowptr = ps.ops.coalesce_u32(o_row, o_col, r0, o_val, rows_local, N, "add", True)
            else:
                lr = ps.ops.add_scalar(o_row, -r0) if r0 else o_row
                c_row, c_col, c_val, c_rowptr = ps.ops.coalesce(lr, o_col, o_val, rows_local, N, "add", True)
            A = ps.SparseTensor(row=c_row, rowptr=c_rowptr, col=c_col, value=c_val, sparse_sizes=(rows_local, N),
                                is_sorted=True, trust_data=True)
            A.storage.fill_cache_(carry=False)
            return A
        del out, shard  # the shard of the timed loop: its memory is what the anchor's own working set needs
        local_only()
        local_only()    # twice: the first call regrows the allocators' pools for this (different) working set
        local_ms, A_full = timed(local_only, max(args.steps // 2, 2))
        # global invariants of the full-size shards (parity gate, second half)
        st = A_full.storage
        n_loc = st.col().numel()
        key = st.row() * N + st.col()
        inv_ok = bool(n_loc == 0 or (key[1:] > key[:-1]).all())
        inv_ok = inv_ok and int(st.rowptr()[-1]) == n_loc and bool((st.col()[st.csr2csc()][1:] >= st.col()[st.csr2csc()][:-1]).all())
        flag = torch.tensor([int(inv_ok), n_loc], device=device, dtype=torch.int64)
        dist.all_reduce(flag, op=dist.ReduceOp.SUM)
        assert int(flag[0]) == world, "full-size shards: sortedness / rowptr / csr2csc invariants violated"
        parity["full_size_invariants"] = {"strictly_sorted_shards": True, "rowptr_closes": True,
                                          "csr2csc_sorts_columns": True, "total_nnz_out": int(flag[1])}
        del A_full, st, key, o_row, o_col, o_val

    line = {
        "metric": "coalesce+CSR/CSC build nnz/s", "value": value, "unit": "nnz/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "int64 keys + f32 values", "data": "synthetic",
        "config": workload_config(world), "clocks": clocks.summary(), "e2e": e2e,
        "gpu_launches": launches, "parity_checked": bool(parity.get("passed")), "parity": parity,
        "kernels_ms_per_step": {k: round(v[0] / args.steps, 4) for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0])},
        "ms_per_step_with_kernel_events": ms_prof,
    }
    if skewed is not None:
        line["skewed_variant"] = skewed
    if world == 1:
        fam_ms, fam_n = prof.get("rs_pass_kernel", (0.0, 0))
        per_launch = rs_pass_launches(nnz, nnz_out, M, N)
        alg = sum(per_launch)  # per step
        achieved = alg * args.steps / (fam_ms * 1e-3) / 1e9 if fam_ms else None
        line["roofline"] = {
            "bound": "hbm", "kernel": "rs_pass_kernel (one-sweep radix pass)", "achieved": achieved, "peak": hbm_peak,
            "unit": "GB/s", "frac": (achieved / hbm_peak) if achieved else None, "traffic": None,
            "peak_source": peak_src, "launches_per_step": fam_n / args.steps,
            "avg_launch_ms": fam_ms / fam_n if fam_n else None,
            "algorithmic_bytes_per_launch": alg / (fam_n / args.steps) if fam_n else None,
            "share_of_step": (fam_ms / args.steps) / ms_prof if ms_prof else None,
            "launches_modelled_per_step": len(per_launch),
            "traffic_source": "profiles/r02y_step_kernels_ncu.txt (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum, average per launch)",
        }
        line["roofline"]["traffic"] = NCU_TRAFFIC_PER_LAUNCH
        line["pipeline_roofline"] = {
            "algorithmic_bytes_per_nnz": PIPELINE_BYTES_PER_NNZ,
            "achieved_GBps": PIPELINE_BYTES_PER_NNZ * value / 1e9,
            "frac_of_measured_hbm": PIPELINE_BYTES_PER_NNZ * value / 1e9 / hbm_peak,
            "frac_of_nominal_8TBps": PIPELINE_BYTES_PER_NNZ * value / 1e9 / 8000.0,
        }
        if cpu_baseline is not None:
            line["cpu_baseline"] = cpu_baseline
    if world > 1:
        # SURVEY.md 8e combined roofline: local figure-A bytes / HBM peak + bytes sent per GPU / NVLink (770 GB/s
        # measured peer copy per direction, B200_PROFILING.md; 900 nominal)
        local_bytes = PIPELINE_LOCAL_BYTES_PER_NNZ * float(sent[1])
        sent_bytes = 20.0 * float(sent[0])
        t_roof_ms = (local_bytes / (hbm_peak * 1e9) + sent_bytes / 770e9) * 1e3
        line["roofline"] = {
            "bound": "hbm", "kernel": "partitioned build (exchange + local pipeline)", "achieved": None, "peak": hbm_peak,
            "unit": "GB/s", "frac": t_roof_ms / ms, "traffic": None, "peak_source": peak_src,
            "t_roof_ms": t_roof_ms, "max_tuples_sent_per_gpu": int(sent[0]), "max_tuples_received_per_gpu": int(sent[1]),
            "nvlink_GBps_per_direction": 770.0,
            "exchange_ms": exchange_ms, "single_gpu_same_shape_ms": local_ms,
            # on the wire the fused exchange moves 12 bytes per tuple (32-bit row / col + value), not SURVEY 8e's 20
            "wire_bytes_per_tuple": 12 if os.environ.get("PS_DIST_EXCHANGE", "fused") == "fused" else 20,
            "exchange_GBps_per_direction": (12.0 if os.environ.get("PS_DIST_EXCHANGE", "fused") == "fused" else 20.0)
            * float(sent[0]) / (exchange_ms * 1e-3) / 1e9,
            "scaling_anchor": "single_gpu_same_shape_ms = one GPU building one shard of this shape (the tuples the rank "
                              "owns, no exchange); efficiency = single_gpu_same_shape_ms / ms_per_step",
            "note": "frac = (local figure-A bytes / HBM peak + bytes sent / NVLink peak) / measured step time",
        }
    if rank == 0:
        assert line["parity_checked"], "no bench line without a passed parity gate"
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true",
                    help="profiling runs: no cpu_baseline; the parity gate runs on a 4 M-entry prefix")
    ap.add_argument("--quick-cpu", action="store_true",
                    help="CPU legs (parity gate, cpu_baseline, reference arm) on an 8 M-entry prefix instead of the "
                         "full configuration")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
