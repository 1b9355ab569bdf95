"""The reference's own CUDA kernels as the checker: csrc/cuda/convert_cuda.cu compiled UNMODIFIED for sm_100a
(oracle/_ref/libps_ref_cuda.so, `make -C oracle refcuda`) is run on the device next to ps_ind2ptr / ps_ptr2ind -
bit-exact comparison on the configs[1] / configs[2] shapes plus the edge cases of test/test_storage.py:16-32 -
and timed side by side (the "reference GPU baseline" of SURVEY.md 8d).  With PS_REFCUDA_REPORT=<path> the timing
lines are also written to that file (profiles/ evidence); a time measured here is not a bench value."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import native
from tests.golden import reference_vectors as G

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not native.have_ref_cuda(), reason="oracle/_ref/libps_ref_cuda.so not built")]


def _ref_call(fn, src, n_arg, out):
    lib = native.ref_cuda_lib()
    rc = fn(src.data_ptr(), src.numel(), n_arg, out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.refcuda_last_error().decode()
    return out


def ref_ind2ptr(row, M):
    out = torch.empty(M + 1, dtype=torch.int64, device=row.device)
    return _ref_call(native.ref_cuda_lib().refcuda_ind2ptr, row, M, out)


def ref_ptr2ind(ptr, E):
    out = torch.empty(max(E, 1), dtype=torch.int64, device=ptr.device)
    return _ref_call(native.ref_cuda_lib().refcuda_ptr2ind, ptr, E, out)[:E]


def _time(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def test_golden_and_edges(ops):
    g = G.IND2PTR
    row = torch.tensor(g["row"], dtype=torch.int64, device="cuda")
    want = ref_ind2ptr(row, g["M"])
    assert want.tolist() == g["rowptr"]
    assert torch.equal(ops.ind2ptr(row, g["M"]), want)
    assert torch.equal(ops.ptr2ind(want, row.numel()), ref_ptr2ind(want, row.numel()))
    assert ref_ptr2ind(want, row.numel()).tolist() == g["row"]
    # empty input (convert_cuda.cu:26-28) and a single row holding everything
    empty = torch.empty(0, dtype=torch.int64, device="cuda")
    assert torch.equal(ops.ind2ptr(empty, 8), ref_ind2ptr(empty, 8))
    hub = torch.full((5000,), 3, dtype=torch.int64, device="cuda")
    assert torch.equal(ops.ind2ptr(hub, 7), ref_ind2ptr(hub, 7))
    assert torch.equal(ops.ptr2ind(ref_ind2ptr(hub, 7), 5000), ref_ptr2ind(ref_ind2ptr(hub, 7), 5000))


@pytest.mark.parametrize("shape", ["random-small", "skewed", "configs[1]", "configs[2]"])
def test_matches_reference_cuda_kernels(ops, shape):
    g = torch.Generator(device="cuda").manual_seed(11)
    if shape == "random-small":
        M, n = 1000, 20000
        row = torch.randint(0, M, (n,), device="cuda", generator=g).sort().values
    elif shape == "skewed":  # long empty stretches and hub rows: every fill path of ind2ptr
        M, n = 3_000_000, 2_000_000
        row = (torch.rand(n, device="cuda", generator=g) ** 6 * (M - 1000)).long().sort().values
    else:
        M, n = (2449029, 61859140) if shape == "configs[1]" else (232965, 114615892)
        row = torch.randint(0, M, (n,), device="cuda", generator=g).sort().values
    want_ptr = ref_ind2ptr(row, M)
    got_ptr = ops.ind2ptr(row, M)
    assert torch.equal(got_ptr, want_ptr)
    want_row = ref_ptr2ind(want_ptr, n)
    assert torch.equal(want_row, row)
    assert torch.equal(ops.ptr2ind(got_ptr, n), want_row)
    if not shape.startswith("configs"):
        return
    lines = []
    for name, ours, ref, nbytes in (
            ("ind2ptr", lambda: ops.ind2ptr(row, M), lambda: ref_ind2ptr(row, M), 8 * n + 8 * (M + 1)),
            ("ptr2ind", lambda: ops.ptr2ind(want_ptr, n), lambda: ref_ptr2ind(want_ptr, n), 8 * n + 8 * (M + 1))):
        t_ours, t_ref = _time(ours), _time(ref)
        lines.append({"op": name, "shape": shape, "M": M, "nnz": n, "ms_this_repo": round(t_ours, 4),
                      "ms_reference_kernel_sm100a": round(t_ref, 4), "speedup": round(t_ref / t_ours, 2),
                      "GBps_this_repo": round(nbytes / t_ours / 1e6, 1),
                      "GBps_reference_kernel": round(nbytes / t_ref / 1e6, 1)})
        print(json.dumps(lines[-1]))
    path = os.environ.get("PS_REFCUDA_REPORT")
    if path:
        with open(path, "a") as f:
            for ln in lines:
                f.write(json.dumps(ln) + "\n")
