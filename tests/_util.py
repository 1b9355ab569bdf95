"""Shared helpers of the GPU parity tests."""
import numpy as np
import torch

I64 = np.int64
REDUCES = ["sum", "add", "mean", "min", "max"]
DT = [torch.float32, torch.float64, torch.int32, torch.int64, torch.float16, torch.bfloat16]


def dev(x, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(x))
    if dtype is not None:
        t = t.to(dtype)
    return t.cuda()


def host(t):
    if t.dtype == torch.bfloat16:
        return t.float().cpu().numpy()
    return t.cpu().numpy()


def assert_values(got, want, dtype, exact=False):
    got, want = np.asarray(got, dtype=np.float64), np.asarray(want, dtype=np.float64)
    assert got.shape == want.shape
    if exact:
        assert np.array_equal(got, want)
        return
    rtol = {torch.bfloat16: 1e-3 * 8, torch.float16: 1e-3, torch.float32: 1e-6, torch.float64: 1e-12}.get(dtype, 0)
    # bf16 keeps 8 significant bits: neighbouring values are 2^-8 .. 2^-7 apart relative to their size, so ONE
    # rounding of the fp32-accumulated result is already up to 2^-9 = 2e-3 from the exact sum and north_star's 1e-3
    # cannot be met by any bf16 output.  The op-level tests therefore allow 8e-3 (one ulp, either neighbour); the
    # config-level test (tests/test_gpu_configs.py) is sharper: within 1 ulp of the fp32 accumulation everywhere and
    # EQUAL to round_bf16(oracle) on > 99.99 % of the outputs.
    np.testing.assert_allclose(got, want, rtol=rtol, atol=0)


def _random_coo(rng, M, N, nnz, dup_frac=0.2):
    base = max(int(nnz * (1 - dup_frac)), 1)
    row = rng.integers(0, M, base)
    col = rng.integers(0, N, base)
    pick = rng.integers(0, max(base, 1), nnz - base)
    row = np.concatenate([row, row[pick]]).astype(I64)
    col = np.concatenate([col, col[pick]]).astype(I64)
    p = rng.permutation(nnz)
    return row[p], col[p]
