"""ps_ptr2ind_noalloc (include/ps_abi.h): the ptr2ind entry point for callers that may not allocate.  It has no
counterpart in the reference; its checker is the existing entry point
ps_ptr2ind, which the other GPU tests hold against the oracle.  Index output: compared bit-exactly, like every index
result in this suite.  Collected last on purpose."""
import ctypes

import numpy as np
import pytest
import torch

from tests._util import I64, dev, host

pytestmark = pytest.mark.gpu

MAX_E = 2147483648  # PS_PTR2IND_NOALLOC_MAX_E


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _stream(t):
    return ctypes.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


def _both(ops, rowptr, E):
    """(ps_ptr2ind, ps_ptr2ind_noalloc) on the same rowptr, straight through the C ABI."""
    lib = ops._lib.lib()
    M = rowptr.numel() - 1
    want = torch.full((E,), -7, dtype=torch.int64, device=rowptr.device)
    got = torch.full((E,), -9, dtype=torch.int64, device=rowptr.device)
    with torch.cuda.device(rowptr.device):
        assert lib.ps_ptr2ind(_ptr(rowptr), M, E, _ptr(want), _stream(rowptr)) == 0, ops._lib.last_error()
        assert lib.ps_ptr2ind_noalloc(_ptr(rowptr), M, E, _ptr(got), _stream(rowptr)) == 0, ops._lib.last_error()
    return want, got


def _rowptr(rng, M, E, hub=False):
    ind = np.sort(rng.integers(0, M, E)).astype(I64)
    if hub and E > 1000:  # a hub row and a long run of empty rows
        ind[: E // 3] = ind[E // 3]
    ptr = np.zeros(M + 1, I64)
    np.cumsum(np.bincount(ind, minlength=M), out=ptr[1:])
    return ptr, ind


def test_two_calls_in_a_row_match_ptr2ind(ops):
    rng = np.random.default_rng(20261105)
    # the first call fills many words of the shared array, the second (smaller, other row structure) must not be
    # disturbed by what the first one left there; then a larger one again
    for M, E, hub in ((100000, 3000017, True), (777, 50001, False), (1 << 20, 4100000, True), (3, 5, False)):
        ptr, ind = _rowptr(rng, M, E, hub)
        want, got = _both(ops, dev(ptr), E)
        assert torch.equal(got, want)
        assert np.array_equal(host(got), ind)
    # the same input twice, back to back, without a synchronisation in between
    ptr, ind = _rowptr(rng, 5000, 1000003, True)
    d = dev(ptr)
    lib = ops._lib.lib()
    a = torch.empty(ind.size, dtype=torch.int64, device="cuda")
    b = torch.empty(ind.size, dtype=torch.int64, device="cuda")
    assert lib.ps_ptr2ind_noalloc(_ptr(d), 5000, ind.size, _ptr(a), _stream(d)) == 0
    assert lib.ps_ptr2ind_noalloc(_ptr(d), 5000, ind.size, _ptr(b), _stream(d)) == 0
    assert np.array_equal(host(a), ind) and np.array_equal(host(b), ind)


def test_empty_and_rejected_sizes(ops):
    lib = ops._lib.lib()
    ptr = dev(np.zeros(9, I64))
    out = torch.empty(0, dtype=torch.int64, device="cuda")
    assert lib.ps_ptr2ind_noalloc(_ptr(ptr), 8, 0, _ptr(out), _stream(ptr)) == 0
    # beyond the fixed array: refused before anything is launched (the pointers are never touched)
    n0 = int(lib.ps_launch_count())
    rc = lib.ps_ptr2ind_noalloc(_ptr(ptr), 8, MAX_E + 1, None, _stream(ptr))
    assert rc != 0 and "exceeds" in ops._lib.last_error() and "ps_ptr2ind" in ops._lib.last_error()
    assert int(lib.ps_launch_count()) == n0
    assert lib.ps_ptr2ind_noalloc(_ptr(ptr), 8, -1, None, _stream(ptr)) != 0


def test_device_memory_is_untouched(ops):
    """No cudaMalloc / pool growth: the free device memory is the same before and after (module loading and the
    address look-up happen in the warm-up call)."""
    rng = np.random.default_rng(5)
    ptr, ind = _rowptr(rng, 4096, 2000000)
    d = dev(ptr)
    out = torch.empty(ind.size, dtype=torch.int64, device="cuda")
    lib = ops._lib.lib()
    assert lib.ps_ptr2ind_noalloc(_ptr(d), 4096, ind.size, _ptr(out), _stream(d)) == 0
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(2):
        assert lib.ps_ptr2ind_noalloc(_ptr(d), 4096, ind.size, _ptr(out), _stream(d)) == 0
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0
    assert np.array_equal(host(out), ind)


def test_second_device_has_its_own_array(ops):
    if torch.cuda.device_count() < 2:
        pytest.skip("one GPU")
    rng = np.random.default_rng(6)
    ptr, ind = _rowptr(rng, 999, 300001, True)
    for name in ("cuda:1", "cuda:0", "cuda:1"):
        want, got = _both(ops, torch.from_numpy(ptr).to(name), ind.size)
        assert torch.equal(got, want) and np.array_equal(got.cpu().numpy(), ind)
