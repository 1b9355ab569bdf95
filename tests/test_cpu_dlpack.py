"""CPU-side checks of the DLPack boundary (paddle_sparse_b200/paddle_sparse_ops.py): the capsule parser reads
pointer / shape / dtype / device of foreign producers, refuses host tensors like CHECK_CUDA does, the module imports
no array library, and `sparse_cuda_version()` behaves like the reference's op (paddle_sparse/__init__.py:17)."""
import gc
import os
import re

import numpy as np
import pytest
import torch

from paddle_sparse_b200 import paddle_sparse_ops as pso

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_module_exposes_the_reference_ops_and_no_array_library():
    for name in ("ind2ptr", "ptr2ind", "sample_adj", "sparse_cuda_version"):  # csrc/convert.cpp:37,70; sample.cpp:35; version.cpp:35
        assert callable(getattr(pso, name))
    src = open(os.path.join(ROOT, "paddle_sparse_b200", "paddle_sparse_ops.py")).read()
    assert not re.search(r"^\s*(import|from)\s+(torch|numpy|paddle|cupy)\b", src, flags=re.M)


def test_sparse_cuda_version_is_a_host_int64_scalar():
    v = pso.sparse_cuda_version()
    assert v.shape == (1,) and v.item() >= 12000 and v.__dlpack_device__() == (pso.kDLCPU, 0)
    a = np.from_dlpack(v)  # a foreign consumer takes it zero-copy
    assert a.dtype == np.int64 and a.tolist() == [v.item()]
    t = torch.from_dlpack(v)
    assert t.dtype == torch.int64 and t.tolist() == [v.item()]
    del a, t, v
    gc.collect()
    assert not pso._live  # every exported DLManagedTensor was released by its consumer


def test_capsule_parser_reads_foreign_producers():
    x = np.arange(7, dtype=np.int64)
    v = pso._import(x, "x", 0)
    assert (v.shape, v.code, v.bits, v.device_type) == ((7,), pso.kDLInt, 64, pso.kDLCPU) and v.ptr == x.ctypes.data
    v.release()
    t = torch.arange(9, dtype=torch.int64)
    cap = torch.utils.dlpack.to_dlpack(t[2:])  # a bare capsule with a byte offset / shifted pointer
    v = pso._import(cap, "t", 0)
    assert v.shape == (7,) and v.ptr == t.data_ptr() + 16
    v.release()
    with pytest.raises(TypeError):  # a capsule is single-use (renamed used_dltensor)
        pso._import(cap, "t", 0)
    with pytest.raises(RuntimeError, match="contiguous"):
        pso._import(torch.arange(10, dtype=torch.int64)[::2], "strided", 0)
    with pytest.raises(TypeError, match="__dlpack__"):
        pso._import([1, 2, 3], "list", 0)


def test_host_tensors_are_refused():
    # the reference dispatches on ind.is_gpu() (convert.cpp:14); this library has no CPU path
    with pytest.raises(RuntimeError, match="GPU-only"):
        pso.ind2ptr(np.arange(4, dtype=np.int64), 4)
    with pytest.raises(RuntimeError, match="GPU-only"):
        pso.ptr2ind(torch.arange(4, dtype=torch.int64), 3)
    one = torch.zeros(2, dtype=torch.int64)
    with pytest.raises(RuntimeError, match="GPU-only"):
        pso.sample_adj(one, one, one, 2, False)


def test_device_array_slices_share_one_block():
    import sys
    buf = (pso.ctypes.c_int64 * 6)(*range(6))
    a = pso.DeviceArray(pso.ctypes.addressof(buf), (6,), "int64", pso.kDLCPU, 0, host_buffer=buf)
    block = a._block
    b = a[2:5]
    assert b.tolist() == [2, 3, 4] and len(b) == 3 and b._block is block
    t = torch.from_dlpack(b)
    assert t.tolist() == [2, 3, 4] and len(pso._live) == 1
    del a, b
    gc.collect()
    assert not block.freed and t.tolist() == [2, 3, 4]  # the consumer's capsule keeps the block alive
    del t
    gc.collect()
    assert not pso._live and sys.getrefcount(block) == 2  # only `block` itself (+ the call's temporary) is left
    with pytest.raises(IndexError):
        pso.DeviceArray(pso.ctypes.addressof(buf), (6,), "int64", pso.kDLCPU, 0, host_buffer=buf)[::2]
