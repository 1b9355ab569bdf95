"""The drop-in boundary as the reference binds it: `paddle_sparse_ops.{ind2ptr, ptr2ind, sample_adj}` fed through
DLPack by foreign producers (torch tensors as producers and as bare capsules, and a hand-made producer over a raw
device allocation that involves no array library), results taken back through DLPack, compared with the reference's
own CPU code compiled unmodified (oracle/_ref) and its golden vectors (test/test_storage.py:20-25,
test/test_sample.py:17-37)."""
import ctypes
import gc

import numpy as np
import pytest
import torch

from oracle import native
from paddle_sparse_b200 import _lib
from paddle_sparse_b200 import paddle_sparse_ops as pso
from tests._util import I64, dev
from tests.golden import reference_vectors as G

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _defaults(cuda):
    pso.set_allocator(None)
    pso.set_stream(lambda: torch.cuda.current_stream().cuda_stream)
    pso.set_seed(0)
    yield
    pso.set_allocator(None)
    pso.set_stream(0)


def back(x):
    return torch.from_dlpack(x).cpu().numpy()


def test_golden_through_capsules():
    g = G.IND2PTR
    row = dev(np.array(g["row"], I64))
    ptr = pso.ind2ptr(torch.utils.dlpack.to_dlpack(row), g["M"])  # a bare capsule in
    assert isinstance(ptr, pso.DeviceArray) and ptr.tolist() == g["rowptr"]
    ind = pso.ptr2ind(ptr, len(g["row"]))  # our own producer straight back in
    assert back(ind).tolist() == g["row"]


def test_random_against_the_compiled_reference():
    rng = np.random.default_rng(5)
    for M, nnz in ((1, 0), (7, 3), (1000, 50_000), (300_000, 2_000_000)):
        row = np.sort(rng.integers(0, M, nnz)).astype(I64)
        want = native.ind2ptr(row, M)
        t = dev(row)
        got = pso.ind2ptr(t, M)  # a __dlpack__ producer in
        assert np.array_equal(back(got), want)
        assert np.array_equal(back(pso.ptr2ind(got, nnz)), row)


class RawProducer:
    """A DLPack producer over a raw device block (ps_alloc + ps_memcpy): no torch, no cupy, its own struct layout."""

    class _DLTensor(ctypes.Structure):
        _fields_ = [("data", ctypes.c_void_p), ("device_type", ctypes.c_int32), ("device_id", ctypes.c_int32),
                    ("ndim", ctypes.c_int32), ("code", ctypes.c_uint8), ("bits", ctypes.c_uint8),
                    ("lanes", ctypes.c_uint16), ("shape", ctypes.POINTER(ctypes.c_int64)),
                    ("strides", ctypes.POINTER(ctypes.c_int64)), ("byte_offset", ctypes.c_uint64)]

    class _Managed(ctypes.Structure):
        pass

    def __init__(self, values):
        self.n = len(values)
        host = (ctypes.c_int64 * max(self.n, 1))(*values)
        p = ctypes.c_void_p()
        _lib.call("ps_alloc", 8 * self.n, ctypes.byref(p), None)
        if self.n:
            _lib.call("ps_memcpy", p, host, 8 * self.n, 1, None)
        self.ptr = p.value
        self.deleted = 0
        self._keep = []

    def __dlpack_device__(self):
        return (2, int(_lib.lib().ps_current_device()))

    def __dlpack__(self, stream=None):
        M = RawProducer._Managed
        if not hasattr(M, "_fields_"):
            M._DEL = ctypes.CFUNCTYPE(None, ctypes.POINTER(M))
            M._fields_ = [("t", RawProducer._DLTensor), ("ctx", ctypes.c_void_p), ("deleter", M._DEL)]
        shape = (ctypes.c_int64 * 1)(self.n)

        def deleter(_m, self=self):
            self.deleted += 1
        cb = M._DEL(deleter)
        m = M()
        m.t.data, m.t.device_type, m.t.device_id, m.t.ndim = self.ptr, 2, self.__dlpack_device__()[1], 1
        m.t.code, m.t.bits, m.t.lanes = 0, 64, 1
        m.t.shape = ctypes.cast(shape, ctypes.POINTER(ctypes.c_int64))
        m.t.strides = None
        m.t.byte_offset = 0
        m.deleter = cb
        self._keep.append((m, shape, cb))
        new = ctypes.pythonapi.PyCapsule_New
        new.restype, new.argtypes = ctypes.py_object, [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_void_p]
        return new(ctypes.addressof(m), b"dltensor", None)

    def free(self):
        _lib.call("ps_free", ctypes.c_void_p(self.ptr), None)


def test_foreign_producer_without_any_array_library():
    pso.set_stream(0)
    rng = np.random.default_rng(11)
    M, nnz = 5000, 123_457
    row = np.sort(rng.integers(0, M, nnz)).astype(I64)
    prod = RawProducer(row.tolist())
    ptr = pso.ind2ptr(prod, M)
    assert prod.deleted == 1  # the borrowed view was handed back to its producer
    assert ptr.tolist() == native.ind2ptr(row, M).tolist()
    assert pso.ptr2ind(ptr, nnz).tolist() == row.tolist()
    prod.free()


def test_allocator_returns_the_callers_own_tensors():
    made = []

    def alloc(shape, dtype, device_id):
        t = torch.empty(shape, dtype=getattr(torch, dtype), device=torch.device("cuda", device_id))
        made.append(t)
        return t
    pso.set_allocator(alloc)
    g = G.IND2PTR
    out = pso.ind2ptr(dev(np.array(g["row"], I64)), g["M"])
    assert out is made[0] and out.tolist() == g["rowptr"]
    s = G.SAMPLE_ADJ
    rowptr = dev(native.ind2ptr(np.array(s["row"], I64), 6))
    res = pso.sample_adj(rowptr, dev(np.array(s["col"], I64)), dev(np.array(s["subset"], I64)), -1, False)
    assert all(isinstance(t, torch.Tensor) for t in res)
    out_rowptr, out_col, n_id, e_id = res  # unpacked as at paddle_sparse/sample.py:37
    assert n_id.tolist() == s["all_n_id"] and out_col.tolist() == s["all_col"] and e_id.tolist() == s["all_val"]
    assert native.ptr2ind(out_rowptr.cpu().numpy(), out_col.numel()).tolist() == s["all_row"]

    pso.set_allocator(lambda shape, dtype, dev_: torch.empty(3, dtype=torch.int64, device="cuda"))
    with pytest.raises(RuntimeError, match="allocator returned"):
        pso.ind2ptr(dev(np.array(g["row"], I64)), g["M"])


def test_sample_adj_matches_the_operator_layer(ops):
    rng = np.random.default_rng(3)
    M = 20_000
    deg = rng.poisson(12, M)
    rowptr = np.concatenate([[0], np.cumsum(deg)]).astype(I64)
    col = np.concatenate([np.sort(rng.choice(M, d, replace=False)) for d in deg]).astype(I64)
    idx = rng.choice(M, 512, replace=False).astype(I64)
    pso.set_seed(42)
    for k, replace in ((5, False), (5, True), (-1, False)):
        want = [t.cpu().numpy() for t in ops.sample_adj(dev(rowptr), dev(col), dev(idx), k, replace, 42)]
        port = native.sample_adj_port(rowptr, col, idx, k, replace, 42)
        got = pso.sample_adj(dev(rowptr), dev(col), dev(idx), k, replace)
        for a, b, c in zip(got, want, port):
            assert np.array_equal(back(a), b) and np.array_equal(b, c)


def test_outputs_release_their_blocks():
    t = dev(np.arange(1000, dtype=I64))
    out = pso.ind2ptr(t, 1000)
    block = out._block
    view = torch.from_dlpack(out)
    del out
    gc.collect()
    assert not block.freed and view[-1].item() == 1000  # the consumer's capsule keeps the block alive
    del view
    gc.collect()
    assert not pso._live and not block.freed
    del block
    gc.collect()  # the last reference: _Block.__del__ hands the memory back (ps_free, stream-ordered)
