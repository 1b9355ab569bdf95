"""``paddle_sparse_ops`` — the module object the reference imports (paddle_sparse/__init__.py:8-17) and calls at
storage.py:202 (``ptr2ind``), storage.py:218,393 (``ind2ptr``), sample.py:37 (``sample_adj``) and __init__.py:17
(``sparse_cuda_version``), rebuilt on libpaddle_sparse_b200.so.

The reference registers these four as Paddle custom operators (PD_BUILD_OP, csrc/convert.cpp:37-43,70-76,
csrc/sample.cpp:35-40, csrc/version.cpp:35-40).  Here they are plain Python callables with the same names, the same
positional signature and the same results, and tensors cross the boundary by **DLPack**: an argument is any object
with ``__dlpack__`` (a Paddle, PyTorch or CuPy tensor, or a bare ``dltensor`` capsule); pointer, shape, dtype and
device are read from the ``DLManagedTensor`` with ctypes and handed to the C ABI (include/ps_abi.h) as raw
pointers.  This file imports no array library — not torch, not paddle, not numpy.

Outputs (the reference allocates them inside the op: convert_cpu.cpp:10,13, convert_cuda.cu:27,30,
sample_cpu.cpp:25-26,123-130):
  * with an allocator installed (``set_allocator``), every output is a tensor of the CALLER's framework:
    ``allocator(shape, "int64", device_id)`` must return an object with ``__dlpack__`` (for Paddle:
    ``lambda shape, dtype, dev: paddle.empty(shape, dtype, place=paddle.CUDAPlace(dev))``); the op fills it and
    returns it (``sample_adj`` returns ``out[:E]`` slices of blocks allocated at the upper bound);
  * without one, outputs are ``DeviceArray`` objects owning a stream-ordered block of the shim's allocator
    (ps_alloc / ps_free); they implement ``__dlpack__`` / ``__dlpack_device__`` themselves, so
    ``torch.from_dlpack(out)`` / ``paddle.utils.dlpack.from_dlpack(out.__dlpack__())`` take them zero-copy.
Kernels are enqueued on the stream ``set_stream`` names (default: the legacy default stream) and producers are
told so through ``__dlpack__(stream=...)``.  GPU only: a host tensor raises, where the reference's sample_adj raises
on a GPU one (csrc/sample.cpp:13-15).
"""
from __future__ import annotations

import ctypes
from typing import Callable, Optional, Sequence, Tuple

from . import _lib

__all__ = ["ind2ptr", "ptr2ind", "sample_adj", "sparse_cuda_version", "set_allocator", "set_stream", "set_seed",
           "DeviceArray"]

# ------------------------------------------------------------------------------------------------ DLPack (dlpack.h)
kDLCPU, kDLCUDA, kDLCUDAHost, kDLCUDAManaged = 1, 2, 3, 13
kDLInt, kDLUInt, kDLFloat, kDLBfloat, kDLBool = 0, 1, 2, 4, 6


class DLDevice(ctypes.Structure):
    _fields_ = [("device_type", ctypes.c_int32), ("device_id", ctypes.c_int32)]


class DLDataType(ctypes.Structure):
    _fields_ = [("code", ctypes.c_uint8), ("bits", ctypes.c_uint8), ("lanes", ctypes.c_uint16)]


class DLTensor(ctypes.Structure):
    _fields_ = [("data", ctypes.c_void_p), ("device", DLDevice), ("ndim", ctypes.c_int32), ("dtype", DLDataType),
                ("shape", ctypes.POINTER(ctypes.c_int64)), ("strides", ctypes.POINTER(ctypes.c_int64)),
                ("byte_offset", ctypes.c_uint64)]


class DLManagedTensor(ctypes.Structure):
    pass


_DELETER = ctypes.CFUNCTYPE(None, ctypes.POINTER(DLManagedTensor))
DLManagedTensor._fields_ = [("dl_tensor", DLTensor), ("manager_ctx", ctypes.c_void_p), ("deleter", _DELETER)]


class DLPackVersion(ctypes.Structure):
    _fields_ = [("major", ctypes.c_uint32), ("minor", ctypes.c_uint32)]


class DLManagedTensorVersioned(ctypes.Structure):
    pass


_DELETER_V = ctypes.CFUNCTYPE(None, ctypes.POINTER(DLManagedTensorVersioned))
DLManagedTensorVersioned._fields_ = [("version", DLPackVersion), ("manager_ctx", ctypes.c_void_p),
                                     ("deleter", _DELETER_V), ("flags", ctypes.c_uint64), ("dl_tensor", DLTensor)]

_NAME, _USED = b"dltensor", b"used_dltensor"
_NAME_V, _USED_V = b"dltensor_versioned", b"used_dltensor_versioned"

_py = ctypes.pythonapi
_py.PyCapsule_IsValid.restype = ctypes.c_int
_py.PyCapsule_IsValid.argtypes = [ctypes.py_object, ctypes.c_char_p]
_py.PyCapsule_GetPointer.restype = ctypes.c_void_p
_py.PyCapsule_GetPointer.argtypes = [ctypes.py_object, ctypes.c_char_p]
_py.PyCapsule_SetName.restype = ctypes.c_int
_py.PyCapsule_SetName.argtypes = [ctypes.py_object, ctypes.c_char_p]
_py.PyCapsule_GetName.restype = ctypes.c_char_p
_py.PyCapsule_GetName.argtypes = [ctypes.py_object]
_CAPSULE_DTOR = ctypes.CFUNCTYPE(None, ctypes.c_void_p)
_py.PyCapsule_New.restype = ctypes.py_object
_py.PyCapsule_New.argtypes = [ctypes.c_void_p, ctypes.c_char_p, _CAPSULE_DTOR]
# the same API on raw PyObject* (a capsule destructor runs on an object whose reference count is already zero: it
# must not be wrapped in a py_object)
_py_raw = ctypes.PyDLL(None)
_py_raw.PyCapsule_IsValid.restype = ctypes.c_int
_py_raw.PyCapsule_IsValid.argtypes = [ctypes.c_void_p, ctypes.c_char_p]
_py_raw.PyCapsule_GetPointer.restype = ctypes.c_void_p
_py_raw.PyCapsule_GetPointer.argtypes = [ctypes.c_void_p, ctypes.c_char_p]

_DTYPES = {"int64": (kDLInt, 64, 8), "int32": (kDLInt, 32, 4), "float32": (kDLFloat, 32, 4),
           "float64": (kDLFloat, 64, 8), "float16": (kDLFloat, 16, 2), "bfloat16": (kDLBfloat, 16, 2),
           "uint8": (kDLUInt, 8, 1), "bool": (kDLBool, 8, 1)}


# ------------------------------------------------------------------------------------------------ configuration
_allocator: Optional[Callable] = None
_stream: int = 0
_seed: int = 0


def set_allocator(fn: Optional[Callable]) -> None:
    """``fn(shape: tuple, dtype: str, device_id: int) -> tensor with __dlpack__`` (None: the shim's own blocks)."""
    global _allocator
    _allocator = fn


def set_stream(stream) -> None:
    """The CUDA stream kernels are enqueued on: a ``cudaStream_t`` handle as an int (0 = legacy default stream)
    or a callable returning one (e.g. ``lambda: torch.cuda.current_stream().cuda_stream``)."""
    global _stream
    _stream = stream


def set_seed(seed: int) -> None:
    """Philox key of ``sample_adj`` (the reference draws from Paddle's global generator, csrc/cpu/utils.h:22-34;
    here the draw of (seed row, draw index) is a counter-based function of this seed: SURVEY 8c)."""
    global _seed
    _seed = int(seed) & (2**64 - 1)


def _current_stream() -> int:
    s = _stream() if callable(_stream) else _stream
    return int(s or 0)


def _dlpack_stream_arg(stream: int):
    # array-API convention for CUDA consumers: 1 = legacy default stream, 2 = per-thread default, else the handle
    return 1 if stream == 0 else stream


# ------------------------------------------------------------------------------------------------ importing
class _View:
    """A borrowed device array: what one DLPack capsule describes.  `release()` runs the producer's deleter."""

    __slots__ = ("ptr", "shape", "code", "bits", "device_type", "device_id", "_capsule", "_managed", "_versioned",
                 "_keep")

    def release(self) -> None:
        m, self._managed = self._managed, None
        if m is not None and m.contents.deleter:
            m.contents.deleter(m)
        self._capsule = None
        self._keep = None

    @property
    def numel(self) -> int:
        n = 1
        for d in self.shape:
            n *= d
        return n


def _capsule_of(obj, stream: int):
    if _py.PyCapsule_IsValid(obj, _NAME) or _py.PyCapsule_IsValid(obj, _NAME_V):
        return obj
    if not hasattr(obj, "__dlpack__"):
        raise TypeError(f"paddle_sparse_ops: expected a tensor with __dlpack__ or a DLPack capsule, got {type(obj).__name__}")
    dev = obj.__dlpack_device__() if hasattr(obj, "__dlpack_device__") else (kDLCUDA, 0)
    if int(dev[0]) in (kDLCUDA, kDLCUDAManaged):
        try:
            return obj.__dlpack__(stream=_dlpack_stream_arg(stream))
        except TypeError:  # pre-array-API producers take no stream argument
            return obj.__dlpack__()
    return obj.__dlpack__()


def _import(obj, name: str, stream: int) -> _View:
    """DLManagedTensor fields of `obj` (consumes the capsule: renamed to used_dltensor, deleter owed to release())."""
    cap = _capsule_of(obj, stream)
    v = _View()
    v._keep = obj
    if _py.PyCapsule_IsValid(cap, _NAME_V):
        raw = _py.PyCapsule_GetPointer(cap, _NAME_V)
        m = ctypes.cast(raw, ctypes.POINTER(DLManagedTensorVersioned))
        _py.PyCapsule_SetName(cap, _USED_V)
    elif _py.PyCapsule_IsValid(cap, _NAME):
        raw = _py.PyCapsule_GetPointer(cap, _NAME)
        m = ctypes.cast(raw, ctypes.POINTER(DLManagedTensor))
        _py.PyCapsule_SetName(cap, _USED)
    else:
        raise TypeError(f"paddle_sparse_ops: `{name}` is not an unconsumed DLPack capsule")
    v._capsule, v._managed = cap, m
    t = m.contents.dl_tensor
    v.shape = tuple(int(t.shape[i]) for i in range(t.ndim))
    v.code, v.bits = int(t.dtype.code), int(t.dtype.bits)
    v.device_type, v.device_id = int(t.device.device_type), int(t.device.device_id)
    v.ptr = (t.data or 0) + int(t.byte_offset)
    if t.dtype.lanes != 1:
        v.release()
        raise TypeError(f"paddle_sparse_ops: `{name}` has a vector dtype")
    if t.strides and v.numel > 0:  # NULL strides = compact row-major; an empty array has no layout to check
        expect = 1
        for i in range(t.ndim - 1, -1, -1):
            if v.shape[i] != 1 and int(t.strides[i]) != expect:
                v.release()
                raise RuntimeError(f"paddle_sparse_ops: `{name}` must be contiguous")
            expect *= v.shape[i]
    return v


def _check_index(v: _View, name: str, ndim: int = 1) -> None:
    # CHECK_CUDA (csrc/cuda/utils.cuh:5-6) + data<int64_t>() (convert_cuda.cu:31-32)
    if v.device_type not in (kDLCUDA, kDLCUDAManaged):
        raise RuntimeError(f"paddle_sparse_ops is GPU-only (no CPU fallback): `{name}` must be a CUDA tensor")
    if (v.code, v.bits) != (kDLInt, 64):
        raise TypeError(f"`{name}` must be int64")
    if len(v.shape) != ndim:
        raise RuntimeError(f"`{name}` must be {ndim}-D")


# ------------------------------------------------------------------------------------------------ exporting
_live = {}  # address of a handed-out DLManagedTensor -> everything it points at (kept until its deleter runs)


@_DELETER
def _release_export(mptr):
    _live.pop(ctypes.addressof(mptr.contents), None)  # drops the record's reference to the block


@_CAPSULE_DTOR
def _capsule_dropped(capsule_ptr):
    # a capsule nobody consumed still owns its DLManagedTensor (name unchanged): run the deleter
    if _py_raw.PyCapsule_IsValid(capsule_ptr, _NAME):
        m = ctypes.cast(_py_raw.PyCapsule_GetPointer(capsule_ptr, _NAME), ctypes.POINTER(DLManagedTensor))
        if m.contents.deleter:
            m.contents.deleter(m)


class _Block:
    """One allocation of the shim's allocator (or a host buffer).  Released - ps_free, ordered on the stream it was
    produced on - when the last DeviceArray and the last exported DLPack capsule that refer to it are gone."""

    def __init__(self, ptr: int, device_type: int, stream: int, host_buffer=None):
        self.ptr, self.device_type, self.stream, self.host_buffer, self.freed = int(ptr), device_type, stream, host_buffer, False

    def __del__(self):
        if not self.freed:
            self.freed = True
            if self.device_type == kDLCUDA and self.host_buffer is None and self.ptr:
                try:
                    _lib.lib().ps_free(ctypes.c_void_p(self.ptr), ctypes.c_void_p(self.stream))
                except Exception:  # interpreter shutdown
                    pass


class DeviceArray:
    """A view of a block owned by the shim, exported zero-copy through the DLPack producer protocol
    (``__dlpack__`` / ``__dlpack_device__``); 1-D contiguous slices share the block."""

    def __init__(self, ptr: int, shape: Sequence[int], dtype: str, device_type: int, device_id: int, stream: int = 0,
                 base: "Optional[DeviceArray]" = None, host_buffer=None):
        self.ptr, self.shape, self.dtype = int(ptr), tuple(int(s) for s in shape), dtype
        self.device_type, self.device_id, self.stream = device_type, device_id, stream
        self._block = base._block if base is not None else _Block(ptr, device_type, stream, host_buffer)

    @property
    def nbytes(self) -> int:
        n = _DTYPES[self.dtype][2]
        for d in self.shape:
            n *= d
        return n

    def __len__(self) -> int:
        return self.shape[0]

    def __getitem__(self, key):
        if not (isinstance(key, slice) and len(self.shape) == 1 and key.step in (None, 1)):
            raise IndexError("DeviceArray supports contiguous 1-D slices only")
        lo, hi, _ = key.indices(self.shape[0])
        hi = max(hi, lo)
        return DeviceArray(self.ptr + lo * _DTYPES[self.dtype][2], (hi - lo,), self.dtype, self.device_type,
                           self.device_id, self.stream, base=self)

    def tolist(self):
        """Host copy as a Python list (synchronises; for tests and small results)."""
        if self.dtype != "int64":
            raise TypeError("tolist() is implemented for int64 arrays")
        n = self.nbytes // 8
        buf = (ctypes.c_int64 * max(n, 1))()
        if n:
            if self.device_type == kDLCUDA:
                _lib.call("ps_memcpy", buf, ctypes.c_void_p(self.ptr), self.nbytes, 2, ctypes.c_void_p(self.stream))
            else:
                ctypes.memmove(buf, self.ptr, self.nbytes)
        return list(buf[:n])

    def item(self):
        vals = self.tolist()
        if len(vals) != 1:
            raise ValueError("only one-element arrays convert to a Python scalar")
        return vals[0]

    # -- the DLPack producer protocol
    def __dlpack_device__(self) -> Tuple[int, int]:
        return (self.device_type, self.device_id)

    def __dlpack__(self, stream=None, **_):
        # the data was produced on self.stream; a consumer on another stream is ordered behind it here
        if self.device_type == kDLCUDA and stream is not None and stream != -1:
            want = 0 if stream in (1, 2) else int(stream)
            if want != self.stream:
                _lib.call("ps_stream_sync", ctypes.c_void_p(self.stream))
        code, bits, _size = _DTYPES[self.dtype]
        shape = (ctypes.c_int64 * max(len(self.shape), 1))(*self.shape)
        m = DLManagedTensor()
        m.dl_tensor.data = ctypes.c_void_p(self.ptr)
        m.dl_tensor.device = DLDevice(self.device_type, self.device_id)
        m.dl_tensor.ndim = len(self.shape)
        m.dl_tensor.dtype = DLDataType(code, bits, 1)
        m.dl_tensor.shape = ctypes.cast(shape, ctypes.POINTER(ctypes.c_int64))
        m.dl_tensor.strides = None
        m.dl_tensor.byte_offset = 0
        m.manager_ctx = None
        m.deleter = _release_export
        _live[ctypes.addressof(m)] = {"m": m, "shape": shape, "block": self._block}
        return _py.PyCapsule_New(ctypes.addressof(m), _NAME, _capsule_dropped)


class _Out:
    """One output being produced: a pointer to write through, and what to hand back."""

    __slots__ = ("ptr", "result", "_view")


def _new_output(shape: Tuple[int, ...], device_id: int, stream: int) -> _Out:
    o = _Out()
    o._view = None
    if _allocator is not None:
        t = _allocator(tuple(shape), "int64", device_id)
        v = _import(t, "allocator result", stream)
        try:
            _check_index(v, "allocator result", len(shape))
            if v.shape != tuple(shape) or v.device_id != device_id:
                raise RuntimeError(f"allocator returned shape {v.shape} on device {v.device_id}, "
                                   f"asked for {tuple(shape)} on device {device_id}")
        except Exception:
            v.release()
            raise
        o.ptr, o.result, o._view = v.ptr, t, v
        return o
    n = 8
    for d in shape:
        n *= d
    p = ctypes.c_void_p()
    _lib.call("ps_alloc", n, ctypes.byref(p), ctypes.c_void_p(stream))
    o.ptr = p.value or 0
    o.result = DeviceArray(o.ptr, shape, "int64", kDLCUDA, device_id, stream)
    return o


def _finish(outs, views) -> None:
    for o in outs:
        if o._view is not None:
            o._view.release()
    for v in views:
        v.release()


class _on_device:
    """Kernels run on the device the inputs live on (convert_cuda.cu:24: the tensor's place)."""

    def __init__(self, dev: int):
        self.dev, self.prev = dev, -1

    def __enter__(self):
        lib = _lib.lib()
        self.prev = int(lib.ps_current_device())
        if self.prev != self.dev:
            _lib.call("ps_set_device", self.dev)

    def __exit__(self, *exc):
        if self.prev not in (-1, self.dev):
            _lib.call("ps_set_device", self.prev)


def _vp(p: int):
    return ctypes.c_void_p(p)


# ------------------------------------------------------------------------------------------------ the four operators
def ind2ptr(ind, M: int):
    """csrc/convert.cpp:13-23: ``out[r] = #{e : ind[e] < r}``, r in [0, M]; ``ind`` sorted int64 [nnz] -> int64 [M+1]."""
    stream = _current_stream()
    v = _import(ind, "ind", stream)
    outs = []
    try:
        _check_index(v, "ind")
        with _on_device(v.device_id):
            out = _new_output((int(M) + 1,), v.device_id, stream)
            outs.append(out)
            _lib.call("ps_ind2ptr", _vp(v.ptr), v.numel, int(M), _vp(out.ptr), _vp(stream))
        return out.result
    finally:
        _finish(outs, [v])


def ptr2ind(ptr, E: int):
    """csrc/convert.cpp:46-56: ``out[e] = r`` for e in [ptr[r], ptr[r+1]); int64 [M+1] -> int64 [E]."""
    stream = _current_stream()
    v = _import(ptr, "ptr", stream)
    outs = []
    try:
        _check_index(v, "ptr")
        with _on_device(v.device_id):
            out = _new_output((int(E),), v.device_id, stream)
            outs.append(out)
            _lib.call("ps_ptr2ind", _vp(v.ptr), v.numel - 1, int(E), _vp(out.ptr), _vp(stream))
        return out.result
    finally:
        _finish(outs, [v])


def sample_adj(rowptr, col, idx, num_neighbors: int, replace: bool):
    """csrc/sample.cpp:8-24 -> ``[out_rowptr [S+1], out_col [E], out_n_id [N'], out_e_id [E]]`` (the reference runs
    this on the CPU only and throws on GPU tensors, :13-15; this is the GPU implementation, seeded by set_seed)."""
    stream = _current_stream()
    views = []
    outs = []
    try:
        for t, name in ((rowptr, "rowptr"), (col, "col"), (idx, "idx")):
            views.append(_import(t, name, stream))
            _check_index(views[-1], name)  # CHECK_INPUT(idx.dim() == 1), sample_cpu.cpp:19
        vr, vc, vi = views
        if not (vr.device_id == vc.device_id == vi.device_id):
            raise RuntimeError("sample_adj: rowptr, col and idx must live on the same device")
        S = vi.numel
        with _on_device(vr.device_id):
            hb = (ctypes.c_int64 * 1)()
            _lib.call("ps_sample_adj_bound", _vp(vr.ptr), _vp(vi.ptr), S, int(num_neighbors), int(bool(replace)), hb,
                      _vp(stream))
            bound = int(hb[0])
            o_rowptr = _new_output((S + 1,), vr.device_id, stream)
            o_col = _new_output((bound,), vr.device_id, stream)
            o_nid = _new_output((S + bound,), vr.device_id, stream)
            o_eid = _new_output((bound,), vr.device_id, stream)
            outs += [o_rowptr, o_col, o_nid, o_eid]
            sizes = (ctypes.c_int64 * 2)()
            _lib.call("ps_sample_adj", _vp(vr.ptr), _vp(vc.ptr), vr.numel - 1, vc.numel, _vp(vi.ptr), S,
                      int(num_neighbors), int(bool(replace)), ctypes.c_uint64(_seed), bound, _vp(o_rowptr.ptr),
                      _vp(o_col.ptr), _vp(o_nid.ptr), _vp(o_eid.ptr), sizes, _vp(stream))
        E, Nn = int(sizes[0]), int(sizes[1])
        return [o_rowptr.result, o_col.result[:E], o_nid.result[:Nn], o_eid.result[:E]]
    finally:
        _finish(outs, views)


def sparse_cuda_version():
    """csrc/version.cpp:8-33: int64 [1] on the HOST holding the CUDA_VERSION the library was built with
    (``paddle_sparse_ops.sparse_cuda_version().item()``, paddle_sparse/__init__.py:17)."""
    buf = (ctypes.c_int64 * 1)(int(_lib.lib().ps_cuda_version()))
    return DeviceArray(ctypes.addressof(buf), (1,), "int64", kDLCPU, 0, host_buffer=buf)
