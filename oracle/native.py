"""TEST INFRASTRUCTURE ONLY (oracle).

ctypes access to the two native oracle libraries:

* ``ref``  = oracle/_ref/libps_ref_cpu.so — the reference's own
  ``csrc/cpu/convert_cpu.cpp`` + ``csrc/cpu/sample_cpu.cpp`` compiled UNMODIFIED
  (recipe: oracle/Makefile, target ``ref``).  Present only where it was built
  (this container, and GPU boxes that received the snapshot).
* ``port`` = oracle/_build/libps_oracle_port.so — plain-C restatement
  (oracle/oracle_port.c).  Always buildable (needs only gcc).

``ind2ptr`` / ``ptr2ind`` prefer ``ref`` and fall back to ``port``; the tests pin
one against the other and both against the reference's golden vectors.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_REF_PATH = os.path.join(_HERE, "_ref", "libps_ref_cpu.so")
_PORT_PATH = os.path.join(_HERE, "_build", "libps_oracle_port.so")
_REF_CUDA_PATH = os.path.join(_HERE, "_ref", "libps_ref_cuda.so")

_i64p = ctypes.POINTER(ctypes.c_int64)
_ref = None
_port = None


def _ptr(a: np.ndarray):
    assert a.dtype == np.int64 and a.flags.c_contiguous
    return a.ctypes.data_as(_i64p)


def build(ref: bool = True) -> None:
    """Compile the oracle libraries (make port; make ref when /root/reference exists)."""
    subprocess.check_call(["make", "-s", "-C", _HERE, "port"])
    if ref and os.path.isdir("/root/reference/csrc"):
        subprocess.check_call(["make", "-s", "-C", _HERE, "ref"])
        subprocess.check_call(["make", "-s", "-C", _HERE, "refcuda"])


def have_ref() -> bool:
    return os.path.exists(_REF_PATH)


def have_ref_cuda() -> bool:
    return os.path.exists(_REF_CUDA_PATH)


_ref_cuda = None


def ref_cuda_lib():
    """The reference's own CUDA ind2ptr / ptr2ind (csrc/cuda/convert_cuda.cu, unmodified, sm_100a) behind
    oracle/ref_cuda_shim.cu.  Arguments are raw device pointers (ints) and a stream handle."""
    global _ref_cuda
    if _ref_cuda is None:
        if not have_ref_cuda():
            raise FileNotFoundError(
                f"{_REF_CUDA_PATH} not built (run `make -C oracle refcuda` where /root/reference exists)"
            )
        lib = ctypes.CDLL(_REF_CUDA_PATH)
        lib.refcuda_last_error.restype = ctypes.c_char_p
        for f in (lib.refcuda_ind2ptr, lib.refcuda_ptr2ind):
            f.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p]
            f.restype = ctypes.c_int
        _ref_cuda = lib
    return _ref_cuda


def port_lib():
    global _port
    if _port is None:
        if not os.path.exists(_PORT_PATH):
            build(ref=False)
        lib = ctypes.CDLL(_PORT_PATH)
        lib.port_ind2ptr.argtypes = [_i64p, ctypes.c_int64, ctypes.c_int64, _i64p]
        lib.port_ind2ptr.restype = None
        lib.port_ptr2ind.argtypes = [_i64p, ctypes.c_int64, ctypes.c_int64, _i64p]
        lib.port_ptr2ind.restype = None
        lib.port_sample_adj_bound.argtypes = [_i64p, _i64p, ctypes.c_int64, ctypes.c_int64]
        lib.port_sample_adj_bound.restype = ctypes.c_int64
        lib.port_sample_adj.argtypes = [
            _i64p, _i64p, _i64p, ctypes.c_int64, ctypes.c_int64, ctypes.c_int,
            ctypes.c_uint64, _i64p, _i64p, _i64p, _i64p, _i64p,
        ]
        lib.port_sample_adj.restype = None
        lib.port_philox_u64.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint32]
        lib.port_philox_u64.restype = ctypes.c_uint64
        _port = lib
    return _port


def ref_lib():
    global _ref
    if _ref is None:
        if not have_ref():
            raise FileNotFoundError(
                f"{_REF_PATH} not built (run `make -C oracle ref` where /root/reference exists)"
            )
        lib = ctypes.CDLL(_REF_PATH)
        lib.ref_last_error.restype = ctypes.c_char_p
        lib.ref_ind2ptr.argtypes = [_i64p, ctypes.c_int64, ctypes.c_int64, _i64p]
        lib.ref_ind2ptr.restype = ctypes.c_int
        lib.ref_ptr2ind.argtypes = [_i64p, ctypes.c_int64, ctypes.c_int64, _i64p]
        lib.ref_ptr2ind.restype = ctypes.c_int
        lib.ref_sample_adj.argtypes = [
            _i64p, ctypes.c_int64, _i64p, ctypes.c_int64, _i64p, ctypes.c_int64,
            ctypes.c_int64, ctypes.c_int, ctypes.c_uint64, _i64p,
        ]
        lib.ref_sample_adj.restype = ctypes.c_void_p
        lib.ref_sample_adj_fetch.argtypes = [ctypes.c_void_p, _i64p, _i64p, _i64p, _i64p]
        lib.ref_sample_adj_fetch.restype = None
        lib.ref_sample_adj_free.argtypes = [ctypes.c_void_p]
        lib.ref_sample_adj_free.restype = None
        _ref = lib
    return _ref


def _c64(a) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(a, dtype=np.int64))


# --------------------------------------------------------------------------- ind2ptr / ptr2ind
def ind2ptr(ind, M: int, which: str = "auto") -> np.ndarray:
    """csrc/convert.cpp:13-23 -> csrc/cpu/convert_cpu.cpp:6-30."""
    ind = _c64(ind)
    out = np.empty(M + 1, dtype=np.int64)
    if which == "ref" or (which == "auto" and have_ref()):
        lib = ref_lib()
        if lib.ref_ind2ptr(_ptr(ind), ind.size, M, _ptr(out)) != 0:
            raise RuntimeError(lib.ref_last_error().decode())
    else:
        port_lib().port_ind2ptr(_ptr(ind), ind.size, M, _ptr(out))
    return out


def ptr2ind(ptr, E: int, which: str = "auto") -> np.ndarray:
    """csrc/convert.cpp:46-56 -> csrc/cpu/convert_cpu.cpp:32-48."""
    ptr = _c64(ptr)
    out = np.empty(E, dtype=np.int64)
    if which == "ref" or (which == "auto" and have_ref()):
        lib = ref_lib()
        if lib.ref_ptr2ind(_ptr(ptr), ptr.size - 1, E, _ptr(out)) != 0:
            raise RuntimeError(lib.ref_last_error().decode())
    else:
        port_lib().port_ptr2ind(_ptr(ptr), ptr.size - 1, E, _ptr(out))
    return out


# --------------------------------------------------------------------------- sample_adj
def sample_adj_ref(rowptr, col, idx, num_neighbors: int, replace: bool, seed: int = 0):
    """The reference's own sample_adj_cpu (csrc/cpu/sample_cpu.cpp:10-148), RNG injected
    per oracle/philox.h.  Returns (out_rowptr, out_col, n_id, e_id)."""
    rowptr, col, idx = _c64(rowptr), _c64(col), _c64(idx)
    lib = ref_lib()
    sizes = np.zeros(2, dtype=np.int64)
    h = lib.ref_sample_adj(_ptr(rowptr), rowptr.size - 1, _ptr(col), col.size, _ptr(idx),
                           idx.size, int(num_neighbors), int(bool(replace)),
                           ctypes.c_uint64(seed & (2**64 - 1)), _ptr(sizes))
    if not h:
        raise RuntimeError(lib.ref_last_error().decode())
    E, Nn = int(sizes[0]), int(sizes[1])
    out_rowptr = np.empty(idx.size + 1, dtype=np.int64)
    out_col = np.empty(E, dtype=np.int64)
    n_id = np.empty(Nn, dtype=np.int64)
    e_id = np.empty(E, dtype=np.int64)
    lib.ref_sample_adj_fetch(h, _ptr(out_rowptr), _ptr(out_col), _ptr(n_id), _ptr(e_id))
    lib.ref_sample_adj_free(h)
    return out_rowptr, out_col, n_id, e_id


def sample_adj_port(rowptr, col, idx, num_neighbors: int, replace: bool, seed: int = 0):
    """Plain-C restatement (oracle/oracle_port.c: port_sample_adj); fixes the n_id
    order to insertion order and the in-row tie order to (local id, e_id)."""
    rowptr, col, idx = _c64(rowptr), _c64(col), _c64(idx)
    lib = port_lib()
    bound = int(lib.port_sample_adj_bound(_ptr(rowptr), _ptr(idx), idx.size, int(num_neighbors)))
    out_rowptr = np.empty(idx.size + 1, dtype=np.int64)
    out_col = np.empty(max(bound, 1), dtype=np.int64)
    e_id = np.empty(max(bound, 1), dtype=np.int64)
    n_id = np.empty(idx.size + max(bound, 1), dtype=np.int64)
    sizes = np.zeros(2, dtype=np.int64)
    lib.port_sample_adj(_ptr(rowptr), _ptr(col), _ptr(idx), idx.size, int(num_neighbors),
                        int(bool(replace)), ctypes.c_uint64(seed & (2**64 - 1)),
                        _ptr(out_rowptr), _ptr(out_col), _ptr(n_id), _ptr(e_id), _ptr(sizes))
    E, Nn = int(sizes[0]), int(sizes[1])
    return out_rowptr, out_col[:E].copy(), n_id[:Nn].copy(), e_id[:E].copy()


def philox_u64(seed: int, pos: int, draw: int) -> int:
    return int(port_lib().port_philox_u64(ctypes.c_uint64(seed), ctypes.c_uint64(pos),
                                          ctypes.c_uint32(draw)))


def canonical_sample(out_rowptr, out_col, n_id, e_id):
    """Canonical form used to compare two sample_adj results:
    per seed row, the sorted list of (global col id, e_id) pairs."""
    rows = []
    for i in range(len(out_rowptr) - 1):
        lo, hi = int(out_rowptr[i]), int(out_rowptr[i + 1])
        rows.append(sorted(zip(n_id[out_col[lo:hi]].tolist(), e_id[lo:hi].tolist())))
    return rows
