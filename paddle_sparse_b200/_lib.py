"""ctypes doorway onto libpaddle_sparse_b200.so (the C ABI of include/ps_abi.h).

This is the only place the shared library is opened.  There is no CPU fallback: when
the library is missing, ``lib()`` raises with the build command.  Every entry point
returns an int status; ``call`` turns a non-zero status into ``RuntimeError`` carrying
``ps_last_error()`` (the role PD_CHECK / PD_THROW play in the reference,
csrc/cpu/utils.h:6-9).
"""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# PS_B200_LIB: a differently built copy of the same library (debug / timeline builds of the tools); never a fallback
LIB_PATH = os.environ.get("PS_B200_LIB") or os.path.join(_HERE, "libpaddle_sparse_b200.so")

c_int = ctypes.c_int
c_i64 = ctypes.c_int64
c_u64 = ctypes.c_uint64
c_ptr = ctypes.c_void_p

# name -> argtypes, in the order of include/ps_abi.h.  restype is int unless listed in _RESTYPES.
SIGNATURES = {
    "ps_cuda_version": [],
    "ps_last_error": [],
    "ps_launch_count": [],
    "ps_prof_enable": [c_int],
    "ps_prof_kinds": [],
    "ps_prof_name": [c_int],
    "ps_prof_read": [c_int, c_ptr, c_ptr],
    "ps_alloc": [c_i64, c_ptr, c_ptr],
    "ps_free": [c_ptr, c_ptr],
    "ps_memcpy": [c_ptr, c_ptr, c_i64, c_int, c_ptr],
    "ps_stream_sync": [c_ptr],
    "ps_current_device": [],
    "ps_set_device": [c_int],
    "ps_selftest_atom_order": [c_ptr, c_ptr],
    "ps_rank_mode": [],
    "ps_ind2ptr": [c_ptr, c_i64, c_i64, c_ptr, c_ptr],
    "ps_ptr2ind": [c_ptr, c_i64, c_i64, c_ptr, c_ptr],
    "ps_ptr2ind_noalloc": [c_ptr, c_i64, c_i64, c_ptr, c_ptr],
    "ps_ptr2count": [c_ptr, c_i64, c_ptr, c_ptr],
    "ps_count2ptr": [c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_ptr],
    "ps_histogram": [c_ptr, c_i64, c_i64, c_ptr, c_ptr],
    "ps_index_max": [c_ptr, c_i64, c_ptr, c_i64, c_ptr, c_ptr],
    "ps_coo_order": [c_ptr, c_ptr, c_i64, c_ptr, c_ptr],
    "ps_sort_coo": [c_ptr, c_ptr, c_i64, c_i64, c_i64, c_ptr, c_int, c_i64, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr],
    "ps_coalesce_sorted": [c_ptr, c_ptr, c_i64, c_i64, c_i64, c_ptr, c_int, c_i64, c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr],
    "ps_coo_inspect_bytes": [],
    "ps_coo_inspect": [c_ptr, c_ptr, c_i64, c_i64, c_i64, c_ptr, c_ptr, c_ptr],
    "ps_coalesce": [c_ptr, c_ptr, c_i64, c_i64, c_i64, c_ptr, c_int, c_i64, c_int, c_ptr, c_ptr, c_ptr, c_ptr,
                    c_ptr, c_ptr, c_i64, c_ptr],
    "ps_csr2csc": [c_ptr, c_i64, c_i64, c_ptr, c_ptr, c_ptr, c_ptr],
    "ps_invert_perm": [c_ptr, c_i64, c_ptr, c_ptr],
    "ps_csc_build": [c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_i64, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr],
    "ps_gather_rows": [c_ptr, c_i64, c_ptr, c_i64, c_ptr, c_ptr],
    "ps_transpose_gather": [c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_i64, c_ptr, c_ptr, c_ptr, c_ptr],
    "ps_segment_reduce": [c_ptr, c_int, c_i64, c_i64, c_ptr, c_i64, c_int, c_ptr, c_ptr],
    "ps_gather_csr": [c_ptr, c_i64, c_ptr, c_i64, c_i64, c_ptr, c_ptr],
    "ps_scatter_reduce": [c_ptr, c_int, c_i64, c_ptr, c_i64, c_i64, c_int, c_ptr, c_ptr, c_ptr],
    "ps_reduce_all": [c_ptr, c_int, c_i64, c_int, c_ptr, c_ptr],
    "ps_mul_row": [c_ptr, c_int, c_i64, c_ptr, c_i64, c_i64, c_ptr, c_ptr, c_ptr],
    "ps_mul_col": [c_ptr, c_int, c_i64, c_ptr, c_i64, c_ptr, c_ptr, c_ptr],
    "ps_mul_nnz": [c_ptr, c_int, c_i64, c_ptr, c_ptr, c_ptr],
    "ps_add_row": [c_ptr, c_int, c_i64, c_ptr, c_i64, c_i64, c_ptr, c_ptr, c_ptr],
    "ps_add_col": [c_ptr, c_int, c_i64, c_ptr, c_i64, c_ptr, c_ptr, c_ptr],
    "ps_add_nnz": [c_ptr, c_int, c_i64, c_ptr, c_ptr, c_ptr],
    "ps_pad_i64": [c_ptr, c_i64, c_i64, c_i64, c_ptr, c_ptr],
    "ps_reshape_index": [c_ptr, c_ptr, c_i64, c_i64, c_i64, c_ptr, c_ptr, c_ptr],
    "ps_concat_add_i64": [c_ptr, c_ptr, c_ptr, c_i64, c_ptr, c_ptr],
    "ps_concat_bytes": [c_ptr, c_ptr, c_i64, c_ptr, c_ptr],
    "ps_coo_match": [c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_ptr],
    "ps_mask_to_index": [c_ptr, c_i64, c_ptr, c_ptr, c_ptr],
    "ps_mask_entries": [c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_ptr],
    "ps_narrow_cols": [c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_i64, c_i64, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr],
    "ps_index_select_fill": [c_ptr, c_ptr, c_ptr, c_i64, c_ptr, c_i64, c_ptr, c_i64, c_ptr, c_ptr, c_ptr, c_ptr,
                             c_ptr],
    "ps_sample_adj_bound": [c_ptr, c_ptr, c_i64, c_i64, c_int, c_ptr, c_ptr],
    "ps_sample_adj": [c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_i64, c_i64, c_int, c_u64, c_i64, c_ptr, c_ptr, c_ptr, c_ptr,
                      c_ptr, c_ptr],
    "ps_dist_row_histogram": [c_ptr, c_i64, c_int, c_i64, c_ptr, c_ptr],
    "ps_dist_partition": [c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr],
    "ps_dist_dest": [c_ptr, c_i64, c_ptr, c_int, c_ptr, c_ptr, c_ptr],
    "ps_sample_uniform": [c_ptr, c_ptr, c_i64, c_ptr, c_i64, c_i64, ctypes.c_uint64, c_ptr, c_ptr],
    "ps_dist_scatter_peer": [c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_int, c_ptr, c_ptr, c_ptr, c_ptr, c_int, c_ptr],
    "ps_coalesce_u32": [c_ptr, c_ptr, c_i64, c_i64, c_i64, c_i64, c_ptr, c_int, c_i64, c_int, c_ptr, c_ptr, c_ptr, c_ptr,
                        c_ptr, c_ptr],
    "ps_add_scalar": [c_ptr, c_i64, c_i64, c_ptr, c_ptr],
    "ps_synth_coo": [c_u64, c_i64, c_i64, c_i64, c_i64, c_ptr, c_ptr, c_ptr, c_ptr],
    "ps_synth_rmat": [c_u64, c_i64, c_i64, c_i64, c_i64, c_ptr, c_ptr, c_ptr, c_ptr],
}
_RESTYPES = {"ps_cuda_version": c_i64, "ps_last_error": ctypes.c_char_p, "ps_launch_count": c_i64, "ps_coo_inspect_bytes": c_i64,
             "ps_prof_enable": None, "ps_prof_name": ctypes.c_char_p}

_lib = None


def lib() -> ctypes.CDLL:
    """Open the library once; fail loudly (no fallback) when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: the CUDA library has not been built. Run "
                "`python -c 'import __graft_entry__ as g; g.build()'` (or `make -C paddle_sparse_b200/csrc`). "
                "paddle_sparse_b200 has no CPU fallback."
            )
        handle = ctypes.CDLL(LIB_PATH)
        for name, argtypes in SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError here = header/library out of step
            fn.argtypes = argtypes
            fn.restype = _RESTYPES.get(name, c_int)
        _lib = handle
    return _lib


def last_error() -> str:
    msg = lib().ps_last_error()
    return msg.decode("utf-8", "replace") if msg else ""


def call(name: str, *args) -> None:
    """Invoke an int-status entry point; raise RuntimeError(ps_last_error()) on failure."""
    rc = getattr(lib(), name)(*args)
    if rc != 0:
        raise RuntimeError(f"{name} failed (status {rc}): {last_error()}")
