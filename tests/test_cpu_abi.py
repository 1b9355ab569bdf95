"""CPU-side checks of the drop-in boundary: the C-ABI library loads without a GPU, exports every
symbol include/ps_abi.h declares, the ctypes signature table covers exactly those symbols, and the
operator layer refuses CPU tensors loudly (no fallback).  No compute calls."""
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "ps_abi.h")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ps_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_hot_path():
    names = declared_symbols()
    for must in ("ps_ind2ptr", "ps_ptr2ind", "ps_sample_adj", "ps_cuda_version", "ps_coalesce", "ps_sort_coo",
                 "ps_csr2csc", "ps_csc_build", "ps_invert_perm", "ps_transpose_gather", "ps_segment_reduce", "ps_scatter_reduce",
                 "ps_gather_csr", "ps_mul_row", "ps_mul_col", "ps_index_select_fill", "ps_mask_to_index",
                 "ps_dist_partition", "ps_dist_row_histogram"):
        assert must in names, must


def test_library_exports_every_declared_symbol():
    from paddle_sparse_b200 import _lib
    lib = _lib.lib()
    for name in declared_symbols():
        assert hasattr(lib, name), f"{name} declared in include/ps_abi.h but not exported"
    assert sorted(_lib.SIGNATURES) == declared_symbols()
    assert int(lib.ps_cuda_version()) >= 12000
    assert int(lib.ps_launch_count()) == 0  # nothing has been launched on this CPU box


def test_header_cites_the_reference_interface():
    text = open(HEADER).read()
    for cite in ("csrc/convert.cpp:13-23", "csrc/convert.cpp:46-56", "csrc/sample.cpp:8-24", "csrc/version.cpp",
                 "storage.py:454-486", "storage.py:425-434", "storage.py:211-222", "transpose.py:9-33", "reduce.py:42", "mul.py:76-93"):
        assert cite in text, cite


def test_no_cpu_fallback():
    import paddle_sparse_b200 as ps
    cpu = torch.tensor([0, 1], dtype=torch.int64)
    with pytest.raises(RuntimeError, match="GPU-only"):
        ps.ops.ind2ptr(cpu, 4)
    with pytest.raises(RuntimeError, match="GPU-only"):
        ps.SparseTensor(row=cpu, col=cpu)
    with pytest.raises(RuntimeError, match="GPU-only"):
        ps.coalesce(torch.stack([cpu, cpu]), None, 2, 2)


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from paddle_sparse_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _lib.lib()


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "paddle_sparse_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f
                assert '#include "../../oracle' not in src and "oracle/" not in src.replace("oracle/oracle_port.c", ""), f


def test_convert_round_trips_golden():
    """test/test_convert.py:9-30: the free converters are host-side glue and run without a GPU."""
    import torch
    import paddle_sparse_b200 as ps
    from tests.golden import reference_vectors as G
    g = G.CONVERT
    index, value, N = torch.tensor(g["index"]), torch.tensor(g["value"]), g["N"]
    out = ps.from_scipy(ps.to_scipy(index, value, N, N))
    assert out[0].tolist() == index.tolist() and out[1].tolist() == value.tolist()
    out = ps.from_torch_sparse(ps.to_torch_sparse(index, value, N, N).coalesce())
    assert out[0].tolist() == index.tolist() and out[1].tolist() == value.tolist()
    assert ps.to_paddle_sparse is ps.to_torch_sparse and ps.from_paddle_sparse is ps.from_torch_sparse
