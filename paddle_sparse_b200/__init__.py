"""paddle_sparse_b200 — B200-native (sm_100a) drop-in for the SparseStorage hot path of
PFCCLab/paddle_sparse: same `SparseTensor` / `SparseStorage` API and free functions
(paddle_sparse/__init__.py:34-85), every operation a hand-written CUDA kernel behind the C ABI of
include/ps_abi.h.  GPU only: there is no CPU fallback and importing fails loudly when the
shared library has not been built."""
from . import _lib

_lib.lib()  # fail at import time, not at first use, when libpaddle_sparse_b200.so is missing

from . import ops  # noqa: E402
from . import paddle_sparse_ops  # noqa: E402  (the module object of paddle_sparse/__init__.py:8-17, over DLPack)
from .storage import SparseStorage  # noqa: E402
from .tensor import SparseTensor  # noqa: E402
from .coalesce import coalesce  # noqa: E402
from .transpose import t, transpose  # noqa: E402
from .reduce import sum, mean, min, max, reduction  # noqa: E402,A004
from .mul import mul, mul_, mul_nnz, mul_nnz_  # noqa: E402
from .add import add, add_, add_nnz, add_nnz_  # noqa: E402
from .index_select import index_select, index_select_nnz  # noqa: E402
from .masked_select import masked_select, masked_select_nnz  # noqa: E402
from .narrow import narrow, select, __narrow_diag__  # noqa: E402
from .permute import permute  # noqa: E402
from .cat import cat  # noqa: E402
from .convert import (to_scipy, from_scipy, to_torch_sparse, from_torch_sparse, to_paddle_sparse,  # noqa: E402
                      from_paddle_sparse)
from .sample import sample, sample_adj, manual_seed  # noqa: E402

__version__ = "0.1.0"
cuda_version = paddle_sparse_ops.sparse_cuda_version().item()  # paddle_sparse/__init__.py:17, verbatim

__all__ = [
    "SparseStorage", "SparseTensor", "coalesce", "t", "transpose", "sum", "mean", "min", "max", "reduction",
    "mul", "mul_", "mul_nnz", "mul_nnz_", "add", "add_", "add_nnz", "add_nnz_", "index_select", "index_select_nnz", "masked_select",
    "masked_select_nnz", "narrow", "__narrow_diag__", "select", "permute", "cat", "to_scipy", "from_scipy", "to_torch_sparse", "from_torch_sparse",
    "to_paddle_sparse", "from_paddle_sparse", "sample", "sample_adj", "manual_seed", "ops", "paddle_sparse_ops", "__version__",
]
