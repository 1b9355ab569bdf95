"""`permute` (reference: paddle_sparse/permute.py:8-13): the same row and column relabelling of a square
matrix, as two `index_select`s (each one load-balanced fill launch, index_select.py)."""
from __future__ import annotations

import torch

from .tensor import SparseTensor


def permute(src: SparseTensor, perm: torch.Tensor) -> SparseTensor:
    assert src.is_quadratic()
    return src.index_select(0, perm).index_select(1, perm)


SparseTensor.permute = lambda self, perm: permute(self, perm)
