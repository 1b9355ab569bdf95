"""Neighbour sampling (reference: paddle_sparse/sample.py:10-57)."""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import ops
from .tensor import SparseTensor

_SEED = [0]


def manual_seed(seed: int) -> None:
    """Seed of the Philox stream behind `sample_adj` (the reference draws from Paddle's global
    generator, csrc/cpu/utils.h:22-34; call counter advances so successive batches differ)."""
    _SEED[0] = int(seed)


def sample(src: SparseTensor, num_neighbors: int, subset: Optional[torch.Tensor] = None,
           seed: Optional[int] = None) -> torch.Tensor:
    """sample.py:10-28: `num_neighbors` uniform draws with replacement per row of `subset` (every row when None), as
    ONE kernel (ops.sample_uniform) where the reference chains five tensor ops on the framework's generator.  `seed`
    (extension) pins the Philox stream of this call; otherwise the module's call counter advances it."""
    rowptr, col, _ = src.csr()
    if seed is None:
        seed = _SEED[0]
        _SEED[0] += 1
    return ops.sample_uniform(rowptr, col, subset, num_neighbors, seed)


def sample_adj(src: SparseTensor, subset: torch.Tensor, num_neighbors: int, replace: bool = False,
               seed: Optional[int] = None) -> Tuple[SparseTensor, torch.Tensor]:
    """sample.py:31-53.  `seed` (extension) pins the Philox stream of this call."""
    rowptr, col, value = src.csr()
    if seed is None:
        seed = _SEED[0]
        _SEED[0] += 1
    rowptr, col, n_id, e_id = ops.sample_adj(rowptr, col, subset, num_neighbors, replace, seed)
    if value is not None:
        value = ops.gather_rows(value, e_id)
    out = SparseTensor(rowptr=rowptr, row=None, col=col, value=value, sparse_sizes=(subset.shape[0], n_id.shape[0]),
                       is_sorted=True, trust_data=True)
    return out, n_id


SparseTensor.sample = sample
SparseTensor.sample_adj = sample_adj
