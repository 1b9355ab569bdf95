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


def sample(src: SparseTensor, num_neighbors: int, subset: Optional[torch.Tensor] = None) -> torch.Tensor:
    """sample.py:10-28: uniform draws with replacement through the framework's RNG."""
    rowptr, col, _ = src.csr()
    rowcount = src.storage.rowcount()
    if subset is not None:
        rowcount = rowcount[subset]
        rowptr = rowptr[subset]
    else:
        rowptr = rowptr[:-1]
    rand = torch.rand((rowcount.shape[0], num_neighbors), device=col.device)
    rand.mul_(rowcount.to(rand.dtype).view(-1, 1))
    rand = rand.to(torch.int64)
    rand.add_(rowptr.view(-1, 1))
    return col[rand]


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
