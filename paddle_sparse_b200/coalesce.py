"""Free function `coalesce` (reference: paddle_sparse/coalesce.py:8-29)."""
from __future__ import annotations

import torch

from . import ops


def coalesce(index, value, m: int, n: int, op: str = "add"):
    """Row-wise sorts `index` and merges duplicate entries with `op` (add/sum/mean/min/max).
    Returns `(index[2, nnz'], value')`.  The reference builds a SparseStorage (argsort + 3 gathers)
    and calls `.coalesce` on it (key rebuild, 2 mask selects, nonzero, segment_csr); here it is ONE
    fused call: packed-key radix sort, run detection and segmented reduce."""
    row, col, value = ops.coalesce(index[0], index[1], value, m, n, op)
    return torch.stack([row, col], dim=0), value
