"""Row-range-partitioned build for matrices that span several GPUs of one NVSwitch box.

The reference has no distributed code; this is the sample sort named by the north star:

  1. every rank histograms its local tuples over coarse row buckets            (ps_dist_row_histogram)
  2. all-reduce of the histogram; identical, nnz-balanced, ROW-ALIGNED splitters on every rank
  3. stable partition of the local tuples by destination rank                  (ps_dist_partition)
  4. ONE all-to-all-v over NVLink: counts by NCCL all-gather, then row / col / value by device-to-device copies
     straight into the destination ranks' symmetric-memory receive buffers (all_to_all_single as the fall-back)
  5. shard-local coalesce + CSR/CSC build on local row ids                     (ps_coalesce, SparseStorage)

Duplicates of a key always meet on the rank that owns the row, so dedup, rowptr, row reductions and
row/col-broadcast mul need no further communication.  One process per GPU (torch.distributed).
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Optional, Tuple

import torch
import torch.distributed as dist

MAX_BUCKETS = 32768  # histogram counters live in shared memory (128 KB) up to this many buckets


def bucket_shift(M: int, max_buckets: int = MAX_BUCKETS) -> Tuple[int, int]:
    """Smallest shift s such that ceil(M / 2^s) <= max_buckets.  Returns (shift, buckets)."""
    shift = 0
    while ((M + (1 << shift) - 1) >> shift) > max_buckets:
        shift += 1
    return shift, max((M + (1 << shift) - 1) >> shift, 1)


def choose_splitters(hist: torch.Tensor, G: int, shift: int, M: int) -> torch.Tensor:
    """Row-aligned splitters r_0 = 0 <= r_1 <= ... <= r_G = M balancing nnz: r_g is the first bucket
    boundary at which the running count reaches g * total / G.  Deterministic given the (all-reduced)
    histogram, so every rank computes the same answer without another exchange."""
    h = hist.to("cpu", torch.int64)
    cum = torch.cumsum(h, 0)
    total = int(cum[-1]) if cum.numel() else 0
    targets = torch.tensor([(total * g + G - 1) // G for g in range(1, G)], dtype=torch.int64)
    # first bucket b with cum[b] >= target  ->  boundary after bucket b
    idx = torch.searchsorted(cum, targets, right=False) + 1 if G > 1 else targets
    rows = torch.clamp(idx << shift, max=M)
    out = torch.empty(G + 1, dtype=torch.int64)
    out[0], out[G] = 0, M
    if G > 1:
        out[1:G] = rows
    return torch.cummax(out, 0)[0]


@dataclass
class RowShard:
    """The rows [row_start, row_end) of the global matrix as a local SparseTensor with LOCAL row ids."""
    tensor: "object"
    row_start: int
    row_end: int
    sparse_sizes: Tuple[int, int]
    splitters: torch.Tensor
    sent: int = 0      # tuples this rank sent to other ranks
    received: int = 0  # tuples this rank received (including its own)
    transposed: "object" = None  # local N x rows transpose (column ids global) when requested

    def global_row(self) -> torch.Tensor:
        from . import ops
        return ops.add_scalar(self.tensor.storage.row(), self.row_start)


def exchange(row, col, value, counts: torch.Tensor, group=None):
    """all-to-all-v of the destination-grouped tuples: counts first, then one collective per array."""
    G = dist.get_world_size(group)
    send_counts = counts.to(torch.int64)
    recv_counts = torch.empty_like(send_counts)
    dist.all_to_all_single(recv_counts, send_counts, group=group)
    s = send_counts.tolist() if send_counts.device.type == "cpu" else send_counts.cpu().tolist()
    r = recv_counts.tolist() if recv_counts.device.type == "cpu" else recv_counts.cpu().tolist()
    n_recv = int(sum(r))

    def a2a(x):
        if x is None:
            return None
        out = torch.empty((n_recv,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
        dist.all_to_all_single(out, x.contiguous(), output_split_sizes=r, input_split_sizes=s, group=group)
        return out

    rank = dist.get_rank(group)
    sent = int(sum(s)) - int(s[rank])
    return a2a(row), a2a(col), a2a(value), sent, n_recv, G


class _PeerBuffers:
    """Receive buffers for the exchange in SYMMETRIC memory (torch.distributed._symmetric_memory): every rank of the
    group allocates the same capacity and maps its peers' buffers, so the exchange is device-to-device copies
    straight into the destination GPU's memory over NVLink / NVSwitch - no NCCL on the data path (measured at two
    B200s: 1.43 ms against 4.45 ms for all_to_all_single on the same 1.6 GB, tools/p2pbench.py, tools/a2abench.py).
    One set per process group, regrown collectively when a step needs more room."""

    def __init__(self, group, capacity: int, value_row_bytes: int, device):
        import torch.distributed._symmetric_memory as symm_mem
        self.capacity, self.value_row_bytes = capacity, value_row_bytes
        G = dist.get_world_size(group)
        pg = group if group is not None else dist.group.WORLD
        self.row = symm_mem.empty(capacity, dtype=torch.int64, device=device)
        self.col = symm_mem.empty(capacity, dtype=torch.int64, device=device)
        self.val = symm_mem.empty(max(capacity * value_row_bytes, 1), dtype=torch.uint8, device=device)
        self.h_row = symm_mem.rendezvous(self.row, pg)
        self.h_col = symm_mem.rendezvous(self.col, pg)
        self.h_val = symm_mem.rendezvous(self.val, pg)
        self.peer_row = [self.h_row.get_buffer(g, (capacity,), torch.int64) for g in range(G)]
        self.peer_col = [self.h_col.get_buffer(g, (capacity,), torch.int64) for g in range(G)]
        self.peer_val = [self.h_val.get_buffer(g, (max(capacity * value_row_bytes, 1),), torch.uint8) for g in range(G)]
        # the same base addresses as device tables, for the kernel that scatters straight into the peers' buffers

        def table(bufs):
            return torch.tensor([b.data_ptr() for b in bufs], dtype=torch.int64, device=device)
        self.tab_row, self.tab_col, self.tab_val = table(self.peer_row), table(self.peer_col), table(self.peer_val)


class PeerMemoryUnavailable(RuntimeError):
    """Raised on EVERY rank of the group when the symmetric receive buffers could not be set up on at least one."""


_peer_buffers = {}       # group -> _PeerBuffers
_peer_unavailable = []   # non-empty once symmetric memory failed to come up: the reason, and NCCL from then on


def _peer_plan(counts: torch.Tensor, value, device, group):
    """Counts of every (source, destination) pair on the host (one all-gather, one sync) and receive buffers large
    enough for this step.  Returns (c[src][dst], buffers, value bytes per entry)."""
    G = dist.get_world_size(group)
    mine = counts.to(torch.int64)
    table = torch.empty(G * G, dtype=torch.int64, device=mine.device)
    dist.all_gather_into_tensor(table, mine, group=group)
    c = table.view(G, G).cpu()
    need = int(c.sum(0).max())
    vb = 0 if value is None else value.element_size() * int(torch.Size(value.shape[1:]).numel())
    buf = _peer_buffers.get(group)
    if buf is None or buf.capacity < need or buf.value_row_bytes < vb:
        # every rank takes this branch together (need, capacity and vb are the same everywhere), so whether the
        # buffers came up is agreed COLLECTIVELY: one rank failing (no P2P, an allocation failure while regrowing)
        # must not leave the others waiting in the first barrier of the exchange
        cap = max(int(need * 1.125) + 1024, 0 if buf is None else buf.capacity)
        new, why = None, ""
        try:
            new = _PeerBuffers(group, cap, max(vb, 0 if buf is None else buf.value_row_bytes), device)
        except (RuntimeError, ImportError, AttributeError) as e:
            why = repr(e)
        ok = torch.tensor([0 if new is None else 1], dtype=torch.int32, device=device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
        if int(ok.item()) == 0:
            _peer_buffers.pop(group, None)
            raise PeerMemoryUnavailable(why or "symmetric memory did not come up on another rank")
        buf = _peer_buffers[group] = new
    return c, buf, vb


def partition_exchange_fused(row, col, value, splitters: torch.Tensor, group=None):
    """Steps 3 + 4 as ONE kernel over peer memory: the stable counting pass on the destination id stores each
    destination's run directly into that rank's symmetric receive buffer, so the tuples cross NVLink tile by tile
    while the pass is still ranking the next tile - nothing is staged locally and nothing is copied afterwards.
    Values: absent or 4 bytes per entry.  Same result and validity rule as exchange_peer."""
    from . import ops
    G, me = dist.get_world_size(group), dist.get_rank(group)
    dest, counts = ops.dist_dest(row, splitters)
    c, buf, vb = _peer_plan(counts, value, row.device, group)
    recv_base = torch.cumsum(c, 0) - c  # recv_base[src][dst]: where src's tuples start in dst's buffer
    bases = recv_base[me].to(row.device)
    n_recv = int(c[:, me].sum())
    buf.h_row.barrier()  # every rank has consumed what the previous exchange left in its buffers
    # row and col cross NVLink as 32-bit words (ids < 2^32 on this path: route_tuples checks): 12 instead of 20 bytes
    # per tuple; they land in the first half of the int64 receive buffers and the shard's coalesce reads them as
    # such (ops.coalesce_u32), with the shard's first row as offset
    ops.dist_scatter_peer(dest, row, col, value, bases, buf.tab_row, buf.tab_col, None if value is None else buf.tab_val,
                          narrow=True)
    buf.h_row.barrier()  # every rank's pass has finished writing into my buffers
    r_val = None
    if value is not None:
        r_val = buf.val[:n_recv * vb].view(value.dtype).reshape((n_recv,) + tuple(value.shape[1:]))
    sent = int(c[me].sum()) - int(c[me][me])
    return buf.row.view(torch.int32)[:n_recv], buf.col.view(torch.int32)[:n_recv], r_val, sent, n_recv


def exchange_peer(row, col, value, counts: torch.Tensor, group=None):
    """The all-to-all-v of `exchange` over peer memory.  Same result (tuples grouped by source rank, source order
    kept).  The returned row / col / value are views of this rank's symmetric receive buffers: they stay valid until
    the next exchange on the same group (build_partitioned consumes them at once)."""
    G, me = dist.get_world_size(group), dist.get_rank(group)
    c, buf, vb = _peer_plan(counts, value, row.device, group)  # c[src][dst]; the one host sync of the exchange
    send_off = torch.cumsum(c[me], 0) - c[me]       # where each destination's tuples start in my partitioned arrays
    recv_base = torch.cumsum(c, 0) - c              # recv_base[src][dst]: where src's tuples start in dst's buffer
    n_recv = int(c[:, me].sum())
    vbytes = None if value is None else value.contiguous().view(torch.uint8).reshape(-1)
    buf.h_row.barrier()  # every rank has consumed what the previous exchange left in its buffers
    for k in range(G):
        g = (me + k) % G  # staggered: at any moment every rank writes to a different peer
        n, so, do = int(c[me][g]), int(send_off[g]), int(recv_base[me][g])
        if n == 0:
            continue
        buf.peer_row[g][do:do + n].copy_(row[so:so + n], non_blocking=True)
        buf.peer_col[g][do:do + n].copy_(col[so:so + n], non_blocking=True)
        if vbytes is not None:
            buf.peer_val[g][do * vb:(do + n) * vb].copy_(vbytes[so * vb:(so + n) * vb], non_blocking=True)
    buf.h_row.barrier()  # every write into my buffers has landed
    r_val = None
    if value is not None:
        r_val = buf.val[:n_recv * vb].view(value.dtype).reshape((n_recv,) + tuple(value.shape[1:]))
    sent = int(c[me].sum()) - int(c[me][me])
    return buf.row[:n_recv], buf.col[:n_recv], r_val, sent, n_recv, G


def _exchange_auto(row, col, value, counts, group):
    """Peer-memory exchange on CUDA tensors across several ranks; NCCL / gloo all_to_all_single otherwise, and from
    then on if symmetric memory cannot be set up on this machine."""
    if row.is_cuda and dist.get_world_size(group) > 1 and not _peer_unavailable and \
            os.environ.get("PS_DIST_EXCHANGE", "fused") in ("peer", "fused"):
        try:
            return exchange_peer(row, col, value, counts, group)
        except (RuntimeError, ImportError, AttributeError) as e:  # no P2P / no symmetric memory on this box
            _peer_unavailable.append(repr(e))
            import warnings
            warnings.warn(f"paddle_sparse_b200.dist: peer-memory exchange unavailable ({e}); using all_to_all_single")
    return exchange(row, col, value, counts, group)


def route_tuples(row, col, value, sparse_sizes, group=None, _ops=None):
    """Steps 1-4: returns (row, col, value) of the tuples this rank owns (GLOBAL row ids, arrival order:
    grouped by source rank, source order kept) plus (splitters, sent, received).  `_ops` lets the CPU
    (gloo) tests drive this host logic with stand-in kernels; the product path uses the CUDA operators."""
    if _ops is None:
        from . import ops as _ops
    M = int(sparse_sizes[0])
    if max(int(sparse_sizes[0]), int(sparse_sizes[1])) > 2 ** 32:
        # ps_dist_partition / ps_dist_scatter_peer move row and col as 32-bit payload words (csrc/dist.cu)
        raise ValueError(f"partitioned build: sparse sizes {tuple(sparse_sizes)} exceed the 2^32 ids per dimension "
                         "the exchange carries")
    G = dist.get_world_size(group)
    shift, buckets = bucket_shift(M)
    hist = _ops.dist_row_histogram(row, shift, buckets)
    dist.all_reduce(hist, op=dist.ReduceOp.SUM, group=group)
    splitters = choose_splitters(hist, G, shift, M)
    mode = os.environ.get("PS_DIST_EXCHANGE", "fused")
    narrow = value is None or value.element_size() * int(torch.Size(value.shape[1:]).numel()) == 4
    if row.is_cuda and G > 1 and narrow and mode == "fused" and not _peer_unavailable:
        try:
            r_row, r_col, r_val, sent, n_recv = partition_exchange_fused(row, col, value, splitters.to(row.device), group)
            return r_row, r_col, r_val, splitters, sent, n_recv
        except (PeerMemoryUnavailable, ImportError, AttributeError) as e:  # no symmetric memory / no P2P: all ranks
            _peer_unavailable.append(repr(e))                                # land here together (_peer_plan)
            import warnings
            warnings.warn(f"paddle_sparse_b200.dist: fused peer-memory exchange unavailable ({e}); using NCCL")
    p_row, p_col, p_val, counts = _ops.dist_partition(row, col, value, splitters.to(row.device))
    r_row, r_col, r_val, sent, n_recv, _ = _exchange_auto(p_row, p_col, p_val, counts, group)
    return r_row, r_col, r_val, splitters, sent, n_recv


def build_partitioned(row, col, value, sparse_sizes, group=None, reduce: str = "add", fill_cache: bool = True,
                      transpose: bool = False) -> RowShard:
    """Build this rank's row shard of the coalesced matrix from its local, unsorted tuples."""
    from . import ops
    from .tensor import SparseTensor
    M, N = int(sparse_sizes[0]), int(sparse_sizes[1])
    rank = dist.get_rank(group)
    r_row, r_col, r_val, splitters, sent, n_recv = route_tuples(row, col, value, (M, N), group)
    r0, r1 = int(splitters[rank]), int(splitters[rank + 1])
    if r_row.dtype == torch.int32:  # the fused exchange: 32-bit global ids, made local inside the coalesce's first read
        c_row, c_col, c_val, c_rowptr = ops.coalesce_u32(r_row, r_col, r0, r_val, max(r1 - r0, 0), N, reduce, True)
    else:
        local_row = ops.add_scalar(r_row, -r0) if r0 else r_row
        c_row, c_col, c_val, c_rowptr = ops.coalesce(local_row, r_col, r_val, max(r1 - r0, 0), N, reduce, True)
    A = SparseTensor(row=c_row, rowptr=c_rowptr, col=c_col, value=c_val, sparse_sizes=(max(r1 - r0, 0), N),
                     is_sorted=True, trust_data=True)
    if fill_cache:
        # a shard that will not be transposed does not carry row / value through its CSC sort (20 bytes per entry of
        # writes and two payload words per pass saved)
        A.storage.fill_cache_(carry=transpose)
    At = A.t() if transpose else None
    shard = RowShard(A, r0, r1, (M, N), splitters, sent, n_recv)
    shard.transposed = At
    return shard


# ------------------------------------------------------------------ second exchange
def combine_column_partials(partial: torch.Tensor, count: torch.Tensor, reduce: str, group=None) -> torch.Tensor:
    """Global `reduction(dim=0)` (reduce.py:40-42) from per-shard results.  `partial[c]` is this shard's
    reduction over its own entries of column c - the SUM for sum / mean, the min / max otherwise, 0 where the
    shard has no entry (as scatter leaves it) - and `count[c]` how many entries that was.
    sum: all-reduce(SUM).  mean: all-reduce(SUM) of sums and counts, one division (floor for integers, as
    scatter's mean).  min / max: columns a shard does not touch are neutralised before all-reduce(MIN / MAX) and
    columns nobody touches come back as 0, the value scatter gives an empty segment."""
    reduce = "sum" if reduce == "add" else reduce
    if reduce not in ("sum", "mean", "min", "max"):
        raise ValueError(reduce)
    wide = (slice(None),) + (None,) * (partial.dim() - 1)
    if reduce == "sum":
        acc = partial.clone()
        dist.all_reduce(acc, op=dist.ReduceOp.SUM, group=group)
        return acc
    total = count.clone()
    dist.all_reduce(total, op=dist.ReduceOp.SUM, group=group)
    if reduce == "mean":
        acc = partial.clone()
        dist.all_reduce(acc, op=dist.ReduceOp.SUM, group=group)
        den = total.clamp(min=1)[wide]
        if acc.dtype.is_floating_point:
            return acc / den.to(acc.dtype)
        return torch.div(acc, den, rounding_mode="floor")
    info = torch.finfo(partial.dtype) if partial.dtype.is_floating_point else torch.iinfo(partial.dtype)
    neutral = info.max if reduce == "min" else info.min
    acc = torch.where((count > 0)[wide], partial, torch.full_like(partial, neutral))
    dist.all_reduce(acc, op=dist.ReduceOp.MIN if reduce == "min" else dist.ReduceOp.MAX, group=group)
    return torch.where((total > 0)[wide], acc, torch.zeros_like(acc))


def reduce_columns(shard: RowShard, reduce: str = "sum", group=None) -> torch.Tensor:
    """`A.sum(dim=0)` / mean / min / max of the GLOBAL matrix, identical on every rank: the shard-local scatter
    reduction over `col` ([N] partials, ps_scatter_reduce) followed by one all-reduce over NVLink."""
    from .reduce import reduction
    reduce = "sum" if reduce == "add" else reduce
    A = shard.tensor
    partial = reduction(A, dim=0, reduce="sum" if reduce == "mean" else reduce)
    return combine_column_partials(partial, A.storage.colcount(), reduce, group)


def transpose_partitioned(shard: RowShard, group=None, fill_cache: bool = True) -> RowShard:
    """Global `A.t()` (transpose.py:9-33) under row partitioning: the rows of the transpose are the columns of
    A, which every shard holds a slice of, so this is the second all-to-all - keyed on col - through the same
    route / coalesce / CSR build.  Returns this rank's row shard of the N x M transpose."""
    st = shard.tensor.storage
    M, N = shard.sparse_sizes
    return build_partitioned(st.col(), shard.global_row(), st.value(), (N, M), group, "add", fill_cache, False)
