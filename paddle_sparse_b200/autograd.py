"""Gradients of `value` through the kernels.

The reference gets value gradients for free because its hot path is a chain of differentiable Paddle /
paddle_scatter ops: `value[csr2csc]` (transpose.py:15), `segment_csr` (reduce.py:51, storage.py:471), `scatter`
(reduce.py:42), `gather_csr(...) * value` (mul.py:79-90).  Here those are raw kernels behind a C ABI, so each
value-producing operator is wrapped in a `torch.autograd.Function` whose backward is again one of this library's
kernels - the backward of a gather is a scatter-sum and vice versa, the backward of a segment reduction is a CSR
expansion.  `ops.py` routes through these wrappers only when a floating-point input requires grad and grad mode is
on; everything else keeps calling the kernels directly.  Fused entry points that move values along with the indices
(sort, coalesce, CSC build, index_select fill) are decomposed there into "indices by the kernel, values by a
differentiable gather / segment reduction", so no operator drops a gradient silently.
Ties in min / max share the gradient evenly (the convention of torch.amax / amin).
"""
from __future__ import annotations

import torch


def needs_grad(*tensors) -> bool:
    return torch.is_grad_enabled() and any(isinstance(t, torch.Tensor) and t.requires_grad for t in tensors)


def _ops():
    from . import ops
    return ops


def _wide(v: torch.Tensor, like: torch.Tensor) -> torch.Tensor:
    """[n] -> [n, 1, ...] to broadcast against `like`."""
    return v.reshape((v.shape[0],) + (1,) * (like.dim() - 1))


class GatherRows(torch.autograd.Function):
    """out = src[index] along dim 0 (storage.py:166-169, transpose.py:15)."""

    @staticmethod
    def forward(ctx, src, index):
        ctx.save_for_backward(index)
        ctx.n = src.shape[0]
        return _ops().gather_rows(src.detach(), index)

    @staticmethod
    def backward(ctx, grad):
        (index,) = ctx.saved_tensors
        return _ops().scatter(grad.contiguous(), index, 0, None, ctx.n, "sum"), None


class SegmentCSR(torch.autograd.Function):
    """paddle_scatter.segment_csr(src, indptr, None, reduce) (reduce.py:51, storage.py:471)."""

    @staticmethod
    def forward(ctx, src, indptr, reduce):
        ops = _ops()
        out = ops.segment_csr(src.detach(), indptr, None, reduce)
        ctx.reduce = "sum" if reduce == "add" else reduce
        ctx.E = src.shape[0]
        if ctx.reduce in ("min", "max"):
            ctx.save_for_backward(indptr, src.detach(), out)
        else:
            ctx.save_for_backward(indptr)
        return out

    @staticmethod
    def backward(ctx, grad):
        ops = _ops()
        grad = grad.contiguous()
        if ctx.reduce == "sum":
            (indptr,) = ctx.saved_tensors
            return ops.gather_csr(grad, indptr, ctx.E), None, None
        if ctx.reduce == "mean":
            (indptr,) = ctx.saved_tensors
            count = ops.ptr2count(indptr).clamp_(min=1).to(grad.dtype)
            return ops.gather_csr((grad / _wide(count, grad)).contiguous(), indptr, ctx.E), None, None
        indptr, src, out = ctx.saved_tensors
        hit = (src == ops.gather_csr(out, indptr, ctx.E)).to(grad.dtype)
        ties = ops.segment_csr(hit, indptr, None, "sum").clamp_(min=1)
        return ops.gather_csr((grad / ties).contiguous(), indptr, ctx.E) * hit, None, None


class GatherCSR(torch.autograd.Function):
    """paddle_scatter.gather_csr(src, indptr): src[s] repeated over segment s (mul.py:16,79)."""

    @staticmethod
    def forward(ctx, src, indptr, E):
        ctx.save_for_backward(indptr)
        return _ops().gather_csr(src.detach(), indptr, E)

    @staticmethod
    def backward(ctx, grad):
        (indptr,) = ctx.saved_tensors
        return _ops().segment_csr(grad.contiguous(), indptr, None, "sum"), None, None


class Scatter(torch.autograd.Function):
    """paddle_scatter.scatter(src, index, 0, None, dim_size, reduce) (reduce.py:42)."""

    @staticmethod
    def forward(ctx, src, index, dim_size, reduce, count):
        ops = _ops()
        out = ops.scatter(src.detach(), index, 0, None, dim_size, reduce, count)
        ctx.reduce = "sum" if reduce == "add" else reduce
        ctx.dim_size = dim_size
        if ctx.reduce in ("min", "max"):
            ctx.save_for_backward(index, src.detach(), out)
        else:
            ctx.save_for_backward(index)
        return out

    @staticmethod
    def backward(ctx, grad):
        ops = _ops()
        grad = grad.contiguous()
        if ctx.reduce == "sum":
            (index,) = ctx.saved_tensors
            return ops.gather_rows(grad, index), None, None, None, None
        if ctx.reduce == "mean":
            (index,) = ctx.saved_tensors
            count = ops.scatter_add_ones(index, ctx.dim_size).clamp_(min=1).to(grad.dtype)
            return ops.gather_rows((grad / _wide(count, grad)).contiguous(), index), None, None, None, None
        index, src, out = ctx.saved_tensors
        hit = (src == ops.gather_rows(out, index)).to(grad.dtype)
        ties = ops.scatter(hit, index, 0, None, ctx.dim_size, "sum").clamp_(min=1)
        return ops.gather_rows((grad / ties).contiguous(), index) * hit, None, None, None, None


class BroadcastRow(torch.autograd.Function):
    """value (op) other[row]  with op = * (mul.py:76-93) or + (add.py:50-66); `other` is [M, ...]."""

    @staticmethod
    def forward(ctx, value, rowptr, other, mul):
        ops = _ops()
        ctx.mul = mul
        ctx.save_for_backward(value.detach(), rowptr, other.detach())
        fn = ops.mul_row if mul else ops.add_row
        return fn(value.detach(), rowptr, other.detach())

    @staticmethod
    def backward(ctx, grad):
        ops = _ops()
        value, rowptr, other = ctx.saved_tensors
        grad = grad.contiguous()
        if ctx.mul:
            g_value = ops.mul_row(grad, rowptr, other) if ctx.needs_input_grad[0] else None
            g_other = ops.segment_csr((grad * value).contiguous(), rowptr, None, "sum") if ctx.needs_input_grad[2] else None
        else:
            g_value = grad if ctx.needs_input_grad[0] else None
            g_other = ops.segment_csr(grad, rowptr, None, "sum") if ctx.needs_input_grad[2] else None
        return g_value, None, g_other, None


class BroadcastCol(torch.autograd.Function):
    """value (op) other[col]; `other` is [N, ...]."""

    @staticmethod
    def forward(ctx, value, col, other, mul):
        ops = _ops()
        ctx.mul = mul
        ctx.save_for_backward(value.detach(), col, other.detach())
        fn = ops.mul_col if mul else ops.add_col
        return fn(value.detach(), col, other.detach())

    @staticmethod
    def backward(ctx, grad):
        ops = _ops()
        value, col, other = ctx.saved_tensors
        grad = grad.contiguous()
        N = other.shape[0]
        if ctx.mul:
            g_value = ops.mul_col(grad, col, other) if ctx.needs_input_grad[0] else None
            g_other = ops.scatter((grad * value).contiguous(), col, 0, None, N, "sum") if ctx.needs_input_grad[2] else None
        else:
            g_value = grad if ctx.needs_input_grad[0] else None
            g_other = ops.scatter(grad, col, 0, None, N, "sum") if ctx.needs_input_grad[2] else None
        return g_value, None, g_other, None


def run_starts(row: torch.Tensor, col: torch.Tensor) -> torch.Tensor:
    """Start of every run of equal (row, col) in a sorted sequence, plus the end: the `ptr` storage.py:462-470 builds
    (index arithmetic only; used on the gradient path, where the fused dedup kernel cannot be)."""
    n = row.numel()
    head = torch.ones(n, dtype=torch.bool, device=row.device)
    if n > 1:
        head[1:] = (row[1:] != row[:-1]) | (col[1:] != col[:-1])
    starts = head.nonzero().squeeze(1)
    return torch.cat([starts, torch.tensor([n], dtype=torch.int64, device=row.device)])
