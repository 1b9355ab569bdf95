"""Every consumer kernel of the built storage, launched a few times on the configs[2] shape (Reddit: broadcast mul,
reductions, ptr2ind / ind2ptr) and the configs[3] shape (sample_adj on the products graph): the command line the
per-kernel ncu captures of the consumers are taken from (tools/opbench.py times the same calls).
   python tools/prof_consumers.py [--small] [--rmat] [--once]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import paddle_sparse_b200 as ps  # noqa: E402
from paddle_sparse_b200 import ops  # noqa: E402

small = "--small" in sys.argv
rmat = "--rmat" in sys.argv
M = N = 232965
nnz = 114615892 if not small else 8000000
row, col, val = ops.synth_coo(0x5EED0003, 0, nnz, M, N, "cuda", rmat=rmat)
A = ps.SparseTensor(row=row, col=col, value=val, sparse_sizes=(M, N)).coalesce("add")
st = A.storage
n = st.col().numel()
rowptr, colv = st.rowptr(), st.col()
colcount = ops.scatter_add_ones(colv, N)
del row, col, val


def run():
    for dt in (torch.float32, torch.bfloat16):
        v = st.value().to(dt)
        o_row = torch.rand(M, device="cuda").to(dt)
        o_col = torch.rand(N, device="cuda").to(dt)
        ops.mul_row(v, rowptr, o_row)
        ops.mul_col(v, colv, o_col)
        for red in ("sum", "mean", "max"):
            ops.segment_csr(v, rowptr, None, red)
            ops.scatter(v, colv, 0, None, N, red)
            if red != "sum":
                ops.scatter(v, colv, 0, None, N, red, colcount)
    ops.ptr2ind(rowptr, n)
    ops.ind2ptr(st.row(), M)
    torch.cuda.synchronize()


run()
if "--once" not in sys.argv:
    run()
del A, st, rowptr, colv, colcount
torch.cuda.empty_cache()

M = N = 2449029
nnz = 61859140 if not small else 6000000
row, col, val = ops.synth_coo(0x5EED0002, 0, nnz, M, N, "cuda", rmat=rmat)
A = ps.SparseTensor(row=row, col=col, value=val, sparse_sizes=(M, N)).coalesce("add")
rowptr, colv = A.storage.rowptr(), A.storage.col()
g = torch.Generator(device="cuda").manual_seed(1)
seeds = torch.randperm(M, device="cuda", generator=g)[:1024].contiguous()
for _ in range(2):
    out1 = ops.sample_adj(rowptr, colv, seeds, 25, False, 42)
    out2 = ops.sample_adj(rowptr, colv, out1[2], 10, False, 42)
torch.cuda.synchronize()
print("ok", n, out1[1].numel(), out2[1].numel())
