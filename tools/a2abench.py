"""all_to_all_single bandwidth between the GPUs of one box (run under torchrun): what the exchange step can expect."""
import os
import sys

import torch
import torch.distributed as dist

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
n = 201960734 // world * world
for dtype, label in ((torch.int64, "int64 [n]"), (torch.float32, "f32 [n]")):
    x = torch.ones(n, dtype=dtype, device="cuda")
    y = torch.empty_like(x)
    for _ in range(3):
        dist.all_to_all_single(y, x)
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        dist.all_to_all_single(y, x)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    sent = x.numel() * x.element_size() * (world - 1) / world
    if rank == 0:
        print(f"{label}: {ms:.3f} ms per all_to_all, {sent / 1e9:.2f} GB sent per GPU -> {sent / ms / 1e6:.0f} GB/s per direction "
              f"(NCCL_* env: { {k: v for k, v in os.environ.items() if k.startswith('NCCL_')} })")
dist.destroy_process_group()
