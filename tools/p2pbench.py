"""Peer-memory exchange between the GPUs of one box without NCCL on the data path (run under torchrun):
symmetric-memory receive buffers, device-to-device copies into the peer's buffer over NVLink, device-side barrier."""
import os

import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
n = 201960734 // world * world
per = n // world
recv = symm_mem.empty(n, dtype=torch.int64, device="cuda")
hdl = symm_mem.rendezvous(recv, dist.group.WORLD)
src = torch.full((n,), rank + 1, dtype=torch.int64, device="cuda")
peers = [hdl.get_buffer(g, (n,), torch.int64) for g in range(world)]


def exchange():
    hdl.barrier()  # every rank is done reading its receive buffer
    for k in range(world):
        g = (rank + k) % world  # staggered start: no two ranks hammer the same peer first
        peers[g][rank * per:(rank + 1) * per].copy_(src[g * per:(g + 1) * per], non_blocking=True)
    hdl.barrier()  # every write has landed


for _ in range(3):
    exchange()
torch.cuda.synchronize()
ok = all(bool((recv[g * per:(g + 1) * per] == g + 1).all()) for g in range(world))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    exchange()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
sent = 8 * per * (world - 1)
if rank == 0:
    print(f"peer-copy exchange of int64[{n}]: {ms:.3f} ms, {sent / 1e9:.2f} GB sent per GPU -> {sent / ms / 1e6:.0f} GB/s per direction, "
          f"data ok: {ok}")
dist.destroy_process_group()
