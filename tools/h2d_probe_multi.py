"""Pinned host -> device copy rate with all GPUs of the box copying at once: the ceiling of the
end-to-end bench line at N GPUs.  Run under torchrun like bench.py:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29512 tools/h2d_probe_multi.py
Every rank copies its own 5.5 GB pinned buffer (the bench's per-GPU batch) to its own GPU in
64 MiB pieces, five times, between barriers; prints per-rank and summed GB/s."""
import os
import time

import torch
import torch.distributed as dist

rank = int(os.environ.get("RANK", 0))
world = int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))

n = 5_502_926_848
chunk = 64 << 20
h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
h.fill_(rank + 1)
d = torch.empty(n, dtype=torch.uint8, device="cuda")


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def once():
    for o in range(0, n, chunk):
        e = min(n, o + chunk)
        d[o:e].copy_(h[o:e], non_blocking=True)
    torch.cuda.synchronize()


once()
barrier()
t0 = time.perf_counter()
reps = 5
for _ in range(reps):
    once()
mine = time.perf_counter() - t0
barrier()
t = torch.tensor([mine], dtype=torch.float64, device="cuda")
if world > 1:
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
slowest = float(t.item())
print(f"rank {rank}: {reps * n / mine / 1e9:6.2f} GB/s", flush=True)
barrier()
if rank == 0:
    print(f"{world} GPUs together: {world * reps * n / slowest / 1e9:7.2f} GB/s "
          f"({world * reps * n / slowest / 2e9:7.2f} Gsample/s of uint8 IQ), slowest rank {slowest / reps * 1e3:.1f} ms per 5.5 GB")
if world > 1:
    dist.destroy_process_group()
