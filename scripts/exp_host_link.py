"""What the host side of an N-GPU box can move: pinned host <-> device copies on all ranks at once (torchrun)."""
import os
import time

import torch
import torch.distributed as dist

rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
n = 256 << 20
hin = [torch.empty(n, dtype=torch.uint8).pin_memory() for _ in range(3)]
hout = [torch.empty(n, dtype=torch.uint8).pin_memory() for _ in range(3)]
din, dout = torch.empty(n, dtype=torch.uint8, device=dev), torch.ones(n, dtype=torch.uint8, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(mode, reps=6):
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for i in range(reps):
        if mode in ("h2d", "both"):
            with torch.cuda.stream(s1):
                din.copy_(hin[i % 3], non_blocking=True)
        if mode in ("d2h", "both"):
            with torch.cuda.stream(s2):
                hout[i % 3].copy_(dout, non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    gbs = reps * n * (2 if mode == "both" else 1) / dt * 1e-9
    t = torch.tensor([gbs], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t)
    if rank == 0:
        print("%-5s: %.1f GB/s on rank 0, %.1f GB/s over %d ranks" % (mode, gbs, float(t.item()), world), flush=True)


for m in ("h2d", "d2h", "both"):
    run(m, 2)
    run(m)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
