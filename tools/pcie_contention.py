#!/usr/bin/env python
"""Do the ranks of one box slow each other's host<->device copies down?  Every rank moves the e2e frame's bytes
(52 MB up, 108 MB down, page-locked) in a loop: first rank 0 alone, then all ranks at once.  Run under torchrun.
Prints GB/s per rank for both cases -- the explanation (or not) of the e2e arm's efficiency at N = 8."""
import os, sys, time
import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
UP, DOWN, REP = 52 << 20, 108 << 20, 40
hu = torch.empty(UP, dtype=torch.uint8, pin_memory=True); du = torch.empty(UP, dtype=torch.uint8, device=dev)
hd = torch.empty(DOWN, dtype=torch.uint8, pin_memory=True); dd = torch.empty(DOWN, dtype=torch.uint8, device=dev)
s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()


def run(active):
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    if active:
        for _ in range(REP):
            with torch.cuda.stream(s_up):
                du.copy_(hu, non_blocking=True)
            with torch.cuda.stream(s_dn):
                hd.copy_(dd, non_blocking=True)
        torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if world > 1:
        dist.barrier()
    return (UP + DOWN) * REP / dt / 1e9 if active else 0.0


run(True)                                     # warm-up
alone = run(rank == 0)
together = run(True)
vals = torch.tensor([alone, together], dtype=torch.float64, device=dev)
if world > 1:
    out = [torch.zeros_like(vals) for _ in range(world)]
    dist.all_gather(out, vals)
else:
    out = [vals]
if rank == 0:
    print("rank 0 alone (up + down at once): %.1f GB/s" % out[0][0].item())
    print("all %d ranks at once, GB/s per rank: %s" % (world, " ".join("%.1f" % o[1].item() for o in out)))
    print("sum %.1f GB/s" % sum(o[1].item() for o in out))
if world > 1:
    dist.destroy_process_group()
