"""Pinned host -> device copy bandwidth of this box (what bounds bench.py's e2e number): event-timed
cudaMemcpyAsync through the repo's staging entry point.  Under torchrun every rank copies at the same time
(after a barrier), so the printed per-rank and summed figures are the CONCURRENT host->device bandwidth.

    python bench/h2d_probe.py            |  torchrun --nproc-per-node 8 bench/h2d_probe.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from crossscale.ops import _lib

world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
torch.cuda.set_device(dev)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
lib = _lib.cuda_lib()
cs = torch.cuda.Stream()
for mb in (4, 12.9, 64):
    n = int(mb * 1e6)
    h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    h.fill_(1)
    d = torch.empty(n, dtype=torch.uint8, device=dev)
    for _ in range(3):
        lib.cs_stage_h2d(d.data_ptr(), h.data_ptr(), n, cs.cuda_stream)
    cs.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
    reps = 40
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(cs)
    for _ in range(reps):
        lib.cs_stage_h2d(d.data_ptr(), h.data_ptr(), n, cs.cuda_stream)
    b.record(cs)
    cs.synchronize()
    ms = a.elapsed_time(b) / reps
    gbs = torch.tensor([n / ms / 1e6], device=dev)
    if world > 1:
        allg = [torch.zeros_like(gbs) for _ in range(world)]
        dist.all_gather(allg, gbs)
        vals = [float(t) for t in allg]
    else:
        vals = [float(gbs)]
    if rank == 0:
        print(f"H2D {mb:6.1f} MB x {world} rank(s) concurrently: per rank min {min(vals):.1f} / max {max(vals):.1f} GB/s, "
              f"sum {sum(vals):.1f} GB/s", flush=True)
if world > 1:
    dist.destroy_process_group()
