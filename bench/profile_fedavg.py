"""ncu driver: the fused FedAvg kernels with 8 emulated ranks on ONE GPU (one cooperative launch over all
ranks — never separate kernels that wait on one another): one-shot (TinyECG, 1458 floats) and two-shot
(4 M floats per rank)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from crossscale.ops import _lib
from crossscale.parallel.fedavg import aggregate_local_group
from crossscale.parallel.symm import LocalGroup

dev = torch.device("cuda", 0)
world = 8
for n in (1458, 4 << 20):
    grp = LocalGroup(world, int(_lib.cuda_lib().cs_fedavg_arena_bytes(n, 0, world)))
    prev = torch.randn(n, device=dev)
    for it in range(3):
        thetas = [prev + 0.01 * torch.randn(n, device=dev) for _ in range(world)]
        prevs = [prev.clone() for _ in range(world)]
        aggregate_local_group(grp, thetas, prevs, [1.0 / world] * world, n)
    torch.cuda.synchronize()
    print(n, "status", grp.status(), float(thetas[0].mean()))
    grp.close()
