"""Process-group bootstrap: ``torchrun`` replaces ``mpiexec``/``srun``.

One process per GPU; NCCL when CUDA is present (used only for bootstrap, stats gathers and the
"NCCL baseline" aggregation back-end), gloo on CPU.  Rendezvous defaults to 127.0.0.1.
"""
from __future__ import annotations

import datetime
import os
from typing import List, Tuple

import torch
import torch.distributed as dist


def init_distributed(backend: str = "auto") -> Tuple[int, int, int]:
    """Returns ``(rank, world, local_rank)``; a no-op single process when not under torchrun."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        if backend == "auto":
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        kw = {}
        if backend == "nccl":
            torch.cuda.set_device(local_rank % torch.cuda.device_count())
            kw["device_id"] = torch.device("cuda", local_rank % torch.cuda.device_count())
        dist.init_process_group(backend=backend, rank=rank, world_size=world,
                                timeout=datetime.timedelta(seconds=600), **kw)
    return rank, world, local_rank


def barrier() -> None:
    if dist.is_available() and dist.is_initialized():
        dist.barrier()


def gather_rows(rows: List[dict]) -> List[dict]:
    """``comm.gather(rows, root=0)`` equivalent: rank 0 gets the flattened list, others ``[]``."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return list(rows)
    out = [None] * dist.get_world_size() if dist.get_rank() == 0 else None
    dist.gather_object(rows, out, dst=0)
    return [r for part in out for r in part] if out is not None else []


def shutdown() -> None:
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()
