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


def mpi_avg(value: float) -> float:
    """Mean of a python scalar over ranks (``mpi_avg``, ``Module_3/part3_mpi_gpu_train.py:78-79``)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.tensor([float(value)], dtype=torch.float64, device=dev)
    dist.all_reduce(t)
    return float(t.item()) / dist.get_world_size()


def file_allgather(directory: str, tag: str, rank: int, world: int, blob: bytes, timeout_s: float = 60.0) -> List[bytes]:
    """All-gather of a small blob through a shared directory (native ``cs_rendezvous_allgather``): the
    bootstrap for processes that were not started by torchrun."""
    import ctypes as C
    from ..ops import _lib
    lib = _lib.rt_lib()
    out = (C.c_char * (len(blob) * world))()
    rc = lib.cs_rendezvous_allgather(directory.encode(), tag.encode(), rank, world, blob, len(blob), out, float(timeout_s))
    if rc != 0:
        raise RuntimeError(lib.cs_rendezvous_last_error().decode())
    raw = bytes(out.raw)
    return [raw[i * len(blob):(i + 1) * len(blob)] for i in range(world)]


def shutdown() -> None:
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()
