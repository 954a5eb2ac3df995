"""Seeds, device selection, checkpoints.

Reference: ``set_basic_seeds`` ``Module_3/TRUE_FL_M3/part3_fedavg_overlap_mpi_gpu.py:66-70`` (seed ``1234 + rank``, ``:156``);
``setup_device`` ``:60-63`` / ``Module_3/part3_mpi_gpu_train.py:82-86`` (there every rank lands on ``cuda:0``; here
``cuda:LOCAL_RANK``).  The reference has no checkpointing; ``save/load_checkpoint`` store exactly the
``state_dict`` of ``Module_3/tiny_ecg_model.py:8-29``.
"""
from __future__ import annotations

import os
import random
from typing import Optional

import numpy as np
import torch


def set_basic_seeds(seed: int) -> None:
    """numpy + torch (+ all CUDA devices); FedAvg driver uses ``1234 + rank`` (ref ``:66-70,156``)."""
    random.seed(seed)
    np.random.seed(seed % (2 ** 32))
    torch.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)


def setup_device(local_rank: Optional[int] = None) -> torch.device:
    """One GPU per process: ``cuda:LOCAL_RANK`` (the reference puts every rank on ``cuda:0``, C31)."""
    if torch.cuda.is_available():
        if local_rank is None:
            local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        idx = local_rank % torch.cuda.device_count()
        torch.cuda.set_device(idx)
        return torch.device("cuda", idx)
    return torch.device("cpu")


def save_checkpoint(model: torch.nn.Module, path: str) -> None:
    """``torch.save`` of the plain fp32 CPU state_dict (reference key names/shapes)."""
    sd = {k: v.detach().to("cpu", torch.float32).clone() for k, v in model.state_dict().items()}
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    tmp = path + ".tmp"
    torch.save(sd, tmp)
    os.replace(tmp, path)


def load_checkpoint(model: torch.nn.Module, path: str, strict: bool = True):
    sd = torch.load(path, map_location="cpu", weights_only=True)
    with torch.no_grad():
        own = model.state_dict()
        missing = [k for k in own if k not in sd]
        extra = [k for k in sd if k not in own]
        if strict and (missing or extra):
            raise KeyError(f"checkpoint mismatch: missing={missing} unexpected={extra}")
        for k, v in sd.items():
            if k in own:
                own[k].copy_(v.to(own[k].device, own[k].dtype))  # in place: keeps flat-arena views
    return model
