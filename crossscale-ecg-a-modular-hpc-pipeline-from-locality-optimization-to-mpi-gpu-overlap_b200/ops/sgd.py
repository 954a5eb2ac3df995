"""K6: flat SGD-momentum with fused unscale / found-inf skip / GradScaler update.

Replaces ``torch.optim.SGD(lr=1e-2, momentum=0.9)`` + ``scaler.step`` + ``scaler.update``
(``part3_fedavg_overlap_mpi_gpu.py:125-127,181``) with ONE kernel over the model's flat parameter arena
(three small launches when the scaler is on: inf check, update, scale update — all on device).
"""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib

CTRL_STEPS, CTRL_SKIPPED = 2, 3
FS_SCALE = 1


class FlatSGD:
    def __init__(self, params: torch.Tensor, lr: float = 1e-2, momentum: float = 0.9, weight_decay: float = 0.0,
                 use_scaler: bool = False, init_scale: float = 65536.0, growth: float = 2.0, backoff: float = 0.5,
                 growth_interval: int = 2000, bf16_shadow: bool = False):
        assert params.is_cuda and params.dtype == torch.float32 and params.data_ptr() % 16 == 0
        self.params, self.lr, self.momentum, self.wd = params, float(lr), float(momentum), float(weight_decay)
        self.use_scaler, self.growth, self.backoff, self.interval = bool(use_scaler), growth, backoff, growth_interval
        dev = params.device
        self.mom = torch.zeros_like(params)
        self.ctrl = torch.zeros(8, dtype=torch.int32, device=dev)
        self.fstate = torch.zeros(8, dtype=torch.float32, device=dev)
        self.fstate[FS_SCALE] = init_scale if use_scaler else 1.0
        self.shadow: Optional[torch.Tensor] = torch.empty_like(params, dtype=torch.bfloat16) if bf16_shadow else None

    @property
    def scale(self) -> torch.Tensor:
        return self.fstate[FS_SCALE:FS_SCALE + 1]

    def step(self, grad: torch.Tensor) -> None:
        assert grad.numel() == self.params.numel() and grad.dtype == torch.float32
        _lib.check(_lib.cuda_lib().cs_sgd_momentum(
            self.params.data_ptr(), self.mom.data_ptr(), grad.data_ptr(), self.params.numel(), self.lr, self.momentum,
            self.wd, int(self.use_scaler), self.ctrl.data_ptr(), self.fstate.data_ptr(), self.growth, self.backoff,
            self.interval, _lib.ptr(self.shadow), _lib.stream_ptr()), "cs_sgd_momentum")

    def skipped_steps(self) -> int:
        return int(self.ctrl[CTRL_SKIPPED].item())
