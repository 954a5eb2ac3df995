"""Fused TinyECG kernels (K6b): forward, backward and the one-launch training step.

Replaces the ~40 library launches + ``cuda.synchronize`` + ``loss.item()`` per step of the reference
(``train_step_G0/G1``, ``TRUE_FL_M3/part3_fedavg_overlap_mpi_gpu.py:103-132``).  ``amp`` picks the
kernel: ``AMP_FP32`` -> fp32 FFMA (G0), ``AMP_BF16``/``AMP_FP16`` -> tcgen05 implicit GEMM (G1).
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Optional

import torch

from . import _lib

AMP_FP32, AMP_BF16, AMP_FP16 = 0, 1, 2
AMP_BY_NAME = {"none": AMP_FP32, "fp32": AMP_FP32, "bf16": AMP_BF16, "fp16": AMP_FP16}
CTRL_TICKET, CTRL_GROWTH, CTRL_STEPS, CTRL_SKIPPED, CTRL_FOUND_INF = range(5)
FS_LOSS_SUM, FS_SCALE, FS_LAST_LOSS, FS_LOSS_STEP = range(4)
MAX_CLASSES = 8


def use_kernels() -> bool:
    """CUDA kernels are the path whenever a GPU is present (env override for A/B debugging only)."""
    return _lib.have_gpu() and os.environ.get("CROSSSCALE_FORCE_TORCH", "0") != "1"


def _x_args(x: torch.Tensor):
    if x.dim() == 3:
        assert x.shape[1] == 1, "TinyECG takes [B,1,L]"
        x = x[:, 0, :]
    assert x.is_cuda and x.stride(-1) == 1 and x.dtype in (torch.float32, torch.bfloat16)
    if not x.is_contiguous():
        x = x.contiguous()
    return x, int(x.dtype == torch.bfloat16)


def forward_logits(x: torch.Tensor, params: torch.Tensor, num_classes: int, amp: int = AMP_FP32,
                   idx: Optional[torch.Tensor] = None) -> torch.Tensor:
    x2, xb = _x_args(x)
    b = idx.numel() if idx is not None else x2.shape[0]
    out = torch.empty((b, num_classes), dtype=torch.float32, device=x.device)
    _lib.check(_lib.cuda_lib().cs_tiny_ecg_forward(
        x2.data_ptr(), _lib.ptr(idx), xb, b, x2.shape[1], num_classes, params.data_ptr(),
        out.data_ptr(), amp, _lib.stream_ptr()), "cs_tiny_ecg_forward")
    return out


def backward_grads(x: torch.Tensor, params: torch.Tensor, dlogits: torch.Tensor, num_classes: int,
                   amp: int = AMP_FP32, idx: Optional[torch.Tensor] = None,
                   out: Optional[torch.Tensor] = None) -> torch.Tensor:
    x2, xb = _x_args(x)
    b = dlogits.shape[0]
    g = out if out is not None else torch.zeros_like(params)
    _lib.check(_lib.cuda_lib().cs_tiny_ecg_backward(
        x2.data_ptr(), _lib.ptr(idx), xb, b, x2.shape[1], num_classes, params.data_ptr(),
        dlogits.contiguous().data_ptr(), g.data_ptr(), amp, _lib.stream_ptr()), "cs_tiny_ecg_backward")
    return g


class TinyECGLogits(torch.autograd.Function):
    """autograd bridge: ``logits = TinyECG(x; params)``; backward recomputes activations on chip.

    ``flat`` is the arena the kernels read; ``params`` are the six ``nn.Parameter`` views of it, passed
    so that autograd routes the flat gradient back to them.
    """

    @staticmethod
    def forward(ctx, x, flat, num_classes, *params):
        amp = AMP_BF16 if torch.is_autocast_enabled() else AMP_FP32
        ctx.save_for_backward(x, flat)
        ctx.meta = (num_classes, amp, [p.shape for p in params])
        return forward_logits(x, flat, num_classes, amp)

    @staticmethod
    def backward(ctx, dlogits):
        x, flat = ctx.saved_tensors
        num_classes, amp, shapes = ctx.meta
        g = backward_grads(x, flat, dlogits.float(), num_classes, amp)
        outs, off = [], 0
        for shp in shapes:
            n = 1
            for d in shp:
                n *= d
            outs.append(g[off:off + n].view(shp))
            off += n
        return (None, None, None, *outs)


@dataclass
class ScalerConfig:
    enabled: bool = False
    init_scale: float = 65536.0
    growth_factor: float = 2.0
    backoff_factor: float = 0.5
    growth_interval: int = 2000


class TinyTrainState:
    """Device-resident optimiser + AMP-scaler + loss-accumulator state of the fused step."""

    def __init__(self, params: torch.Tensor, num_classes: int, lr: float = 1e-2, momentum: float = 0.9,
                 weight_decay: float = 0.0, amp: int = AMP_FP32, scaler: Optional[ScalerConfig] = None):
        assert params.is_cuda and params.dtype == torch.float32
        if not 1 <= num_classes <= MAX_CLASSES:
            raise ValueError(f"fused TinyECG kernels support 1..{MAX_CLASSES} classes")
        self.params, self.num_classes = params, num_classes
        self.lr, self.momentum, self.wd, self.amp = float(lr), float(momentum), float(weight_decay), int(amp)
        self.scaler = scaler or ScalerConfig(enabled=(amp == AMP_FP16))
        dev = params.device
        self.mom = torch.zeros_like(params)
        self.gacc = torch.zeros(int(_lib.cuda_lib().cs_tiny_ecg_scratch_floats(num_classes)),
                                dtype=torch.float32, device=dev)
        self.ctrl = torch.zeros(8, dtype=torch.int32, device=dev)
        self.fstate = torch.zeros(8, dtype=torch.float32, device=dev)
        self.fstate[FS_SCALE] = self.scaler.init_scale if self.scaler.enabled else 1.0
        self.launches = 0

    def step(self, x: torch.Tensor, y: torch.Tensor, idx: Optional[torch.Tensor] = None,
             batch: Optional[int] = None) -> None:
        """One fused local step (asynchronous, capturable)."""
        x2, xb = _x_args(x)
        b = batch if batch is not None else (idx.numel() if idx is not None else x2.shape[0])
        s = self.scaler
        _lib.check(_lib.cuda_lib().cs_tiny_ecg_train_step(
            x2.data_ptr(), y.data_ptr(), _lib.ptr(idx), xb, b, x2.shape[1], self.num_classes,
            self.params.data_ptr(), self.mom.data_ptr(), self.gacc.data_ptr(), self.ctrl.data_ptr(),
            self.fstate.data_ptr(), self.lr, self.momentum, self.wd, self.amp, int(s.enabled),
            s.growth_factor, s.backoff_factor, s.growth_interval, _lib.stream_ptr()),
            "cs_tiny_ecg_train_step")
        self.launches += 1

    # host-side readers (each one synchronises: call them once per round, not per step)
    def loss_sum(self) -> float:
        return float(self.fstate[FS_LOSS_SUM].item())

    def reset_loss(self) -> None:
        self.fstate[FS_LOSS_SUM].zero_()

    def loss_scale(self) -> float:
        return float(self.fstate[FS_SCALE].item())

    def steps_done(self) -> int:
        return int(self.ctrl[CTRL_STEPS].item())

    def skipped_steps(self) -> int:
        return int(self.ctrl[CTRL_SKIPPED].item())
