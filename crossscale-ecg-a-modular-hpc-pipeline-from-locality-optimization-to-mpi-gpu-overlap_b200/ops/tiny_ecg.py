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
CTRL_TICKET, CTRL_GROWTH, CTRL_STEPS, CTRL_SKIPPED, CTRL_FOUND_INF, CTRL_ERROR, CTRL_POS = range(7)
ERR_LABEL, ERR_MBAR, ERR_GRID, ERR_PEER = 1, 2, 4, 8
SAMPLE_IDX, SAMPLE_ROWS, SAMPLE_PERM = 0, 1, 2
_ERR_TEXT = {ERR_LABEL: "a label outside [0, num_classes) reached the kernel",
             ERR_MBAR: "an mbarrier wait inside the kernel timed out",
             ERR_GRID: "the grid-wide step barrier timed out",
             ERR_PEER: "fused aggregation: a peer never arrived (local model kept)"}
FS_LOSS_SUM, FS_SCALE, FS_LAST_LOSS, FS_LOSS_STEP = range(4)
MAX_CLASSES = 8


def use_kernels() -> bool:
    """CUDA kernels are the path whenever a GPU is present (env override for A/B debugging only)."""
    return _lib.have_gpu() and os.environ.get("CROSSSCALE_FORCE_TORCH", "0") != "1"


_M64 = (1 << 64) - 1


def _splitmix64(z: int) -> int:
    z = (z + 0x9E3779B97F4A7C15) & _M64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _M64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _M64
    return z ^ (z >> 31)


def perm_index(i: int, n: int, seed: int, epoch: int = 0) -> int:
    """Host mirror of the round kernel's sampler (``csrc/cuda/tiny_ecg_round.cu::perm_index``): a bijection on
    ``[0, n)`` keyed by ``(seed, epoch)`` - the stand-in for ``torch.randperm`` of ``make_gpu_batch_iter``
    (``Module_3/shard_dataset.py:133``) that needs no kernel launch and no index array."""
    if n <= 1:
        return 0
    bits = max(1, (n - 1).bit_length())
    mask, sh = (1 << bits) - 1, (bits + 1) >> 1
    k0 = _splitmix64((seed + 0x632BE59BD9B4E019 * epoch) & _M64)
    k1 = _splitmix64(k0); k2 = _splitmix64(k1); k3 = _splitmix64(k2)
    x = i
    while True:
        for mul, k in ((0x9E3779B97F4A7C15, k0), (0xD6E8FEB86659FD93, k1), (0xCA5A826395121157, k2), (0xA24BAED4963EE407, k3)):
            x = (x * mul + k) & mask
            x ^= x >> sh
        if x < n:
            return x


def perm_batches(pos: int, count: int, n: int, batch: int, seed: int):
    """Dataset rows the sampler hands out for draw positions ``pos .. pos+count-1`` (whole batches)."""
    usable = (n // batch) * batch
    return [perm_index((pos + q) % usable, n, seed, (pos + q) // usable) for q in range(count)]


def validate_labels(y: torch.Tensor, num_classes: int) -> None:
    """One synchronising check at set-up time: the fused kernels index logits by label (stock PyTorch raises
    'Target out of bounds' for the same input; the kernels train such a sample as class 0 and set an error bit)."""
    if y.numel() == 0:
        return
    lo, hi = int(y.min().item()), int(y.max().item())
    if lo < 0 or hi >= num_classes:
        raise ValueError(f"labels span [{lo}, {hi}] but the model has {num_classes} classes "
                         f"(shards written with a 5-class .lbl sidecar need --num-classes 5)")


def round_kernel_available() -> bool:
    return _lib.have_gpu() and _lib.is_sm100() and hasattr(_lib.cuda_lib(), "cs_tiny_ecg_train_round")


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
        self.gacc_round: Optional[torch.Tensor] = None

    def round_supported(self, window: Optional[int] = None) -> bool:
        """bf16 / fp16: always.  fp32 runs the round on the tensor core as TF32 (tiny_ecg_round_tf32.cu,
        windows up to 512 samples); CROSSSCALE_G0_TF32=0 keeps G0 on the fp32 FFMA step kernel."""
        if self.amp != AMP_FP32:
            return True
        import os
        return os.environ.get("CROSSSCALE_G0_TF32", "1") != "0" and (window is None or window <= 512)

    def round(self, x: torch.Tensor, y: torch.Tensor, steps: int, batch: int, idx: Optional[torch.Tensor] = None,
              sample_mode: int = SAMPLE_IDX, seed: int = 0, comm=None, prev: Optional[torch.Tensor] = None,
              weight: float = 1.0, wire_bf16: bool = False, server_lr: float = 1.0, reset_momentum: bool = False,
              arena_off: int = 0, channel: int = 0) -> None:
        """``steps`` local steps (+ the FedAvg aggregation when ``comm`` is given) in ONE cooperative launch
        (``cs_tiny_ecg_train_round``).  Asynchronous; tcgen05 path only."""
        x2, xb = _x_args(x)
        lib = _lib.cuda_lib()
        if self.gacc_round is None:
            self.gacc_round = torch.zeros(int(lib.cs_tiny_ecg_round_scratch_floats(self.num_classes)),
                                          dtype=torch.float32, device=self.params.device)
        s = self.scaler
        _lib.check(lib.cs_tiny_ecg_train_round(
            x2.data_ptr(), y.data_ptr(), _lib.ptr(idx), xb, x2.shape[0], batch, x2.shape[1], self.num_classes,
            steps, sample_mode, seed & _M64, self.params.data_ptr(), self.mom.data_ptr(), self.gacc_round.data_ptr(),
            self.ctrl.data_ptr(), self.fstate.data_ptr(), self.lr, self.momentum, self.wd, self.amp, int(s.enabled),
            s.growth_factor, s.backoff_factor, s.growth_interval, comm, _lib.ptr(prev), float(weight), int(wire_bf16),
            float(server_lr), int(reset_momentum), int(arena_off), int(channel), _lib.stream_ptr()),
            "cs_tiny_ecg_train_round")
        self.launches += 1

    def error_bits(self) -> int:
        return int(self.ctrl[CTRL_ERROR].item())

    def check_errors(self) -> None:
        """Raise if a kernel flagged a device-side error (synchronises: call it where the loss is read)."""
        bits = self.error_bits()
        if bits:
            self.ctrl[CTRL_ERROR].zero_()
            raise _lib.NativeError("TinyECG kernel error: " + "; ".join(t for b, t in _ERR_TEXT.items() if bits & b))

    def samples_drawn(self) -> int:
        return int(self.ctrl[CTRL_POS:CTRL_POS + 2].view(torch.int64).item())

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
