"""TinyECG 1D-CNN.

Topology and ``state_dict`` contract of the reference model (``Module_3/tiny_ecg_model.py:8-29``;
``Tiny1D`` in ``Module_1/bench_locality.py:8-21`` is the same network)::

    net.0.weight [16,1,7]  net.0.bias [16]      Conv1d(1,16,k=7,pad=3)  + ReLU
    net.2.weight [16,16,5] net.2.bias [16]      Conv1d(16,16,k=5,pad=2) + ReLU + global avg-pool
    head.weight  [C,16]    head.bias  [C]       Linear(16,C)

all fp32, in that order, no buffers — a checkpoint saved by the stock ``nn.Module`` loads here and
vice versa.  B200 design: the six tensors are *views into one flat fp32 arena* (:class:`FlatParams`)
so that the fused training-step kernel, the SGD kernel and the in-kernel FedAvg aggregation all work
on one contiguous 1 509-float vector (5 classes) instead of six tensors.  On CUDA ``forward`` runs the
hand-written fused kernel (``ops.tiny_ecg``); on CPU it is plain PyTorch and doubles as the oracle.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Dict, List, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

C1_OUT, C1_K, C1_PAD = 16, 7, 3
C2_OUT, C2_K, C2_PAD = 16, 5, 2

PARAM_ORDER = ("net.0.weight", "net.0.bias", "net.2.weight", "net.2.bias",
               "head.weight", "head.bias")


def param_shapes(num_classes: int) -> "OrderedDict[str, Tuple[int, ...]]":
    return OrderedDict([
        ("net.0.weight", (C1_OUT, 1, C1_K)), ("net.0.bias", (C1_OUT,)),
        ("net.2.weight", (C2_OUT, C1_OUT, C2_K)), ("net.2.bias", (C2_OUT,)),
        ("head.weight", (num_classes, C2_OUT)), ("head.bias", (num_classes,)),
    ])


def param_offsets(num_classes: int) -> Dict[str, Tuple[int, int]]:
    """name -> (offset, numel) inside the flat arena (state_dict order, densely packed)."""
    off, out = 0, {}
    for k, shp in param_shapes(num_classes).items():
        n = 1
        for s in shp:
            n *= s
        out[k] = (off, n)
        off += n
    return out


def num_params(num_classes: int) -> int:
    o = param_offsets(num_classes)
    last = o["head.bias"]
    return last[0] + last[1]


class FlatParams:
    """Re-homes a module's parameters (and their grads) as views of flat fp32 buffers."""

    def __init__(self, module: nn.Module, pad_to: int = 4, align: int = 1):
        """``align`` > 1 starts every parameter at a multiple of ``align`` floats (16-byte aligned slices for
        bucketed collectives); the default packs densely — the layout the fused TinyECG kernels index."""
        params = list(module.parameters())
        self.offsets: List[int] = []
        total = 0
        for p in params:
            total = (total + align - 1) // align * align
            self.offsets.append(total)
            total += p.numel()
        padded = (total + pad_to - 1) // pad_to * pad_to
        dev = params[0].device if params else torch.device("cpu")
        self.numel = total
        self.flat = torch.zeros(padded, dtype=torch.float32, device=dev)
        self.grad = torch.zeros(padded, dtype=torch.float32, device=dev)
        for p, off in zip(params, self.offsets):
            n = p.numel()
            self.flat[off:off + n].copy_(p.data.reshape(-1).float())
            p.data = self.flat[off:off + n].view_as(p)
            p.grad = self.grad[off:off + n].view_as(p)
        self.params = params

    def rebind(self, new_flat: torch.Tensor) -> None:
        """Point every parameter at ``new_flat`` (e.g. a symmetric-memory arena)."""
        new_flat[: self.flat.numel()].copy_(self.flat)
        self.flat = new_flat
        for p, off in zip(self.params, self.offsets):
            p.data = new_flat[off:off + p.numel()].view_as(p)


    def rebind_grad(self, new_grad: torch.Tensor) -> None:
        """Point every parameter's gradient at ``new_grad`` (a peer-mapped arena: the bucketed gradient all-reduce then
        works in place on it, without staging copies)."""
        new_grad[: self.grad.numel()].copy_(self.grad)
        self.grad = new_grad
        for p, off in zip(self.params, self.offsets):
            p.grad = new_grad[off:off + p.numel()].view_as(p)


class TinyECG(nn.Module):
    def __init__(self, num_classes: int = 2):
        super().__init__()
        self.num_classes = int(num_classes)
        self.net = nn.Sequential(
            nn.Conv1d(1, C1_OUT, kernel_size=C1_K, padding=C1_PAD),
            nn.ReLU(inplace=True),
            nn.Conv1d(C1_OUT, C2_OUT, kernel_size=C2_K, padding=C2_PAD),
            nn.ReLU(inplace=True),
            nn.AdaptiveAvgPool1d(1),
        )
        self.head = nn.Linear(C2_OUT, self.num_classes)
        self._flat: FlatParams | None = None

    # -- flat arena ---------------------------------------------------------------------------
    def flatten_parameters(self) -> FlatParams:
        """Make the six tensors views of one contiguous buffer (idempotent per device)."""
        dev = self.head.weight.device
        if self._flat is None or self._flat.flat.device != dev:
            self._flat = FlatParams(self)
        return self._flat

    def _apply(self, fn, *a, **kw):  # .to()/.cuda() invalidate the views
        self._flat = None
        return super()._apply(fn, *a, **kw)

    # -- forward ------------------------------------------------------------------------------
    def forward_reference(self, x: torch.Tensor) -> torch.Tensor:
        """Plain PyTorch forward (CPU path and numerical oracle)."""
        return self.head(self.net(x).squeeze(-1))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.is_cuda:
            from ..ops import tiny_ecg as _k
            if _k.use_kernels():
                flat = self.flatten_parameters()
                return _k.TinyECGLogits.apply(x, flat.flat, self.num_classes, *flat.params)
        return self.forward_reference(x)


# the Module-1 scripts call the same network ``Tiny1D`` (bench_locality.py:8, train_ecg_labl:10)
class Tiny1D(TinyECG):
    def __init__(self, in_ch: int = 1, num_classes: int = 2):
        if in_ch != 1:
            raise ValueError("Tiny1D/TinyECG take single-lead windows (in_ch=1)")
        super().__init__(num_classes=num_classes)
