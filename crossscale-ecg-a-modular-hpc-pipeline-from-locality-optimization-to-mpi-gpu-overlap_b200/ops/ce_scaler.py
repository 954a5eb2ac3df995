"""K5: mean cross-entropy with the AMP loss scale folded into the gradient, and the GradScaler state.

Replaces ``F.cross_entropy`` + ``GradScaler.scale`` (``part3_fedavg_overlap_mpi_gpu.py:121-125``) for models
that run through autograd; the fused TinyECG step has the same math inside its kernel.  The scaler's
``unscale_`` / inf check / ``update`` live in the SGD kernel (``ops.sgd``) so no host synchronisation is needed.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F

from . import _lib

FS_SCALE = 1


class ScaledCrossEntropy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, y, scale_t):
        logits = logits.float().contiguous()
        b, c = logits.shape
        loss = torch.empty(1, dtype=torch.float32, device=logits.device)
        dl = torch.empty_like(logits)
        _lib.check(_lib.cuda_lib().cs_cross_entropy_fwd_bwd(logits.data_ptr(), y.data_ptr(), b, c, _lib.ptr(scale_t),
                                                            loss.data_ptr(), dl.data_ptr(), _lib.stream_ptr()), "cs_cross_entropy_fwd_bwd")
        ctx.save_for_backward(dl)
        return loss[0]

    @staticmethod
    def backward(ctx, g):
        (dl,) = ctx.saved_tensors
        return dl * g, None, None


def cross_entropy(logits: torch.Tensor, y: torch.Tensor, scale: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Mean CE (unscaled value); the gradient it back-propagates is multiplied by ``scale`` (a 1-element
    device tensor, e.g. ``fstate[1:2]``).  CPU tensors fall back to PyTorch."""
    if not logits.is_cuda or logits.shape[1] > 64:
        loss = F.cross_entropy(logits.float(), y)
        if scale is None:
            return loss
        sc = scale.detach().to(loss.device).reshape(())
        return loss.detach() + (loss - loss.detach()) * sc      # value unscaled, gradient x scale
    return ScaledCrossEntropy.apply(logits, y, scale)
