"""Fused BatchNorm(+residual)(+ReLU), pooling (K4) on channels-last bf16 activations.

CUDA path: two memory-bound kernels forward (per-channel statistics, then normalise + affine +
residual + ReLU in one pass) and two backward; CPU path: stock PyTorch on ``[B, C, L]`` (oracle).
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib


class _BNAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, residual, relu, eps, mean_in, var_in, running, momentum):
        lib, st = _lib.cuda_lib(), _lib.stream_ptr()
        x = x.contiguous()
        b, l, c = x.shape
        rows = b * l
        if mean_in is None:                                # training: batch statistics + running update, 2 launches
            buf = torch.zeros(4 * c, dtype=torch.float32, device=x.device)
            stats, mean, rstd = buf[:2 * c], buf[2 * c:3 * c], buf[3 * c:]
            _lib.check(lib.cs_bn_stats(x.data_ptr(), stats.data_ptr(), rows, c, st), "cs_bn_stats")
            rm, rv = running if running is not None else (None, None)
            _lib.check(lib.cs_bn_finalize(stats.data_ptr(), mean.data_ptr(), rstd.data_ptr(), _lib.ptr(rm), _lib.ptr(rv), c, rows,
                                          float(eps), float(momentum), st), "cs_bn_finalize")
        else:
            mean = mean_in.float().contiguous()
            rstd = torch.rsqrt(var_in.float() + eps)
        res = None if residual is None else residual.contiguous()
        y = torch.empty_like(x)
        g32, b32 = gamma.detach().float().contiguous(), beta.detach().float().contiguous()
        _lib.check(lib.cs_bn_apply(x.data_ptr(), _lib.ptr(res), y.data_ptr(), mean.contiguous().data_ptr(),
                                   rstd.contiguous().data_ptr(), g32.data_ptr(), b32.data_ptr(), rows, c, int(relu), st),
                   "cs_bn_apply")
        ctx.save_for_backward(x, y, mean, rstd, g32)
        ctx.meta = (relu, residual is not None, mean_in is None)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, y, mean, rstd, gamma = ctx.saved_tensors
        relu, has_res, training = ctx.meta
        lib, st = _lib.cuda_lib(), _lib.stream_ptr()
        b, l, c = x.shape
        rows = b * l
        dy = dy.contiguous()
        dgb = torch.zeros(2 * c, dtype=torch.float32, device=x.device)
        dgamma, dbeta = dgb[:c], dgb[c:]
        _lib.check(lib.cs_bn_bwd_reduce(x.data_ptr(), y.data_ptr(), dy.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                                        dgamma.data_ptr(), dbeta.data_ptr(), rows, c, int(relu), st), "cs_bn_bwd_reduce")
        dx = torch.empty_like(x)
        dres = torch.empty_like(x) if has_res else None
        # eval mode (fixed statistics): the correction terms vanish -> pass zeros for the reductions
        zg = dgamma if training else torch.zeros_like(dgamma)
        zb = dbeta if training else torch.zeros_like(dbeta)
        _lib.check(lib.cs_bn_bwd_apply(x.data_ptr(), y.data_ptr(), dy.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                                       gamma.data_ptr(), zg.data_ptr(), zb.data_ptr(), dx.data_ptr(), _lib.ptr(dres),
                                       rows, c, int(relu), st), "cs_bn_bwd_apply")
        return dx, dgamma, dbeta, dres, None, None, None, None, None, None


class BNReLU1d(nn.Module):
    """BatchNorm1d (+ residual add) (+ ReLU) in one module."""

    def __init__(self, channels: int, relu: bool = True, eps: float = 1e-5, momentum: float = 0.1):
        super().__init__()
        self.relu, self.eps, self.momentum = relu, eps, momentum
        self.weight = nn.Parameter(torch.ones(channels))
        self.bias = nn.Parameter(torch.zeros(channels))
        self.register_buffer("running_mean", torch.zeros(channels))
        self.register_buffer("running_var", torch.ones(channels))

    def forward(self, x: torch.Tensor, residual: Optional[torch.Tensor] = None) -> torch.Tensor:
        if not x.is_cuda:
            y = F.batch_norm(x, self.running_mean, self.running_var, self.weight, self.bias, self.training,
                             self.momentum, self.eps)
            if residual is not None:
                y = y + residual
            return F.relu(y) if self.relu else y
        if self.training:
            return _BNAct.apply(x, self.weight, self.bias, residual, self.relu, self.eps, None, None,
                                (self.running_mean, self.running_var), self.momentum)
        return _BNAct.apply(x, self.weight, self.bias, residual, self.relu, self.eps, self.running_mean, self.running_var,
                            None, self.momentum)


class _GlobalAvgPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        x = x.contiguous()
        b, l, c = x.shape
        out = torch.empty((b, c), dtype=torch.float32, device=x.device)
        _lib.check(_lib.cuda_lib().cs_avgpool_fwd(x.data_ptr(), out.data_ptr(), b, l, c, _lib.stream_ptr()), "cs_avgpool_fwd")
        ctx.shape = (b, l, c)
        return out

    @staticmethod
    def backward(ctx, dout):
        b, l, c = ctx.shape
        dx = torch.empty((b, l, c), dtype=torch.bfloat16, device=dout.device)
        _lib.check(_lib.cuda_lib().cs_avgpool_bwd(dout.contiguous().float().data_ptr(), dx.data_ptr(), b, l, c, _lib.stream_ptr()),
                   "cs_avgpool_bwd")
        return dx


def global_avg_pool1d(x: torch.Tensor) -> torch.Tensor:
    """[B,L,C] bf16 (CUDA) or [B,C,L] fp32 (CPU) -> [B,C] fp32"""
    if x.is_cuda:
        return _GlobalAvgPool.apply(x)
    return x.mean(dim=-1)


class _AvgPool2(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        x = x.contiguous()
        b, l, c = x.shape
        y = torch.empty((b, l // 2, c), dtype=torch.bfloat16, device=x.device)
        _lib.check(_lib.cuda_lib().cs_avgpool2_fwd(x.data_ptr(), y.data_ptr(), b, l, c, _lib.stream_ptr()), "cs_avgpool2_fwd")
        ctx.shape = (b, l, c)
        return y

    @staticmethod
    def backward(ctx, dy):
        b, l, c = ctx.shape
        dx = torch.empty((b, l, c), dtype=torch.bfloat16, device=dy.device)
        _lib.check(_lib.cuda_lib().cs_avgpool2_bwd(dy.contiguous().data_ptr(), dx.data_ptr(), b, l, c, _lib.stream_ptr()),
                   "cs_avgpool2_bwd")
        return dx


def avg_pool2(x: torch.Tensor) -> torch.Tensor:
    """halve the length (mean of position pairs, odd tail dropped); layout-aware"""
    if x.is_cuda:
        return _AvgPool2.apply(x)
    return F.avg_pool1d(x, 2)
