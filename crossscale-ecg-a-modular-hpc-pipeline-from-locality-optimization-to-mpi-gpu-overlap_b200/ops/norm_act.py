"""Fused BatchNorm(+residual)(+ReLU), pooling (K4) on channels-last bf16 activations.

No counterpart in the reference (its only model, ``Module_3/tiny_ecg_model.py:8-29``, has no normalisation layer); this
is the "fused BN/ReLU/pool" of the north-star ResNet-1D config.  CUDA path: the training statistics come out of the
producing convolution's epilogue (``conv_bn``) or one statistics kernel, then one pass does normalise + affine +
residual + ReLU; backward is one reduction and one elementwise pass; all reductions finish inside their own kernel
through a persistent per-layer workspace.  CPU path: stock PyTorch on ``[B, C, L]`` (oracle).
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib


class _BNAct(torch.autograd.Function):
    """``ws`` is the layer's persistent workspace (fp32 [4C] accumulators + int32 ticket, see ``BNReLU1d``): the
    reductions finish inside their own kernel (last block) and leave it zero again, so a training step runs
    2 launches forward and 2 backward per layer with no fill / finalize / add kernels."""

    @staticmethod
    def forward(ctx, x, gamma, beta, residual, relu, eps, mean_in, var_in, running, momentum, ws, module, stats=None):
        lib, st = _lib.cuda_lib(), _lib.stream_ptr()
        x = x.contiguous()
        b, l, c = x.shape
        rows = b * l
        if stats is not None:                              # produced by the convolution that wrote x (fused epilogue)
            g32, mean, rstd, coef = stats["gamma"], stats["mean"], stats["rstd"], stats["coef"]
        else:
            g32, b32 = gamma.detach().float().contiguous(), beta.detach().float().contiguous()
            buf = torch.empty(4 * c, dtype=torch.float32, device=x.device)
            mean, rstd, coef = buf[:c], buf[c:2 * c], buf[2 * c:]
        if stats is not None:
            pass
        elif mean_in is None:                              # training: batch statistics + running update in one launch
            rm, rv = running if running is not None else (None, None)
            acc, ticket = ws
            _lib.check(lib.cs_bn_stats_finalize(x.data_ptr(), acc.data_ptr(), ticket.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                                                _lib.ptr(rm), _lib.ptr(rv), rows, c, float(eps), float(momentum),
                                                g32.data_ptr(), b32.data_ptr(), coef.data_ptr(), st), "cs_bn_stats_finalize")
        else:
            mean.copy_(mean_in.float())
            rstd.copy_(torch.rsqrt(var_in.float() + eps))
            coef[:c] = rstd * g32
            coef[c:] = b32 - mean * coef[:c]
        res = None if residual is None else residual.contiguous()
        y = torch.empty_like(x)
        _lib.check(lib.cs_bn_apply_coef(x.data_ptr(), _lib.ptr(res), y.data_ptr(), coef.data_ptr(), rows, c, int(relu), st),
                   "cs_bn_apply_coef")
        ctx.save_for_backward(x, y, mean, rstd, g32)
        ctx.meta = (relu, residual is not None, mean_in is None, ws, module)
        ctx.params = (gamma, beta)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, y, mean, rstd, g32 = ctx.saved_tensors
        relu, has_res, training, ws, module = ctx.meta
        gamma, beta = ctx.params
        lib, st = _lib.cuda_lib(), _lib.stream_ptr()
        b, l, c = x.shape
        rows = b * l
        dy = dy.contiguous()
        buf = torch.empty(5 * c, dtype=torch.float32, device=x.device)
        dgamma, dbeta, coef = buf[:c], buf[c:2 * c], buf[2 * c:]
        # a trainer that owns a flat gradient arena lets the kernel add the sums straight into gamma.grad / beta.grad
        # (autograd still fires the parameters' post-accumulate-grad hooks when None is returned for them)
        fused = bool(getattr(module, "fused_grad", False)) and gamma.grad is not None and beta.grad is not None \
            and gamma.grad.is_contiguous() and beta.grad.is_contiguous() and gamma.grad.dtype == torch.float32
        acc, ticket = ws
        _lib.check(lib.cs_bn_bwd_reduce_ws(x.data_ptr(), y.data_ptr(), dy.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                                           acc[2 * c:].data_ptr(), ticket[1:].data_ptr(), dgamma.data_ptr(), dbeta.data_ptr(),
                                           gamma.grad.data_ptr() if fused else None, beta.grad.data_ptr() if fused else None,
                                           g32.data_ptr(), coef.data_ptr(), rows, c, int(relu), st), "cs_bn_bwd_reduce_ws")
        if not training:                                    # fixed statistics: the batch-correction terms vanish
            coef[c:].zero_()
        dx = torch.empty_like(x)
        dres = torch.empty_like(x) if has_res else None
        _lib.check(lib.cs_bn_bwd_apply_coef(x.data_ptr(), y.data_ptr(), dy.data_ptr(), coef.data_ptr(), dx.data_ptr(),
                                            _lib.ptr(dres), rows, c, int(relu), st), "cs_bn_bwd_apply_coef")
        if fused:
            return dx, None, None, dres, None, None, None, None, None, None, None, None, None
        return dx, dgamma, dbeta, dres, None, None, None, None, None, None, None, None, None


class BNReLU1d(nn.Module):
    """BatchNorm1d (+ residual add) (+ ReLU) in one module."""

    def __init__(self, channels: int, relu: bool = True, eps: float = 1e-5, momentum: float = 0.1):
        super().__init__()
        self.relu, self.eps, self.momentum = relu, eps, momentum
        self.weight = nn.Parameter(torch.ones(channels))
        self.bias = nn.Parameter(torch.zeros(channels))
        self.register_buffer("running_mean", torch.zeros(channels))
        self.register_buffer("running_var", torch.ones(channels))
        self.fused_grad = False          # accumulate dgamma / dbeta into .grad inside the reduction kernel
        self._ws = None                  # (fp32 [4C] accumulators, int32 [2] tickets): zero, and left zero by every kernel

    def _workspace(self, device):
        if self._ws is None or self._ws[0].device != device:
            c = self.weight.numel()
            self._ws = (torch.zeros(4 * c, dtype=torch.float32, device=device), torch.zeros(2, dtype=torch.int32, device=device))
        return self._ws

    def stats_request(self, device) -> dict:
        """What a producing convolution needs to compute this layer's training statistics in its own epilogue
        (``Conv1dB200.forward(x, bn_req=...)``); pass the dict back to :meth:`forward` as ``stats``."""
        c = self.weight.numel()
        acc, ticket = self._workspace(device)
        buf = torch.empty(4 * c, dtype=torch.float32, device=device)
        return {"acc": acc[:2 * c], "ticket": ticket[:1], "mean": buf[:c], "rstd": buf[c:2 * c], "coef": buf[2 * c:],
                "rm": self.running_mean, "rv": self.running_var, "eps": self.eps, "momentum": self.momentum,
                "gamma": self.weight.detach().float().contiguous(), "beta": self.bias.detach().float().contiguous(), "done": False}

    def forward(self, x: torch.Tensor, residual: Optional[torch.Tensor] = None, stats: Optional[dict] = None) -> torch.Tensor:
        if not x.is_cuda:
            y = F.batch_norm(x, self.running_mean, self.running_var, self.weight, self.bias, self.training,
                             self.momentum, self.eps)
            if residual is not None:
                y = y + residual
            return F.relu(y) if self.relu else y
        ws = self._workspace(x.device)
        if self.training:
            return _BNAct.apply(x, self.weight, self.bias, residual, self.relu, self.eps, None, None,
                                (self.running_mean, self.running_var), self.momentum, ws, self,
                                stats if (stats is not None and stats.get("done")) else None)
        return _BNAct.apply(x, self.weight, self.bias, residual, self.relu, self.eps, self.running_mean, self.running_var,
                            None, self.momentum, ws, self)


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


def conv_bn(conv, bn: "BNReLU1d", x: torch.Tensor, residual: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``bn(conv(x), residual)``; in training on CUDA the convolution's epilogue also produces the BN statistics of
    its output (wide layers), so the separate statistics pass over the activation disappears."""
    if x.is_cuda and bn.training and conv.cin > 1 and torch.is_grad_enabled():
        req = bn.stats_request(x.device)
        return bn(conv(x, bn_req=req), residual, stats=req)
    return bn(conv(x), residual)


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
