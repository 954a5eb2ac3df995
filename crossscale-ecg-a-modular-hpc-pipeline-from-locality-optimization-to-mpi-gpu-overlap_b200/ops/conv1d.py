"""Conv1d on the tcgen05 implicit-GEMM kernels (K1-K3).

Replaces ``nn.Conv1d`` -> cuDNN (``Module_3/tiny_ecg_model.py:17-20``) for the wide models and the conv
micro-benchmark.  Convention: on CUDA, activations between layers are **channels-last bf16**
``[B, L, C]`` (what the TMA plane loads want); the first layer (``Cin = 1``) takes the usual
``[B, 1, L] fp32`` window.  On CPU everything is the stock ``[B, C, L] fp32`` PyTorch path (oracle).
Parameters keep the PyTorch ``[Cout, Cin, K]`` fp32 layout so state_dicts look familiar.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib


def _use_kernels(x: torch.Tensor) -> bool:
    return x.is_cuda


def conv1d_fwd_nlc(x, w_kcoci, bias, pad, relu=False):
    """x [B,L,Cin] bf16, w [K,Cout,Cin] bf16 -> y [B,Lout,Cout] bf16"""
    b, l, cin = x.shape
    k, cout, _ = w_kcoci.shape
    lout = l + 2 * pad - k + 1
    y = torch.empty((b, lout, cout), dtype=torch.bfloat16, device=x.device)
    _lib.check(_lib.cuda_lib().cs_conv1d_fwd(x.data_ptr(), w_kcoci.data_ptr(), _lib.ptr(bias), y.data_ptr(), b, l, cin,
                                             cout, k, pad, 1, int(relu), _lib.stream_ptr()), "cs_conv1d_fwd")
    return y


class _ConvNLC(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, pad):
        w = weight.detach().permute(2, 0, 1).contiguous().to(torch.bfloat16)        # [K, Cout, Cin]
        xc = x.contiguous()
        ctx.save_for_backward(xc, w)
        ctx.meta = (pad, bias is not None, weight.shape)
        return conv1d_fwd_nlc(xc, w, None if bias is None else bias.detach().float().contiguous(), pad)

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        pad, has_bias, wshape = ctx.meta
        lib, st = _lib.cuda_lib(), _lib.stream_ptr()
        dy = dy.contiguous()
        b, l, cin = x.shape
        k, cout, _ = w.shape
        dx = None
        if ctx.needs_input_grad[0]:
            wt = torch.empty((k, cin, cout), dtype=torch.bfloat16, device=x.device)
            _lib.check(lib.cs_conv1d_prepare_dgrad_weight(w.data_ptr(), wt.data_ptr(), k, cout, cin, st))
            dx = torch.empty_like(x)
            _lib.check(lib.cs_conv1d_dgrad(dy.data_ptr(), wt.data_ptr(), dx.data_ptr(), b, l, cin, cout, k, pad, 1, st),
                       "cs_conv1d_dgrad")
        dw = torch.zeros((k, cout, cin), dtype=torch.float32, device=x.device)
        db = torch.zeros(cout, dtype=torch.float32, device=x.device) if has_bias else None
        _lib.check(lib.cs_conv1d_wgrad(x.data_ptr(), dy.data_ptr(), dw.data_ptr(), _lib.ptr(db), b, l, cin, cout, k, pad, 1, st),
                   "cs_conv1d_wgrad")
        return dx, dw.permute(1, 2, 0).contiguous(), db, None


class _ConvCin1(torch.autograd.Function):
    """first layer: x [B,1,L] fp32 -> y [B,L,Cout] bf16 (FFMA kernels)"""

    @staticmethod
    def forward(ctx, x, weight, bias, pad):
        b, _, l = x.shape
        cout, _, k = weight.shape
        x2 = x.reshape(b, l).contiguous().float()
        w2 = weight.detach().reshape(cout, k).contiguous().float()
        lout = l + 2 * pad - k + 1
        y = torch.empty((b, lout, cout), dtype=torch.bfloat16, device=x.device)
        _lib.check(_lib.cuda_lib().cs_conv1d_cin1_fwd(x2.data_ptr(), w2.data_ptr(),
                                                      _lib.ptr(None if bias is None else bias.detach().float().contiguous()),
                                                      y.data_ptr(), b, l, cout, k, pad, 1, 0, _lib.stream_ptr()), "cs_conv1d_cin1_fwd")
        ctx.save_for_backward(x2)
        ctx.meta = (pad, bias is not None, cout, k)
        return y

    @staticmethod
    def backward(ctx, dy):
        (x2,) = ctx.saved_tensors
        pad, has_bias, cout, k = ctx.meta
        b, l = x2.shape
        dw = torch.zeros((cout, k), dtype=torch.float32, device=x2.device)
        db = torch.zeros(cout, dtype=torch.float32, device=x2.device) if has_bias else None
        _lib.check(_lib.cuda_lib().cs_conv1d_cin1_wgrad(x2.data_ptr(), dy.contiguous().data_ptr(), dw.data_ptr(), _lib.ptr(db),
                                                        b, l, cout, k, pad, 1, _lib.stream_ptr()), "cs_conv1d_cin1_wgrad")
        return None, dw.view(cout, 1, k), db, None


class Conv1dB200(nn.Module):
    def __init__(self, cin: int, cout: int, k: int, padding: int = 0, stride: int = 1, bias: bool = True):
        super().__init__()
        if stride != 1:
            raise ValueError("Conv1dB200 is stride-1 (down-sample with pooling; see models/resnet1d_wide.py)")
        self.cin, self.cout, self.k, self.padding = cin, cout, k, padding
        self.weight = nn.Parameter(torch.empty(cout, cin, k))
        self.bias = nn.Parameter(torch.zeros(cout)) if bias else None
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        if bias:
            bound = 1.0 / math.sqrt(cin * k)
            nn.init.uniform_(self.bias, -bound, bound)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if not _use_kernels(x):
            return F.conv1d(x, self.weight, self.bias, padding=self.padding)
        if self.cin == 1:
            return _ConvCin1.apply(x, self.weight, self.bias, self.padding)
        return _ConvNLC.apply(x, self.weight, self.bias, self.padding)
