"""Conv1d on the tcgen05 implicit-GEMM kernels (K1-K3).

Replaces ``nn.Conv1d`` -> cuDNN (``Module_3/tiny_ecg_model.py:17-20``) for the wide models and the conv
micro-benchmark.  Conventions:

* on CUDA, activations between layers are **channels-last bf16** ``[B, L, C]`` (what the TMA plane loads
  want); the first layer (``Cin = 1``) takes the usual ``[B, 1, L] fp32`` window.  On CPU everything is the
  stock ``[B, C, L] fp32`` PyTorch path (oracle).
* the weight parameter is stored **tap-major** ``[K, Cout, Cin]`` fp32 — the layout the kernels read and
  the weight-gradient kernel writes — so neither direction needs a permute pass; ``weight_oik`` gives the
  PyTorch ``[Cout, Cin, K]`` view.
* a trainer may attach ``weight_bf16`` (a view of the optimizer's bf16 shadow arena): then no cast kernel
  runs in forward; and it may set ``fused_wgrad = True``: backward then accumulates the weight gradient
  straight into ``weight.grad`` (the flat gradient arena) from inside the wgrad kernel's reductions
  instead of materialising a gradient tensor for autograd to add (autograd still fires the parameter's
  post-accumulate-grad hooks, so bucketed gradient sync keeps working).
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib


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
    def forward(ctx, x, weight, bias, pad, w_bf16, module, bn_req=None):
        """``bn_req``: optional dict from ``norm_act.BNReLU1d.stats_request`` — the forward kernel then also produces the
        training-mode BatchNorm statistics of its output (``bn_req["done"]`` tells the BN op whether it did)."""
        w = w_bf16 if w_bf16 is not None else weight.detach().to(torch.bfloat16)        # [K, Cout, Cin]
        xc = x.contiguous()
        ctx.save_for_backward(xc, w)
        ctx.meta = (pad, module)
        ctx.params = (weight, bias)
        b32 = None if bias is None else bias.detach().float().contiguous()
        if bn_req is not None:
            bsz, l, cin = xc.shape
            k, cout, _ = w.shape
            y = torch.empty((bsz, l + 2 * pad - k + 1, cout), dtype=torch.bfloat16, device=xc.device)
            r = bn_req
            rc = _lib.cuda_lib().cs_conv1d_fwd_bn(
                xc.data_ptr(), w.data_ptr(), _lib.ptr(b32), y.data_ptr(), bsz, l, cin, cout, k, pad, 1, 0,
                r["acc"].data_ptr(), r["ticket"].data_ptr(), r["mean"].data_ptr(), r["rstd"].data_ptr(), _lib.ptr(r["rm"]),
                _lib.ptr(r["rv"]), float(r["eps"]), float(r["momentum"]), r["gamma"].data_ptr(), r["beta"].data_ptr(),
                r["coef"].data_ptr(), _lib.stream_ptr())
            if rc == _lib.CS_OK:
                r["done"] = True
                return y
            if rc != _lib.CS_ERR_UNSUPPORTED:
                _lib.check(rc, "cs_conv1d_fwd_bn")
        return conv1d_fwd_nlc(xc, w, b32, pad)

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        pad, module = ctx.meta
        weight, bias = ctx.params
        lib, st = _lib.cuda_lib(), _lib.stream_ptr()
        dy = dy.contiguous()
        b, l, cin = x.shape
        k, cout, _ = w.shape
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(x)
            # wide layers: the data gradient reads the forward weights directly (MN-major operand, reversed taps)
            rc = lib.cs_conv1d_dgrad_direct(dy.data_ptr(), w.data_ptr(), dx.data_ptr(), b, l, cin, cout, k, pad, 1, st)
            if rc == _lib.CS_ERR_UNSUPPORTED:               # narrow layers: flipped / transposed copy + forward kernel
                wt = torch.empty((k, cin, cout), dtype=torch.bfloat16, device=x.device)
                _lib.check(lib.cs_conv1d_prepare_dgrad_weight(w.data_ptr(), wt.data_ptr(), k, cout, cin, st))
                rc = lib.cs_conv1d_dgrad(dy.data_ptr(), wt.data_ptr(), dx.data_ptr(), b, l, cin, cout, k, pad, 1, st)
            _lib.check(rc, "cs_conv1d_dgrad")
        fused = bool(getattr(module, "fused_wgrad", False)) and weight.grad is not None and weight.grad.is_contiguous() \
            and (bias is None or bias.grad is not None)
        if fused:                                           # reductions land directly in the gradient arena
            _lib.check(lib.cs_conv1d_wgrad(x.data_ptr(), dy.data_ptr(), weight.grad.data_ptr(),
                                           _lib.ptr(None if bias is None else bias.grad), b, l, cin, cout, k, pad, 1, st),
                       "cs_conv1d_wgrad")
            # autograd still runs the parameter's post-accumulate-grad hooks (e.g. GradSync buckets) after this
            # node returns, even though no gradient tensor is handed back
            return dx, None, None, None, None, None, None
        dw = torch.zeros((k, cout, cin), dtype=torch.float32, device=x.device)
        db = torch.zeros(cout, dtype=torch.float32, device=x.device) if bias is not None else None
        _lib.check(lib.cs_conv1d_wgrad(x.data_ptr(), dy.data_ptr(), dw.data_ptr(), _lib.ptr(db), b, l, cin, cout, k, pad, 1, st),
                   "cs_conv1d_wgrad")
        return dx, dw, db, None, None, None, None


class _ConvCin1(torch.autograd.Function):
    """first layer: x [B,1,L] fp32 -> y [B,L,Cout] bf16 (FFMA kernels); weight [K, Cout, 1]"""

    @staticmethod
    def forward(ctx, x, weight, bias, pad):
        b, _, l = x.shape
        k, cout, _ = weight.shape
        x2 = x.reshape(b, l).contiguous().float()
        w2 = weight.detach().reshape(k, cout).t().contiguous().float()                    # kernel wants [Cout][K]
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
        return None, dw.t().contiguous().view(k, cout, 1), db, None


class Conv1dB200(nn.Module):
    def __init__(self, cin: int, cout: int, k: int, padding: int = 0, stride: int = 1, bias: bool = True):
        super().__init__()
        if stride != 1:
            raise ValueError("Conv1dB200 is stride-1 (down-sample with pooling; see models/resnet1d_wide.py)")
        self.cin, self.cout, self.k, self.padding = cin, cout, k, padding
        w = torch.empty(cout, cin, k)
        nn.init.kaiming_uniform_(w, a=math.sqrt(5))
        self.weight = nn.Parameter(w.permute(2, 0, 1).contiguous())                        # [K, Cout, Cin]
        self.bias = None
        if bias:
            bound = 1.0 / math.sqrt(cin * k)
            self.bias = nn.Parameter(torch.empty(cout).uniform_(-bound, bound))
        self.weight_bf16 = None          # optional [K,Cout,Cin] bf16 view kept fresh by the optimizer (shadow arena)
        self.fused_wgrad = False         # accumulate dW straight into weight.grad inside the kernel

    @property
    def weight_oik(self) -> torch.Tensor:
        """PyTorch-layout view ``[Cout, Cin, K]`` of the tap-major parameter."""
        return self.weight.permute(1, 2, 0)

    def forward(self, x: torch.Tensor, bn_req=None) -> torch.Tensor:
        if not x.is_cuda:
            return F.conv1d(x, self.weight_oik, self.bias, padding=self.padding)
        if self.cin == 1:
            return _ConvCin1.apply(x, self.weight, self.bias, self.padding)
        return _ConvNLC.apply(x, self.weight, self.bias, self.padding, self.weight_bf16, self, bn_req)
