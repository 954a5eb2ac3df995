"""ncu driver: conv1d fwd / dgrad / wgrad at B=256, L=500, C=256, K=7 (bf16)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from crossscale.ops import _lib

B, L, C, K = 256, 500, 256, 7
dev = torch.device("cuda", 0)
lib, st = _lib.cuda_lib(), _lib.stream_ptr()
x = torch.randn(B, L, C, device=dev).to(torch.bfloat16)
w = (torch.randn(K, C, C, device=dev) * 0.05).to(torch.bfloat16)
dy = torch.randn(B, L, C, device=dev).to(torch.bfloat16)
wt, y, dx = torch.empty_like(w), torch.empty_like(x), torch.empty_like(x)
dw = torch.zeros(K, C, C, device=dev)
_lib.check(lib.cs_conv1d_prepare_dgrad_weight(w.data_ptr(), wt.data_ptr(), K, C, C, st))
for it in range(4):
    _lib.check(lib.cs_conv1d_fwd(x.data_ptr(), w.data_ptr(), None, y.data_ptr(), B, L, C, C, K, K // 2, 1, 0, st))
    _lib.check(lib.cs_conv1d_dgrad_direct(dy.data_ptr(), w.data_ptr(), dx.data_ptr(), B, L, C, C, K, K // 2, 1, st))
    _lib.check(lib.cs_conv1d_wgrad(x.data_ptr(), dy.data_ptr(), dw.data_ptr(), None, B, L, C, C, K, K // 2, 1, st))
torch.cuda.synchronize()
print("ok", float(y.float().abs().mean()), float(dw.abs().mean()))
