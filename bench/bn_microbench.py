"""BN kernel micro-benchmark on L2-resident ResNet-1D activations (B=64): device time per launch of the four
BN kernels, CUDA events around 20 back-to-back launches (no L2 flush: inside a training step these tensors were
just written by the previous kernel)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from crossscale.ops import _lib

dev = torch.device("cuda", 0)
lib, st = _lib.cuda_lib(), _lib.stream_ptr()


def t(fn, n=20):
    for _ in range(3):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3


for B, L, C in [(64, 360, 128), (64, 180, 256), (64, 90, 512), (64, 45, 768), (256, 360, 128)]:
    rows = B * L
    x = torch.randn(rows, C, device=dev).to(torch.bfloat16)
    y, dy, dx, res = torch.empty_like(x), torch.randn_like(x), torch.empty_like(x), torch.randn_like(x)
    acc = torch.zeros(4 * C, device=dev); ticket = torch.zeros(2, dtype=torch.int32, device=dev)
    buf = torch.zeros(9 * C, device=dev)
    mean, rstd, coef, dg, db, coefb = buf[:C], buf[C:2 * C], buf[2 * C:4 * C], buf[4 * C:5 * C], buf[5 * C:6 * C], buf[6 * C:]
    g, bt = torch.ones(C, device=dev), torch.zeros(C, device=dev)
    r = {}
    r["stats+finalize"] = t(lambda: lib.cs_bn_stats_finalize(x.data_ptr(), acc.data_ptr(), ticket.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                                                             None, None, rows, C, 1e-5, 0.1, g.data_ptr(), bt.data_ptr(), coef.data_ptr(), st))
    r["apply"] = t(lambda: lib.cs_bn_apply_coef(x.data_ptr(), res.data_ptr(), y.data_ptr(), coef.data_ptr(), rows, C, 1, st))
    r["bwd_reduce"] = t(lambda: lib.cs_bn_bwd_reduce_ws(x.data_ptr(), y.data_ptr(), dy.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                                                        acc[2 * C:].data_ptr(), ticket[1:].data_ptr(), dg.data_ptr(), db.data_ptr(), None, None,
                                                        g.data_ptr(), coefb.data_ptr(), rows, C, 1, st))
    r["bwd_apply"] = t(lambda: lib.cs_bn_bwd_apply_coef(x.data_ptr(), y.data_ptr(), dy.data_ptr(), coefb.data_ptr(), dx.data_ptr(),
                                                        res.data_ptr(), rows, C, 1, st))
    mb = rows * C * 2 / 1e6
    print(f"B={B} L={L} C={C} ({mb:.1f} MB/tensor): " + "  ".join(f"{k} {v:.1f} us" for k, v in r.items()), flush=True)
