"""Conv1d fwd/bwd micro-benchmark sweep (BASELINE.json config 4): batch 32-4096, kernel width 3-31,
channels 16-256, L=500, bf16 — the tcgen05 kernels vs PyTorch/cuDNN (bf16 channels-first autocast-style
and channels-last) on the same GPU.  CUDA-event timing, 3 warm-ups, L2 flushed between timed calls.

    python bench/conv_microbench.py [--quick] [--out results/conv_microbench.csv]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from crossscale.fl.stats import append_results
from crossscale.ops import _lib
from crossscale.ops.conv1d import conv1d_fwd_nlc

DEV = torch.device("cuda", 0)
_flush = None


def flush_l2():
    global _flush
    if _flush is None:
        _flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=DEV)
    _flush.zero_()


def timeit(fn, iters=5, warm=3):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        flush_l2()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]


def bench_case(B, L, C, K):
    pad = K // 2
    lib, st = _lib.cuda_lib(), _lib.stream_ptr()
    x = torch.randn(B, L, C, device=DEV).to(torch.bfloat16)
    w = (torch.randn(K, C, C, device=DEV) * 0.05).to(torch.bfloat16)
    dy = torch.randn(B, L, C, device=DEV).to(torch.bfloat16)
    wt = torch.empty_like(w)
    lib.cs_conv1d_prepare_dgrad_weight(w.data_ptr(), wt.data_ptr(), K, C, C, st)
    dx = torch.empty_like(x)
    dw = torch.zeros(K, C, C, device=DEV)
    y = torch.empty_like(x)

    def ours_fwd():
        _lib.check(lib.cs_conv1d_fwd(x.data_ptr(), w.data_ptr(), None, y.data_ptr(), B, L, C, C, K, pad, 1, 0, st))

    def ours_dgrad():
        # same dispatch as ops/conv1d.py: forward weights directly where the kernel supports it, else the transposed
        # copy (prepared once outside the timed call, as before)
        rc = lib.cs_conv1d_dgrad_direct(dy.data_ptr(), w.data_ptr(), dx.data_ptr(), B, L, C, C, K, pad, 1, st)
        if rc == _lib.CS_ERR_UNSUPPORTED:
            rc = lib.cs_conv1d_dgrad(dy.data_ptr(), wt.data_ptr(), dx.data_ptr(), B, L, C, C, K, pad, 1, st)
        _lib.check(rc)

    def ours_wgrad():
        _lib.check(lib.cs_conv1d_wgrad(x.data_ptr(), dy.data_ptr(), dw.data_ptr(), None, B, L, C, C, K, pad, 1, st))

    xc = x.permute(0, 2, 1).contiguous().requires_grad_(True)          # [B,C,L] bf16 for cuDNN
    wc = w.permute(1, 2, 0).contiguous().requires_grad_(True)          # [Cout,Cin,K]
    dyc = dy.permute(0, 2, 1).contiguous()

    def cudnn_fwd():
        return F.conv1d(xc, wc, padding=pad)

    yc = cudnn_fwd()

    def cudnn_bwd():
        torch.autograd.grad(yc, (xc, wc), dyc, retain_graph=True)

    flops = 2.0 * B * L * C * C * K
    r = {"B": B, "L": L, "C": C, "K": K, "gflop_fwd": flops / 1e9}
    r["ours_fwd_ms"] = timeit(ours_fwd)
    r["ours_dgrad_ms"] = timeit(ours_dgrad)
    r["ours_wgrad_ms"] = timeit(ours_wgrad)
    r["cudnn_fwd_ms"] = timeit(cudnn_fwd)
    r["cudnn_bwd_ms"] = timeit(cudnn_bwd)
    # the memory floor of the forward pass as this harness sees it: a plain device copy of the input into the output
    # (same bytes in, same bytes out, same cold L2, same event bracket)
    r["copy_ms"] = timeit(lambda: y.copy_(x))
    r["fwd_over_copy"] = r["ours_fwd_ms"] / r["copy_ms"]
    r["ours_fwd_tflops"] = flops / r["ours_fwd_ms"] / 1e9
    r["cudnn_fwd_tflops"] = flops / r["cudnn_fwd_ms"] / 1e9
    r["fwd_speedup"] = r["cudnn_fwd_ms"] / r["ours_fwd_ms"]
    r["bwd_speedup"] = r["cudnn_bwd_ms"] / (r["ours_dgrad_ms"] + r["ours_wgrad_ms"])
    return r


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--out", default="results/conv_microbench.csv")
    ap.add_argument("--cases", default="", help='explicit "B,C,K;..." (L = 500) or "B,L,C,K;..." list')
    a = ap.parse_args()
    if a.cases:
        grid = []
        for t in a.cases.split(";"):
            v = [int(z) for z in t.split(",")]
            grid.append((v[0], 500, v[1], v[2]) if len(v) == 3 else tuple(v))
    elif a.quick:
        grid = [(256, 500, 16, 5), (256, 500, 64, 7), (256, 500, 256, 7), (1024, 500, 128, 3), (64, 500, 256, 31)]
    else:
        grid = [(B, 500, C, K) for B in (32, 256, 1024, 4096) for C in (16, 64, 256) for K in (3, 7, 31)]   # all 36 cases
    from crossscale.utils.clocks import ClockSampler
    sampler = ClockSampler(gpu_index=0).start()
    rows = []
    for case in grid:
        r = bench_case(*case)
        rows.append(r)
        print(f"B={r['B']:5d} L={r['L']:4d} C={r['C']:4d} K={r['K']:3d}: fwd ours {r['ours_fwd_ms']:.3f} ms ({r['ours_fwd_tflops']:.1f} TF) "
              f"cuDNN {r['cudnn_fwd_ms']:.3f} ms ({r['cudnn_fwd_tflops']:.1f} TF) x{r['fwd_speedup']:.2f} | "
              f"dgrad {r['ours_dgrad_ms']:.3f} wgrad {r['ours_wgrad_ms']:.3f} | bwd ours {r['ours_dgrad_ms'] + r['ours_wgrad_ms']:.3f} cuDNN {r['cudnn_bwd_ms']:.3f} x{r['bwd_speedup']:.2f} | copy {r['copy_ms']:.3f} (fwd = {r['fwd_over_copy']:.2f} x copy)", flush=True)
    if os.path.exists(a.out):
        os.remove(a.out)
    append_results(rows, a.out)
    clocks = sampler.stop()
    import json
    with open(os.path.splitext(a.out)[0] + "_clocks.json", "w") as fh:
        json.dump(clocks, fh)
    print("wrote", a.out, "clocks", clocks)


if __name__ == "__main__":
    main()
