"""Native conv1d kernel vs PyTorch — ``python -m crossscale.cli.benchmark_part_2``.

Protocol of ``Module_2/benchmark_part_2.py``: grid ``B in {64,128,256,512} x K in {3,5,7}``, ``L=500``,
all threads, 15 trials, each trial = 3 warm-up calls + 1 timed call (``:12-19,61-67``); reports
median/mean/pstdev/p95 ms, samples/s and ``speedup_med = median(torch)/median(native)`` (``:92-108``) into
``results/part2_openmp_results.csv`` and ``..._raw.csv``.  Arms: ``torch`` = ``nn.Conv1d(1,1,K,bias=False)``
forward on CPU, ``omp`` = this repo's native kernel (fresh output allocation inside the timed call, like
the reference's ``run_omp_conv``), optional ``ref`` = the UNMODIFIED reference C kernel built from
``baseline/_ref`` or ``/root/reference``, optional ``gpu`` = the sm_100a conv1d kernels (``--gpu``).
"""
from __future__ import annotations

import argparse
import ctypes
import os
import statistics
import subprocess
import tempfile
import time

import numpy as np
import torch

from ..fl.stats import append_results
from ..ops import cpu_conv

BATCH_SIZES = [64, 128, 256, 512]
KERNEL_SIZES = [3, 5, 7]
L = 500
TRIALS = 15
WARMUP_STEPS = 3


def time_once(fn) -> float:
    for _ in range(WARMUP_STEPS):
        fn()
    t0 = time.perf_counter()
    fn()
    return (time.perf_counter() - t0) * 1e3


def _p95(v):
    s = sorted(v)
    return s[min(len(s) - 1, int(round(0.95 * (len(s) - 1))))]


def _stats(prefix, v):
    return {f"{prefix}_ms_median": statistics.median(v), f"{prefix}_ms_mean": statistics.fmean(v),
            f"{prefix}_ms_std": statistics.pstdev(v), f"{prefix}_ms_p95": _p95(v)}


def load_reference_kernel():
    """Build the unmodified reference C kernel (CPU oracle + timed CPU baseline), or ``None``."""
    here = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    for root in (os.path.join(here, "baseline", "_ref"), "/root/reference"):
        src = os.path.join(root, "Module_2", "conv1d_openmp_simd.c")
        if os.path.exists(src):
            so = os.path.join(tempfile.gettempdir(), f"cs_ref_conv1d_{os.getuid()}.so")
            cc = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else "gcc"
            r = subprocess.run([cc, "-O2", "-fopenmp", "-mavx2", "-mfma", "-shared", "-fPIC", src, "-o", so],
                               capture_output=True)
            if r.returncode != 0:
                return None
            lib = ctypes.CDLL(so)
            lib.conv1d_batch_omp_simd.argtypes = [ctypes.c_void_p] * 3 + [ctypes.c_int] * 4
            return lib
    return None


def gpu_arm(x: np.ndarray, w: np.ndarray, trials: int):
    """The same single-channel valid convolution on the GPU (``cs_conv1d_batch_valid``).  Returns (median ms end to end
    from pinned host memory incl. H2D of x / D2H of y - the host-array contract of the reference's ctypes call -,
    median ms of the kernel alone, max abs difference to the CPU kernel)."""
    from ..ops import _lib
    lib = _lib.cuda_lib()
    dev = torch.device("cuda", 0)
    bs, ln = x.shape
    k = w.shape[0]
    hx = torch.from_numpy(x).pin_memory()
    hy = torch.empty((bs, ln - k + 1), dtype=torch.float32).pin_memory()
    dx, dy = torch.empty_like(hx, device=dev), torch.empty_like(hy, device=dev)
    dw = torch.from_numpy(w).to(dev)
    st = _lib.stream_ptr()

    def e2e():
        dx.copy_(hx, non_blocking=True)
        _lib.check(lib.cs_conv1d_batch_valid(dx.data_ptr(), dw.data_ptr(), dy.data_ptr(), bs, ln, k, st), "cs_conv1d_batch_valid")
        hy.copy_(dy, non_blocking=True)
        torch.cuda.synchronize()

    t_e2e, t_dev = [], []
    for _ in range(trials):
        t_e2e.append(time_once(e2e))
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        _lib.check(lib.cs_conv1d_batch_valid(dx.data_ptr(), dw.data_ptr(), dy.data_ptr(), bs, ln, k, st), "cs_conv1d_batch_valid")
        b.record()
        torch.cuda.synchronize()
        t_dev.append(a.elapsed_time(b))
    err = float(np.abs(hy.numpy() - cpu_conv.conv1d_batch(x, w, 1)).max())
    return statistics.median(t_e2e), statistics.median(t_dev), err


def bench_pair(bs: int, k: int, nthreads: int, rng, ref_lib=None, trials: int = TRIALS, gpu: bool = False):
    x = rng.normal(size=(bs, L)).astype(np.float32)
    w = rng.normal(size=(k,)).astype(np.float32)
    conv = torch.nn.Conv1d(1, 1, k, bias=False)
    with torch.no_grad():
        conv.weight.copy_(torch.from_numpy(w).view(1, 1, k))
    xt = torch.from_numpy(x).unsqueeze(1)
    t_torch, t_omp, t_ref, raw = [], [], [], []
    for trial in range(trials):
        a = time_once(lambda: conv(xt))
        b = time_once(lambda: cpu_conv.conv1d_batch(x, w, nthreads))
        row = {"batch_size": bs, "kernel_size": k, "trial": trial, "torch_ms": a, "omp_ms": b}
        t_torch.append(a)
        t_omp.append(b)
        if ref_lib is not None:
            def ref_call():
                y = np.zeros((bs, L - k + 1), np.float32)
                ref_lib.conv1d_batch_omp_simd(x.ctypes.data, w.ctypes.data, y.ctypes.data, bs, L, k, nthreads)
            c = time_once(ref_call)
            t_ref.append(c)
            row["ref_ms"] = c
        raw.append(row)
    out = {"batch_size": bs, "kernel_size": k, "nthreads": nthreads}
    out.update(_stats("torch", t_torch))
    out.update(_stats("omp", t_omp))
    out["torch_sps"] = bs / (out["torch_ms_median"] / 1e3)
    out["omp_sps"] = bs / (out["omp_ms_median"] / 1e3)
    out["speedup_med"] = out["torch_ms_median"] / out["omp_ms_median"]
    if t_ref:
        out.update(_stats("ref", t_ref))
        out["ref_speedup_med"] = out["torch_ms_median"] / out["ref_ms_median"]
    if gpu:
        g_e2e, g_dev, err = gpu_arm(x, w, trials)
        out["gpu_e2e_ms_median"], out["gpu_kernel_ms_median"], out["gpu_max_abs_err"] = g_e2e, g_dev, err
        out["gpu_e2e_speedup_vs_torch_cpu"] = out["torch_ms_median"] / g_e2e
        if t_ref:
            out["gpu_e2e_speedup_vs_reference_c"] = out["ref_ms_median"] / g_e2e
            out["gpu_kernel_speedup_vs_reference_c"] = out["ref_ms_median"] / g_dev
    return out, raw


def safe_write_csv(rows, path):
    if os.path.exists(path):
        try:
            os.remove(path)
        except PermissionError:
            pass
    return append_results(rows, path)


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--results-dir", default="results")
    ap.add_argument("--trials", type=int, default=TRIALS)
    ap.add_argument("--batch-sizes", nargs="+", type=int, default=BATCH_SIZES)
    ap.add_argument("--kernel-sizes", nargs="+", type=int, default=KERNEL_SIZES)
    ap.add_argument("--no-ref", action="store_true", help="skip the unmodified reference C kernel arm")
    ap.add_argument("--gpu", action="store_true", help="add the GPU arm (cs_conv1d_batch_valid on cuda:0)")
    args = ap.parse_args(argv)
    nthreads = os.cpu_count() or 1
    torch.set_num_threads(nthreads)
    try:
        torch.set_num_interop_threads(max(1, nthreads // 2))
    except RuntimeError:
        pass
    ref_lib = None if args.no_ref else load_reference_kernel()
    rng = np.random.default_rng(1337)
    rows, raws = [], []
    for bs in args.batch_sizes:
        for k in args.kernel_sizes:
            r, raw = bench_pair(bs, k, nthreads, rng, ref_lib, args.trials, gpu=args.gpu and torch.cuda.is_available())
            rows.append(r)
            raws.extend(raw)
            extra = f" | reference C kernel {r['ref_speedup_med']:.2f}x" if "ref_speedup_med" in r else ""
            print(f"B={bs:4d} K={k}: torch {r['torch_ms_median']:.3f} ms, native {r['omp_ms_median']:.3f} ms "
                  f"-> {r['speedup_med']:.2f}x{extra}")
    p1 = safe_write_csv(rows, os.path.join(args.results_dir, "part2_openmp_results.csv"))
    p2 = safe_write_csv(raws, os.path.join(args.results_dir, "part2_openmp_results_raw.csv"))
    print(f"[OK] wrote {p1} and {p2} (ISA path {cpu_conv.isa()}, {nthreads} threads)")
    from . import plots
    plots.plot_part2(args.results_dir)
    return rows


if __name__ == "__main__":
    main()
