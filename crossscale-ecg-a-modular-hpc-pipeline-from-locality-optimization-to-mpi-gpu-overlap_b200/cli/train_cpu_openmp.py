"""OpenMP thread-scaling benchmark of the native conv1d kernel — ``python -m crossscale.cli.train_cpu_openmp``.

``Module_2/train_cpu_openmp.py:19-69``: ``threads in {1,2,4,8,16} x batch in {64,128,256,512}``, ``L=500``,
``K=32``, 5 warm-up + 50 timed calls; writes ``results/part2_openmp_simd_results.csv`` with columns
``threads,batch,compute_ms,samples_per_s`` (forward only, despite the name).
"""
from __future__ import annotations

import argparse
import os
import time

import numpy as np

from ..fl.stats import append_results
from ..ops import cpu_conv

THREADS = [1, 2, 4, 8, 16]
BATCHES = [64, 128, 256, 512]
L, K = 500, 32


def run(threads: int, batch: int, iters: int = 50, warmup: int = 5, rng=None) -> dict:
    rng = rng or np.random.default_rng(1337)
    x = rng.normal(size=(batch, L)).astype(np.float32)
    w = rng.normal(size=(K,)).astype(np.float32)
    y = np.zeros((batch, L - K + 1), np.float32)
    for _ in range(warmup):
        cpu_conv.conv1d_batch(x, w, threads, out=y)
    t0 = time.perf_counter()
    for _ in range(iters):
        cpu_conv.conv1d_batch(x, w, threads, out=y)
    ms = (time.perf_counter() - t0) * 1e3 / iters
    return {"threads": threads, "batch": batch, "compute_ms": ms, "samples_per_s": batch / (ms / 1e3)}


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--results-dir", default="results")
    ap.add_argument("--iters", type=int, default=50)
    ap.add_argument("--threads", nargs="+", type=int, default=THREADS)
    ap.add_argument("--batches", nargs="+", type=int, default=BATCHES)
    args = ap.parse_args(argv)
    rows = [run(t, b, args.iters) for t in args.threads for b in args.batches]
    for r in rows:
        print(f"threads={r['threads']:2d} batch={r['batch']:4d}: {r['compute_ms']:.3f} ms  {r['samples_per_s']:,.0f} samples/s")
    path = os.path.join(args.results_dir, "part2_openmp_simd_results.csv")
    if os.path.exists(path):
        os.remove(path)
    append_results(rows, path)
    print(f"[OK] wrote {len(rows)} rows -> {path}")
    from . import plots
    plots.plot_scaling(args.results_dir)
    return rows


if __name__ == "__main__":
    main()
