"""A4 config: training fed by the LABL pinned-ring prefetcher — ``python -m crossscale.cli.train_ecg_labl``.

Flags of ``Module_1/train_ecg_labl(EXPERIMENTAL).py:99-102``.  Same three-way timing as the locality
benchmark; ``data_ms`` is the prefetch thread's own ``fill_ms``, labels come from the ``.lbl`` sidecars
(zeros when absent, as in the reference), the copy is one coalesced ``non_blocking`` H2D per batch
followed by ``recycle(slot)``.  Writes ``results/part1_labl_results.csv`` (``:91-94,108`` column order).
"""
from __future__ import annotations

import argparse
import os
import time
from glob import glob

import torch

from ..data.pinned_ring import LABLPrefetcher, LABLShardedReader
from ..fl.stats import append_results
from ..models import Tiny1D

COLUMNS = ["config", "batch_size", "step_ms", "samples_per_s", "data_ms", "h2d_ms", "compute_ms"]


def bench_labl(shard_paths, batch_size: int, iters: int, device: torch.device, num_slots: int = 4,
               normalize: bool = True, num_classes: int = 2) -> dict:
    assert shard_paths, "no shards matched"
    pf = LABLPrefetcher(LABLShardedReader(shard_paths), batch_size, num_slots=num_slots, normalize=normalize, loop=True)
    model = Tiny1D(num_classes=num_classes).to(device)
    opt = torch.optim.SGD(model.parameters(), lr=1e-2)
    cuda = device.type == "cuda"
    pf.start()
    data_ms = h2d_ms = comp_ms = 0.0
    n = steps = 0
    try:
        for i in range(iters + 5):
            got = pf.next_batch_cpu()
            if got is None:
                break
            slot, xb, fill_ms = got
            yb = pf.labels_of(slot, xb.shape[0]).clamp_(max=num_classes - 1)
            if cuda:
                torch.cuda.synchronize()
            t0 = time.perf_counter()
            if cuda:
                x = xb.to(device, non_blocking=True)
                y = yb.to(device, non_blocking=True)
            else:                               # no copy on CPU: detach from the slot before recycling it
                x, y = xb.clone(), yb.clone()
            if cuda:
                torch.cuda.synchronize()
            t1 = time.perf_counter()
            pf.recycle(slot)
            loss = torch.nn.functional.cross_entropy(model(x), y)
            loss.backward()
            opt.step()
            opt.zero_grad(set_to_none=True)
            if cuda:
                torch.cuda.synchronize()
            t2 = time.perf_counter()
            if i >= 5:
                data_ms += fill_ms
                h2d_ms += (t1 - t0) * 1e3 if cuda else 0.0
                comp_ms += (t2 - t1) * 1e3
                n += x.shape[0]
                steps += 1
    finally:
        pf.shutdown()
    steps = max(steps, 1)
    step_ms = (data_ms + h2d_ms + comp_ms) / steps
    return dict(config="A4_LABL", batch_size=batch_size, step_ms=step_ms,
                samples_per_s=(n / steps) / (step_ms / 1e3) if step_ms > 0 else 0.0,
                data_ms=data_ms / steps, h2d_ms=h2d_ms / steps, compute_ms=comp_ms / steps)


def build_parser() -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser()
    ap.add_argument("--shards", type=str, default="data/shards/ecg_*.bin")
    ap.add_argument("--batch-sizes", nargs="+", type=int, default=[64, 128, 256, 512])
    ap.add_argument("--iters", type=int, default=100)
    ap.add_argument("--device", type=str, default="cuda")
    ap.add_argument("--results-dir", type=str, default="results")
    return ap


def main(argv=None):
    args = build_parser().parse_args(argv)
    device = torch.device(args.device if torch.cuda.is_available() and args.device.startswith("cuda") else "cpu")
    paths = sorted(glob(args.shards))
    rows = [bench_labl(paths, bs, args.iters, device) for bs in args.batch_sizes]
    for r in rows:
        print(f"[A4_LABL] bs={r['batch_size']}: {r['samples_per_s']:,.0f} samples/s")
    path = os.path.join(args.results_dir, "part1_labl_results.csv")
    if os.path.exists(path):
        os.remove(path)
    append_results([{c: r[c] for c in COLUMNS} for r in rows], path)
    print(f"[OK] wrote {len(rows)} rows -> {path}")
    from . import plots
    plots.plot_labl(args.results_dir)
    return rows


if __name__ == "__main__":
    main()
