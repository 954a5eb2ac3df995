"""Module-1 locality sweep: ``python -m crossscale.cli.bench_locality`` (flags of
``Module_1/bench_locality.py:80-87``).

For each batch size x config it measures per step ``data_ms`` (next(it)), ``h2d_ms`` (the host->device
copy, synchronised on both sides) and ``compute_ms`` (fwd + CE + bwd + SGD), 5 warm-ups, ``iters`` timed
steps; ``samples_per_s = B / (data_ms + h2d_ms + compute_ms)`` (``:73-76``).  Configs A0..A3 as in the
reference (``:111-116``) — except that A0 really samples randomly and A1..A3 contiguously (in the
reference the sampler is a global flag, so A0 == A1).  ``--contiguous/--random`` still
forces one sampler for every config when given explicitly.  Output: ``results/part1_locality_results.csv``
(+ PNGs when matplotlib is installed).
"""
from __future__ import annotations

import argparse
import os
import time

import torch

from ..fl.stats import append_results
from ..models import Tiny1D

CONFIGS = [  # name, pin_memory, non_blocking, contiguous
    ("A0_baseline", False, False, False),
    ("A1_contiguous", False, False, True),
    ("A2_contig_pinned", True, False, True),
    ("A3_contig_pinned_nb", True, True, True),
]
COLUMNS = ["config", "batch_size", "pin_memory", "contiguous", "non_blocking", "data_ms", "h2d_ms", "compute_ms",
           "step_ms", "samples_per_s"]


def _next(it, dl):
    try:
        return next(it), it
    except StopIteration:
        it = iter(dl)
        return next(it), it


def measure_step(dl, device: torch.device, non_blocking: bool, iters: int = 50, num_classes: int = 2) -> dict:
    model = Tiny1D(num_classes=num_classes).to(device)
    opt = torch.optim.SGD(model.parameters(), lr=1e-2)
    model.train()
    cuda = device.type == "cuda"

    def train(x, y):
        loss = torch.nn.functional.cross_entropy(model(x), y)
        loss.backward()
        opt.step()
        opt.zero_grad(set_to_none=True)

    it = iter(dl)
    for _ in range(5):
        (x_cpu, y), it = _next(it, dl)
        train(x_cpu.to(device, non_blocking=non_blocking), y.to(device, non_blocking=non_blocking))
    data_ms = h2d_ms = comp_ms = 0.0
    n = 0
    it = iter(dl)
    for _ in range(iters):
        t0 = time.perf_counter()
        (x_cpu, y), it = _next(it, dl)
        t1 = time.perf_counter()
        if cuda:
            torch.cuda.synchronize()
        th0 = time.perf_counter()
        x = x_cpu.to(device, non_blocking=non_blocking)
        y = y.to(device, non_blocking=non_blocking)
        if cuda:
            torch.cuda.synchronize()
        th1 = time.perf_counter()
        train(x, y)
        if cuda:
            torch.cuda.synchronize()
        t3 = time.perf_counter()
        data_ms += (t1 - t0) * 1e3
        h2d_ms += (th1 - th0) * 1e3 if cuda else 0.0
        comp_ms += (t3 - th1) * 1e3
        n += x.shape[0]
    step_ms = (data_ms + h2d_ms + comp_ms) / iters
    return dict(data_ms=data_ms / iters, h2d_ms=h2d_ms / iters, compute_ms=comp_ms / iters, step_ms=step_ms,
                samples_per_s=(n / iters) / (step_ms / 1e3))


def build_parser() -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser()
    ap.add_argument("--dataset", choices=["mitbih", "synthetic"], default="mitbih")
    ap.add_argument("--device", type=str, default="cuda")
    ap.add_argument("--batch-sizes", nargs="+", type=int, default=[64, 128, 256, 512])
    ap.add_argument("--iters", type=int, default=100)
    ap.add_argument("--num-workers", type=int, default=4)
    ap.add_argument("--contiguous", dest="contiguous", action="store_const", const=True, default=None)
    ap.add_argument("--random", dest="contiguous", action="store_const", const=False)
    ap.add_argument("--results-dir", type=str, default="results")
    ap.add_argument("--num-windows", type=int, default=20000)
    return ap


def main(argv=None):
    args = build_parser().parse_args(argv)
    device = torch.device(args.device if torch.cuda.is_available() and args.device.startswith("cuda") else "cpu")
    from ..data.loaders import make_mitbih_loader, make_synth_loader
    if args.dataset == "mitbih":
        maker, ncls = make_mitbih_loader, 5          # MIT-BIH-shaped beats (no network for PhysioNet here)
    else:
        maker, ncls = make_synth_loader, 2
    rows = []
    for bs in args.batch_sizes:
        for name, pin, nb, contig in CONFIGS:
            contig = contig if args.contiguous is None else args.contiguous
            dl = maker(batch_size=bs, num_workers=args.num_workers, pin_memory=pin, contiguous=contig,
                       num_windows=args.num_windows)
            stats = measure_step(dl, device, nb, args.iters, ncls)
            rows.append(dict(config=name, batch_size=bs, pin_memory=pin, contiguous=contig, non_blocking=nb, **stats))
            print(f"[{name}] bs={bs}: {stats['samples_per_s']:,.0f} samples/s "
                  f"(data {stats['data_ms']:.2f} h2d {stats['h2d_ms']:.2f} compute {stats['compute_ms']:.2f} ms)")
            del dl
    path = os.path.join(args.results_dir, "part1_locality_results.csv")
    if os.path.exists(path):
        os.remove(path)
    append_results([{c: r[c] for c in COLUMNS} for r in rows], path)
    print(f"[OK] wrote {len(rows)} rows -> {path}")
    from . import plots
    plots.plot_locality(args.results_dir, "part1_locality_results.csv", max(args.batch_sizes))
    return rows


if __name__ == "__main__":
    main()
