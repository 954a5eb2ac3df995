"""Module-1 locality sweep: ``python -m crossscale.cli.bench_locality`` (flags of
``Module_1/bench_locality.py:80-87``).

For each batch size x config it measures per step ``data_ms`` (next(it)), ``h2d_ms`` (the host->device
copy, synchronised on both sides) and ``compute_ms`` (fwd + CE + bwd + SGD), 5 warm-ups, ``iters`` timed
steps; ``samples_per_s = B / (data_ms + h2d_ms + compute_ms)`` (``:73-76``).  Configs A0..A3 as in the
reference (``:111-116``) — except that A0 really samples randomly and A1..A3 contiguously (in the
reference the sampler is a global flag, so A0 == A1).  ``--contiguous/--random`` still
forces one sampler for every config when given explicitly.  Output: ``results/part1_locality_results.csv``
(+ PNGs when matplotlib is installed).

On a CUDA device one more row per batch size is measured, **A5_b200_staged_fused**: the B200-native answer to the
reference's "pinned + non-blocking" idea.  Windows live in pinned host memory as bf16 (``PinnedShardStore``), blocks of
``--a5-steps`` batches go host->device with ``cudaMemcpyAsync`` on a copy stream into a double-buffered slot
(``H2DStager``) while the previous block trains, and a block's steps are ONE launch of the persistent fused kernel
(``cs_tiny_ecg_train_round``).  Its ``h2d_ms`` is the copy's own duration per step (it overlaps compute), ``step_ms``
the wall time per step, ``samples_per_s = B / step_ms``.
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


def measure_a5(device: torch.device, batch: int, iters: int, num_classes: int, dataset: str, num_windows: int,
               steps_per_block: int = 10, amp: str = "bf16") -> dict:
    """Pinned store -> copy-stream staging -> fused kernels; wall-clock per step with the copy overlapped."""
    import numpy as np
    from ..data.staging import BlockSampler, H2DStager, PinnedShardStore
    from ..data.synth import make_mitbih_like_beats, make_synth_windows
    from ..models import TinyECG
    from ..ops import tiny_ecg as tk
    if dataset == "mitbih":
        x, y = make_mitbih_like_beats(num_windows, 360, num_classes, 1337)
    else:
        x, y = make_synth_windows(num_windows, 500, 1337), np.zeros(num_windows, dtype=np.int64)
    dt = torch.bfloat16 if amp == "bf16" else torch.float32
    store = PinnedShardStore([], dtype=dt, x=torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32)),
                             y=torch.from_numpy(np.ascontiguousarray(y, dtype=np.int64)))
    rows = steps_per_block * batch
    stager, sampler = H2DStager(rows, store.l, dt, device), BlockSampler(store.n, rows, seed=0)
    model = TinyECG(num_classes).to(device)
    st = tk.TinyTrainState(model.flatten_parameters().flat, num_classes, lr=1e-2, momentum=0.0, amp=tk.AMP_BY_NAME[amp])
    use_round = st.round_supported(store.l) and tk.round_kernel_available()
    blocks = max(2, (iters + steps_per_block - 1) // steps_per_block)
    t_data = 0.0

    def stage(slot):
        nonlocal t_data
        t0 = time.perf_counter()
        hx, hy = store.block(sampler.next_start(), rows)
        t_data += (time.perf_counter() - t0) * 1e3
        stager.stage(slot, hx, hy)

    def train(xb, yb):
        if use_round:
            st.round(xb, yb, steps_per_block, batch, sample_mode=tk.SAMPLE_ROWS)
        else:
            for s_ in range(steps_per_block):
                st.step(xb[s_ * batch:(s_ + 1) * batch], yb[s_ * batch:(s_ + 1) * batch], None, batch=batch)

    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    slot, wall = 0, 0.0
    stage(0)
    for n_blocks in (3, blocks):                              # warm-up, then the timed blocks
        torch.cuda.synchronize()
        t_data, stager.trace = 0.0, []
        t0 = time.perf_counter()
        ev[0].record()
        for _ in range(n_blocks):
            xb, yb = stager.acquire(slot)
            stage(slot ^ 1)                                  # next block's copy runs under this block's compute
            train(xb, yb)
            stager.release(slot)
            slot ^= 1
        ev[1].record()
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) * 1e3
    st.check_errors()
    steps = blocks * steps_per_block
    copy_ms = sum(a.elapsed_time(b) for a, b in stager.trace) / max(1, len(stager.trace)) / steps_per_block
    step_ms = max(wall, ev[0].elapsed_time(ev[1])) / steps
    return dict(data_ms=t_data / steps, h2d_ms=copy_ms, compute_ms=ev[0].elapsed_time(ev[1]) / steps, step_ms=step_ms,
                samples_per_s=batch / (step_ms / 1e3))


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
    ap.add_argument("--a5-steps", type=int, default=10, help="batches per staged block of the A5 row")
    ap.add_argument("--no-a5", action="store_true", help="skip the B200-native A5 row")
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
        if device.type == "cuda" and not args.no_a5:
            from ..ops import tiny_ecg as tk
            if tk.use_kernels():
                stats = measure_a5(device, bs, args.iters, ncls, args.dataset, max(args.num_windows, 2 * args.a5_steps * bs),
                                   args.a5_steps)
                rows.append(dict(config="A5_b200_staged_fused", batch_size=bs, pin_memory=True, contiguous=True,
                                 non_blocking=True, **stats))
                print(f"[A5_b200_staged_fused] bs={bs}: {stats['samples_per_s']:,.0f} samples/s "
                      f"(data {stats['data_ms']:.3f} h2d {stats['h2d_ms']:.3f} [overlapped] device {stats['compute_ms']:.3f} "
                      f"wall {stats['step_ms']:.3f} ms/step)")
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
