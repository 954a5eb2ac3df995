"""Pseudo-federated driver (independent clients, no model exchange).

``torchrun --nproc-per-node N -m crossscale.cli.part3_mpi_gpu_train --batch-size 256 --steps 200``
— flags of ``Module_3/part3_mpi_gpu_train.py:424-430``; ``--data-root`` is honoured (the reference
parses and ignores it).  Runs G0 then (on CUDA) G1 with fresh ``TinyECG(num_classes=2)``,
gathers ``BenchStats`` rows on rank 0 into ``results/part3_mpi_cuda_results.csv``.
"""
from __future__ import annotations

import argparse
import os
from glob import glob

import torch

from ..data.shard_dataset import assign_shards_evenly, load_shards_to_gpu, make_gpu_batch_iter
from ..fl.engine import run_baseline_gpu, run_overlap_gpu
from ..fl.stats import append_results, summarize
from ..models import TinyECG
from ..parallel import launch
from ..utils.misc import setup_device


def build_parser() -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch-size", type=int, default=256)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--max-windows", type=int, default=20000)
    ap.add_argument("--data-root", type=str, default=os.path.join("..", "PART-1", "data", "shards"))
    ap.add_argument("--amp", choices=["bf16", "fp16"], default="bf16")
    ap.add_argument("--results-dir", type=str, default="results")
    return ap


def main(argv=None):
    args = build_parser().parse_args(argv)
    rank, world, local_rank = launch.init_distributed()
    device = setup_device(local_rank)
    shard_paths = sorted(glob(os.path.join(args.data_root, "ecg_*.bin")))
    if not shard_paths:
        raise RuntimeError(f"No ecg_*.bin shards under {args.data_root}")
    launch.barrier()
    mine = assign_shards_evenly(shard_paths, world, rank)
    x_gpu, y_gpu = load_shards_to_gpu(mine, device=device, max_windows=args.max_windows)
    rows = []
    it = make_gpu_batch_iter(x_gpu, y_gpu, args.batch_size)
    rows.append(run_baseline_gpu(TinyECG(num_classes=2).to(device), it, device, args.steps, rank,
                                 args.batch_size, world, x_gpu, y_gpu))
    if device.type == "cuda":
        rows.append(run_overlap_gpu(TinyECG(num_classes=2).to(device), it, device, args.steps, rank,
                                    args.batch_size, world, x_gpu, y_gpu, amp=args.amp))
    elif rank == 0:
        print("[INFO] CUDA not available: G1 (AMP) skipped")
    gathered = launch.gather_rows(rows)
    if rank == 0:
        path = append_results(gathered, os.path.join(args.results_dir, "part3_mpi_cuda_results.csv"))
        for cfg, sps in summarize(gathered).items():
            print(f"[SUMMARY] {cfg}: mean samples/s per client = {sps:,.0f}")
        print(f"[OK] wrote {len(gathered)} rows -> {path}")
    launch.shutdown()
    return gathered


if __name__ == "__main__":
    main()
