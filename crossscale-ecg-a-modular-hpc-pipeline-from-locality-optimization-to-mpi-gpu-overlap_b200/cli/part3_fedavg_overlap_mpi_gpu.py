"""True FedAvg driver — ``torchrun --nproc-per-node W -m crossscale.cli.part3_fedavg_overlap_mpi_gpu``.

Flag-compatible with ``TRUE_FL_M3/part3_fedavg_overlap_mpi_gpu.py:141-146`` (``--data-root --batch-size
--rounds --local-steps --max-windows --config``); same CSV (``results/fedavg_results.csv``, RoundStats
columns first), seeds ``1234 + rank``, fresh model per config.  MPI is replaced by torchrun for the
launch and by the in-kernel NVLink collective for aggregation.
"""
from __future__ import annotations

import argparse
import os
from glob import glob

import torch

from ..data.shard_dataset import assign_shards_evenly, load_shards_to_gpu
from ..fl.engine import FedAvgRunner
from ..fl.stats import append_results
from ..models import build_model
from ..parallel import launch
from ..utils.misc import load_checkpoint, save_checkpoint, set_basic_seeds, setup_device


def build_parser() -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser()
    ap.add_argument("--data-root", required=True, type=str)
    ap.add_argument("--batch-size", type=int, default=256)
    ap.add_argument("--rounds", type=int, default=10)
    ap.add_argument("--local-steps", type=int, default=50)
    ap.add_argument("--max-windows", type=int, default=30000)
    ap.add_argument("--config", choices=["G0", "G1", "both"], default="both")
    # extensions
    ap.add_argument("--amp", choices=["bf16", "fp16"], default="bf16", help="G1 precision")
    ap.add_argument("--model", default="tiny_ecg")
    ap.add_argument("--num-classes", type=int, default=2)
    ap.add_argument("--agg-backend", choices=["auto", "kernel", "nccl", "gloo"], default="auto")
    ap.add_argument("--wire", choices=["fp32", "bf16"], default="fp32")
    ap.add_argument("--server-lr", type=float, default=1.0)
    ap.add_argument("--reset-momentum", action="store_true")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--sync-grads", action="store_true",
                    help="all-reduce gradients every step (bucketed, overlapped with backward) in addition to FedAvg")
    ap.add_argument("--results-dir", type=str, default="results")
    ap.add_argument("--save", type=str, default=None, help="save the final global state_dict here")
    ap.add_argument("--load", type=str, default=None, help="start from this state_dict")
    return ap


def main(argv=None):
    args = build_parser().parse_args(argv)
    rank, world, local_rank = launch.init_distributed()
    torch.set_num_threads(1)
    set_basic_seeds(1234 + rank)
    device = setup_device(local_rank)
    shard_paths = sorted(glob(os.path.join(args.data_root, "ecg_*.bin")))
    if not shard_paths:
        raise RuntimeError("No ecg_*.bin files found")
    local_shards = assign_shards_evenly(shard_paths, world, rank)
    x_gpu, y_gpu = load_shards_to_gpu(local_shards, device=device, max_windows=args.max_windows)
    csv_path = os.path.join(args.results_dir, "fedavg_results.csv")
    configs = [c for c in ("G0", "G1") if args.config in (c, "both")]
    all_rows = []
    for cfg in configs:
        model = build_model(args.model, num_classes=args.num_classes).to(device)
        if args.load:
            load_checkpoint(model, args.load)
        runner = FedAvgRunner(model, device, x_gpu, y_gpu, args.batch_size, args.local_steps, cfg,
                              amp=("none" if cfg == "G0" else args.amp), rank=rank, world=world,
                              agg_backend=args.agg_backend, wire=args.wire, server_lr=args.server_lr,
                              reset_momentum=args.reset_momentum, use_graph=not args.no_graph,
                              model_name=args.model, sync_grads=args.sync_grads)
        rows = runner.run(args.rounds)
        gathered = launch.gather_rows(rows)
        if rank == 0:
            append_results(gathered, csv_path)
            print(f"[OK] wrote {len(gathered)} rows -> {csv_path}")
            if args.save:
                if not runner.fused:
                    runner.trainer.load_flat(runner.flat)
                save_checkpoint(model, args.save if len(configs) == 1 else f"{args.save}.{cfg}")
        all_rows.extend(gathered)
        runner.close()
    launch.shutdown()
    return all_rows


if __name__ == "__main__":
    main()
