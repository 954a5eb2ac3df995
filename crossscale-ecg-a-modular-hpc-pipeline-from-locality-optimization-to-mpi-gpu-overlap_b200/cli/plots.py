"""Result aggregation + plots.

The numeric part (grouping, means, amortised "effective" throughput) needs only the CSVs and always
runs; PNGs are written when matplotlib (Agg backend) is importable — it is not in this image, so the
aggregated tables are also written as CSV next to where the PNG would go.

  plot_locality      Module_1/plot_locality.py:7-73       throughput_vs_batch.png, time_breakdown_stacked.png
  plot_all_results   Module_1/plot_all_results.py:1-129    part1_all_results.csv, effective A4 throughput
  plot_part2         Module_2/benchmark_part_2.py:141-173  part2_throughput.png, part2_speedup.png
  plot_scaling       Module_2/plot_part2.py:1-17           part2_scaling.png
  plot_part3         Module_3/plot_part3.py:1-149          throughput / breakdown vs world size (pseudo-FL)
  plot_part3_fl      TRUE_FL_M3/plot_part3.py:1-164        same for FedAvg rounds (local_train_ms + comm_ms)
"""
from __future__ import annotations

import argparse
import glob
import json
import os
from collections import defaultdict
from typing import Dict, List, Optional

from ..fl.stats import append_results, read_csv

EPOCHS = 10            # plot_all_results.py:8 — epochs over which shard preparation is amortised
BREAKDOWN_BATCH = 512


def _plt():
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
        return plt
    except Exception:
        return None


def _f(v, default=0.0):
    try:
        return float(v)
    except (TypeError, ValueError):
        return default


def group_mean(rows: List[dict], keys: List[str], values: List[str]) -> List[dict]:
    acc: Dict[tuple, Dict[str, List[float]]] = defaultdict(lambda: defaultdict(list))
    for r in rows:
        k = tuple(r[x] for x in keys)
        for v in values:
            if v in r and r[v] != "":
                acc[k][v].append(_f(r[v]))
    out = []
    for k, d in sorted(acc.items(), key=lambda kv: tuple(_f(x, 0) if str(x).replace('.', '', 1).isdigit() else str(x) for x in kv[0])):
        row = dict(zip(keys, k))
        row.update({v: sum(xs) / len(xs) for v, xs in d.items()})
        out.append(row)
    return out


def _write_table(rows, path):
    if os.path.exists(path):
        os.remove(path)
    if rows:
        append_results(rows, path)
    return path


def canonical_config(name: str) -> str:
    """G0* -> G0_baseline, G1* -> G1_overlap_amp (Module_3/plot_part3.py)."""
    if name.startswith("G0"):
        return "G0_baseline"
    if name.startswith("G1"):
        return "G1_overlap_amp"
    return name


# ---- Module 1 ---------------------------------------------------------------------------------------
def plot_locality(results_dir: str, csv_name: str = "part1_locality_results.csv", batch: int = BREAKDOWN_BATCH):
    path = os.path.join(results_dir, csv_name)
    if not os.path.exists(path):
        return None
    rows = read_csv(path)
    table = group_mean(rows, ["config", "batch_size"], ["samples_per_s", "data_ms", "h2d_ms", "compute_ms", "step_ms"])
    _write_table(table, os.path.join(results_dir, "part1_locality_summary.csv"))
    plt = _plt()
    if plt is None:
        return table
    fig, ax = plt.subplots(figsize=(7, 4))
    for cfg in sorted({r["config"] for r in table}):
        pts = sorted(((int(_f(r["batch_size"])), r["samples_per_s"]) for r in table if r["config"] == cfg))
        ax.plot([p[0] for p in pts], [p[1] for p in pts], marker="o", label=cfg)
    ax.set_xlabel("batch size"); ax.set_ylabel("samples / s"); ax.legend(); ax.grid(alpha=.3)
    fig.tight_layout(); fig.savefig(os.path.join(results_dir, "throughput_vs_batch.png"), dpi=130); plt.close(fig)
    sel = [r for r in table if int(_f(r["batch_size"])) == batch]
    if sel:
        fig, ax = plt.subplots(figsize=(7, 4))
        names = [r["config"] for r in sel]
        bottom = [0.0] * len(sel)
        for comp in ("data_ms", "h2d_ms", "compute_ms"):
            vals = [r[comp] for r in sel]
            ax.bar(names, vals, bottom=bottom, label=comp)
            bottom = [b + v for b, v in zip(bottom, vals)]
        ax.set_ylabel(f"ms / step (batch {batch})"); ax.legend()
        fig.tight_layout(); fig.savefig(os.path.join(results_dir, "time_breakdown_stacked.png"), dpi=130); plt.close(fig)
    return table


def plot_labl(results_dir: str):
    path = os.path.join(results_dir, "part1_labl_results.csv")
    if not os.path.exists(path):
        return None
    rows = read_csv(path)
    plt = _plt()
    if plt is not None:
        fig, ax = plt.subplots(figsize=(6, 4))
        ax.plot([int(_f(r["batch_size"])) for r in rows], [_f(r["samples_per_s"]) for r in rows], marker="o", label="A4_LABL")
        ax.set_xlabel("batch size"); ax.set_ylabel("samples / s"); ax.legend(); ax.grid(alpha=.3)
        fig.tight_layout(); fig.savefig(os.path.join(results_dir, "labl_throughput.png"), dpi=130); plt.close(fig)
    return rows


def effective_throughput(sps: float, shard_time_s: float, total_windows: int, epochs: int = EPOCHS) -> float:
    """A4 throughput with the one-off shard preparation amortised over ``epochs`` passes
    (``plot_all_results.py:56-58``): eff = sps / (1 + shard_time / (epochs * N / sps))."""
    if sps <= 0 or total_windows <= 0:
        return 0.0
    return sps / (1.0 + shard_time_s / (epochs * total_windows / sps))


def plot_all_results(results_dir: str, shard_json: Optional[str] = None, epochs: int = EPOCHS):
    a = os.path.join(results_dir, "part1_locality_results.csv")
    b = os.path.join(results_dir, "part1_labl_results.csv")
    rows = (read_csv(a) if os.path.exists(a) else []) + (read_csv(b) if os.path.exists(b) else [])
    if not rows:
        return None
    shard_json = shard_json or os.path.join(results_dir, "shard_prep_metrics.json")
    shard_time, total = 0.0, 0
    if os.path.exists(shard_json):
        with open(shard_json) as f:
            m = json.load(f)
        shard_time, total = float(m.get("total_time_s", 0.0)), int(m.get("total_windows", 0))
    merged = []
    for r in rows:
        sps = _f(r.get("samples_per_s"))
        row = {"config": r["config"], "batch_size": int(_f(r["batch_size"])), "samples_per_s": sps,
               "data_ms": _f(r.get("data_ms")), "h2d_ms": _f(r.get("h2d_ms")), "compute_ms": _f(r.get("compute_ms")),
               "step_ms": _f(r.get("step_ms")), "effective_samples_per_s": sps, "shard_ms_per_step": 0.0}
        if r["config"].startswith("A4") and total > 0:
            row["effective_samples_per_s"] = effective_throughput(sps, shard_time, total, epochs)
            steps = epochs * total / max(1, row["batch_size"])
            row["shard_ms_per_step"] = shard_time * 1e3 / steps
        merged.append(row)
    _write_table(merged, os.path.join(results_dir, "part1_all_results.csv"))
    plt = _plt()
    if plt is not None:
        fig, ax = plt.subplots(figsize=(7, 4))
        for cfg in sorted({r["config"] for r in merged}):
            pts = sorted((r["batch_size"], r["effective_samples_per_s"]) for r in merged if r["config"] == cfg)
            ax.plot([p[0] for p in pts], [p[1] for p in pts], marker="o", label=cfg)
        ax.set_xlabel("batch size"); ax.set_ylabel("effective samples / s"); ax.legend(); ax.grid(alpha=.3)
        fig.tight_layout(); fig.savefig(os.path.join(results_dir, "part1_all_throughput.png"), dpi=130); plt.close(fig)
    return merged


# ---- Module 2 ---------------------------------------------------------------------------------------
def plot_part2(results_dir: str):
    path = os.path.join(results_dir, "part2_openmp_results.csv")
    if not os.path.exists(path):
        return None
    rows = read_csv(path)
    plt = _plt()
    if plt is not None:
        for name, ycols in (("part2_throughput.png", ("torch_sps", "omp_sps")), ("part2_speedup.png", ("speedup_med",))):
            fig, ax = plt.subplots(figsize=(7, 4))
            for k in sorted({int(_f(r["kernel_size"])) for r in rows}):
                sel = sorted((r for r in rows if int(_f(r["kernel_size"])) == k), key=lambda r: _f(r["batch_size"]))
                for c in ycols:
                    ax.plot([_f(r["batch_size"]) for r in sel], [_f(r[c]) for r in sel], marker="o", label=f"{c} K={k}")
            ax.set_xlabel("batch size"); ax.legend(fontsize=7); ax.grid(alpha=.3)
            fig.tight_layout(); fig.savefig(os.path.join(results_dir, name), dpi=130); plt.close(fig)
    return rows


def plot_scaling(results_dir: str):
    path = os.path.join(results_dir, "part2_openmp_simd_results.csv")
    if not os.path.exists(path):
        return None
    rows = read_csv(path)
    plt = _plt()
    if plt is not None:
        fig, ax = plt.subplots(figsize=(6, 4))
        for b in sorted({int(_f(r["batch"])) for r in rows}):
            sel = sorted((r for r in rows if int(_f(r["batch"])) == b), key=lambda r: _f(r["threads"]))
            ax.plot([_f(r["threads"]) for r in sel], [_f(r["samples_per_s"]) for r in sel], marker="o", label=f"batch {b}")
        ax.set_xlabel("threads"); ax.set_ylabel("samples / s"); ax.legend(); ax.grid(alpha=.3)
        fig.tight_layout(); fig.savefig(os.path.join(results_dir, "part2_scaling.png"), dpi=130); plt.close(fig)
    return rows


# ---- Module 3 ---------------------------------------------------------------------------------------
def _world_plots(table, results_dir, prefix, stack_cols):
    plt = _plt()
    if plt is None:
        return
    fig, ax = plt.subplots(figsize=(6, 4))
    for cfg in sorted({r["config"] for r in table}):
        pts = sorted((int(_f(r["world_size"])), r["samples_per_s"]) for r in table if r["config"] == cfg)
        ax.plot([p[0] for p in pts], [p[1] for p in pts], marker="o", label=cfg)
    ax.set_xlabel("world size (clients)"); ax.set_ylabel("samples / s per client"); ax.legend(); ax.grid(alpha=.3)
    fig.tight_layout(); fig.savefig(os.path.join(results_dir, f"{prefix}_throughput_vs_world.png"), dpi=130); plt.close(fig)
    fig, ax = plt.subplots(figsize=(7, 4))
    labels = [f"{r['config']}\nW={int(_f(r['world_size']))}" for r in table]
    bottom = [0.0] * len(table)
    for c in stack_cols:
        vals = [r.get(c, 0.0) for r in table]
        ax.bar(labels, vals, bottom=bottom, label=c)
        bottom = [b + v for b, v in zip(bottom, vals)]
    ax.set_ylabel("ms"); ax.legend(); ax.tick_params(axis="x", labelsize=6)
    fig.tight_layout(); fig.savefig(os.path.join(results_dir, f"{prefix}_breakdown.png"), dpi=130); plt.close(fig)


def plot_part3(results_dir: str, csv_name: str = "part3_mpi_cuda_results.csv"):
    path = os.path.join(results_dir, csv_name)
    if not os.path.exists(path):
        return None
    rows = read_csv(path)
    for r in rows:
        r["config"] = canonical_config(r["config"])
    table = group_mean(rows, ["world_size", "config"], ["samples_per_s", "h2d_ms", "compute_ms", "step_ms", "data_ms"])
    _write_table(table, os.path.join(results_dir, "part3_summary.csv"))
    _world_plots(table, results_dir, "part3", ("h2d_ms", "compute_ms"))
    return table


def plot_part3_fl(results_dir: str, pattern: str = "fedavg_results*.csv"):
    """Mean over ranks, rounds and repeats per (world_size, config); ``step_ms = local_train_ms + comm_ms``
    (``TRUE_FL_M3/plot_part3.py:48,53-62``).  Globs ``fedavg_results*.csv`` so both the driver's
    ``fedavg_results.csv`` and per-run ``fedavg_results_<tag>.csv`` files are found (the reference's
    glob misses the file its own driver writes)."""
    rows = []
    for p in sorted(glob.glob(os.path.join(results_dir, pattern))):
        rows += read_csv(p)
    if not rows:
        return None
    for r in rows:
        r["config"] = canonical_config(r["config"])
        r["step_ms"] = _f(r["local_train_ms"]) + _f(r["comm_ms"])
    table = group_mean(rows, ["world_size", "config"], ["samples_per_s", "local_train_ms", "comm_ms", "step_ms", "avg_loss"])
    _write_table(table, os.path.join(results_dir, "fedavg_summary.csv"))
    _world_plots(table, results_dir, "fedavg", ("local_train_ms", "comm_ms"))
    return table


def main(argv=None):
    ap = argparse.ArgumentParser(description="aggregate result CSVs (and plot when matplotlib is present)")
    ap.add_argument("what", choices=["locality", "all", "part2", "scaling", "part3", "part3_fl"])
    ap.add_argument("--results-dir", default="results")
    ap.add_argument("--csv", default=None)
    ap.add_argument("--batch", type=int, default=BREAKDOWN_BATCH)
    a = ap.parse_args(argv)
    if a.what == "locality":
        out = plot_locality(a.results_dir, a.csv or "part1_locality_results.csv", a.batch)
    elif a.what == "all":
        out = plot_all_results(a.results_dir)
    elif a.what == "part2":
        out = plot_part2(a.results_dir)
    elif a.what == "scaling":
        out = plot_scaling(a.results_dir)
    elif a.what == "part3":
        out = plot_part3(a.results_dir, a.csv or "part3_mpi_cuda_results.csv")
    else:
        out = plot_part3_fl(a.results_dir)
    print(f"[plots] {a.what}: {0 if out is None else len(out)} rows" + ("" if _plt() else " (matplotlib absent: tables only)"))
    return out


if __name__ == "__main__":
    main()
