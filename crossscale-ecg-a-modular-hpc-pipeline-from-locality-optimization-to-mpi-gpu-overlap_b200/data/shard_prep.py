"""Shard preparation: windows -> ``ecg_%05d.bin`` shards + ``shard_prep_metrics.json``.

Behavioural reference: ``Module_1/shard_prep.py:39-94`` (flags, file naming, metrics keys).
Differences by design: importable without side effects (the reference's module-level JSON block
raises on import), output directory / N configurable, MIT-BIH-shaped synthetic beats
with labels (sidecar), and MIT-BIH itself only when ``wfdb`` and a network are present.
"""
from __future__ import annotations

import json
import os
import time
from typing import Optional, Sequence

import numpy as np

from .shard_format import shard_name, write_shard
from .synth import make_mitbih_like_beats, make_synth_windows

DEFAULT_OUT_DIR = os.path.join("data", "shards")
DEFAULT_RECORDS = ("100", "101", "103", "105", "106")


def make_mitbih_windows(records: Sequence[str] = DEFAULT_RECORDS, win_len: int = 500,
                        stride: int = 250, channel: int = 0) -> np.ndarray:
    """Sliding windows over MIT-BIH records fetched with ``wfdb`` (needs network).

    Same windowing rule as ``shard_prep.py:21-33``: starts ``range(0, len - win_len, stride)``,
    one channel, no normalisation.
    """
    try:
        import wfdb  # type: ignore
    except Exception as e:  # pragma: no cover - wfdb is not in this image
        raise RuntimeError("wfdb is not installed; use --dataset synthetic or mitbih_like") from e
    out = []
    for rid in records:  # pragma: no cover - network path
        sig, _ = wfdb.rdsamp(f"mitdb/{rid}", pn_dir="mitdb")
        x = np.asarray(sig[:, channel], dtype=np.float32)
        starts = np.arange(0, len(x) - win_len, stride)
        idx = starts[:, None] + np.arange(win_len)[None, :]
        out.append(x[idx])
    return np.concatenate(out, axis=0).astype(np.float32)


def prepare_shards(dataset: str = "mitbih", win_len: int = 500, stride: int = 250,
                   shard_size: int = 32768, out_dir: str = DEFAULT_OUT_DIR,
                   num_windows: int = 200000, results_dir: Optional[str] = "results",
                   seed: int = 1337, num_classes: int = 5, verbose: bool = True) -> dict:
    """Generate/load windows, write shards, write the metrics JSON, return the metrics dict."""
    t0 = time.perf_counter()
    labels = None
    if dataset == "mitbih":
        windows = make_mitbih_windows(win_len=win_len, stride=stride)
    elif dataset == "synthetic":
        windows = make_synth_windows(N=num_windows, L=win_len, seed=seed)
    elif dataset == "mitbih_like":
        windows, labels = make_mitbih_like_beats(N=num_windows, L=win_len,
                                                 num_classes=num_classes, seed=seed)
    else:
        raise ValueError(f"unknown dataset {dataset!r}")
    t_load = time.perf_counter()
    n, l = windows.shape
    if verbose:
        print(f"[prep] total windows: {n}, L={l}")
    shard_id = 0
    for lo in range(0, n, shard_size):
        hi = min(lo + shard_size, n)
        path = os.path.join(out_dir, shard_name(shard_id))
        write_shard(path, windows[lo:hi], None if labels is None else labels[lo:hi],
                    num_classes=num_classes if labels is not None else 0)
        if verbose:
            print(f"[prep] wrote {path} with {hi - lo} windows")
        shard_id += 1
    t_end = time.perf_counter()
    metrics = {
        "dataset": dataset,
        "total_windows": int(n),
        "window_len": int(l),
        "shard_size_windows": int(shard_size),
        "num_shards": int(shard_id),
        "load_time_s": float(t_load - t0),
        "write_time_s": float(t_end - t_load),
        "total_time_s": float(t_end - t0),
        "timestamp": time.strftime("%Y-%m-%d %H:%M:%S"),
    }
    if verbose:
        print(f"[prep] done. shards in {out_dir} [prep] dataset loaded in "
              f"{metrics['load_time_s']:.2f}s, shards written in {metrics['write_time_s']:.2f}s, "
              f"total = {metrics['total_time_s']:.2f}s")
    if results_dir is not None:
        os.makedirs(results_dir, exist_ok=True)
        mpath = os.path.join(results_dir, "shard_prep_metrics.json")
        with open(mpath, "w") as f:
            json.dump(metrics, f, indent=2)
        if verbose:
            print(f"[prep] metrics -> {mpath}")
    return metrics
