"""Result rows and CSV persistence.

Column sets are the reference's (compat contract):
``BenchStats`` = ``Module_3/part3_mpi_gpu_train.py:64-75``; ``RoundStats`` =
``TRUE_FL_M3/part3_fedavg_overlap_mpi_gpu.py:44-55``.  ``append_results`` keeps the
align-to-existing-header behaviour (``:31-39``) without needing pandas.
"""
from __future__ import annotations

import csv
import os
import time
from dataclasses import asdict, dataclass, fields
from typing import Iterable, List, Mapping, Sequence


@dataclass
class BenchStats:
    config: str
    world_size: int
    rank: int
    batch_size: int
    steps: int
    data_ms: float
    h2d_ms: float
    compute_ms: float
    step_ms: float
    samples_per_s: float


@dataclass
class RoundStats:
    config: str
    world_size: int
    rank: int
    round_idx: int
    batch_size: int
    local_steps: int
    local_train_ms: float
    comm_ms: float
    samples_per_s: float
    avg_loss: float


@dataclass
class RoundStatsExt(RoundStats):
    """Extra device-timed columns appended after the reference's (never reordered)."""
    round_ms_dev: float = 0.0          # device time of the whole round, this rank
    agg_ms_dev: float = 0.0            # device time of the aggregation kernel
    exposed_agg_ms: float = 0.0        # round time minus the no-aggregation round time
    amp: str = "none"
    model: str = "tiny_ecg"


def columns_of(cls) -> List[str]:
    return [f.name for f in fields(cls)]


def rows_to_dicts(rows: Iterable) -> List[dict]:
    return [r if isinstance(r, Mapping) else asdict(r) for r in rows]


def append_results(rows: Sequence, path: str, retries: int = 20, delay_s: float = 0.25) -> str:
    """Append rows to ``path``; if the file exists its header decides column order/selection.

    A locked file (``PermissionError``) is retried; after ``retries`` failures the rows go to a
    time-stamped sibling file (the reference's ``safe_write_csv`` idea, ``benchmark_part_2.py:111-121``).
    """
    dicts = rows_to_dicts(rows)
    if not dicts:
        return path
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    for _ in range(max(1, retries)):
        try:
            header = None
            if os.path.exists(path) and os.path.getsize(path) > 0:
                with open(path, newline="") as f:
                    header = next(csv.reader(f), None)
            cols = header or list(dicts[0].keys())
            with open(path, "a", newline="") as f:
                w = csv.DictWriter(f, fieldnames=cols, extrasaction="ignore", restval="")
                if header is None:
                    w.writeheader()
                w.writerows(dicts)
            return path
        except PermissionError:
            time.sleep(delay_s)
    alt = f"{os.path.splitext(path)[0]}_{time.strftime('%Y%m%d_%H%M%S')}.csv"
    with open(alt, "w", newline="") as f:
        w = csv.DictWriter(f, fieldnames=list(dicts[0].keys()))
        w.writeheader()
        w.writerows(dicts)
    return alt


def read_csv(path: str) -> List[dict]:
    with open(path, newline="") as f:
        return list(csv.DictReader(f))


def summarize(rows: Sequence, key: str = "config", value: str = "samples_per_s") -> dict:
    """Per-config mean (the rank-0 summary print, ``part3_mpi_gpu_train.py:522-528``)."""
    acc: dict = {}
    for r in rows_to_dicts(rows):
        acc.setdefault(r[key], []).append(float(r[value]))
    return {k: sum(v) / len(v) for k, v in acc.items()}
