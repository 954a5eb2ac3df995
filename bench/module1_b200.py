"""Module 1 on the B200: the reference's locality sweep A0-A3 (`Module_1/bench_locality.py:111-123`), the LABL ring A4
(`Module_1/train_ecg_labl(EXPERIMENTAL).py:23-94`) and the B200-native A5 row, with a clock record.
Prints one JSON document (rows + A3/A0 and A5/A0 ratios); `--out` also writes it to a file."""
import argparse
import json
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from crossscale.cli import bench_locality, train_ecg_labl
from crossscale.data.shard_prep import prepare_shards
from crossscale.utils.clocks import ClockSampler

ap = argparse.ArgumentParser()
ap.add_argument("--datasets", nargs="+", default=["synthetic", "mitbih"])
ap.add_argument("--batch-sizes", nargs="+", type=int, default=[64, 128, 256, 512])
ap.add_argument("--iters", type=int, default=100)
ap.add_argument("--num-windows", type=int, default=20000)
ap.add_argument("--out", default=None)
a = ap.parse_args()
tmp = tempfile.mkdtemp(prefix="cs_m1_")
doc = {"gpu": torch.cuda.get_device_name(0) if torch.cuda.is_available() else None, "datasets": {}}
sampler = ClockSampler(gpu_index=0).start()
keep = ("data_ms", "h2d_ms", "compute_ms", "step_ms", "samples_per_s")
for ds in a.datasets:
    res_dir = os.path.join(tmp, ds, "results")
    rows = bench_locality.main(["--dataset", ds, "--batch-sizes", *map(str, a.batch_sizes), "--iters", str(a.iters),
                                "--num-windows", str(a.num_windows), "--results-dir", res_dir])
    # A4: the pinned-ring prefetcher over shards of the same data
    kind = "synthetic" if ds == "synthetic" else "mitbih_like"
    shard_dir = os.path.join(tmp, ds, "shards")
    prepare_shards(kind, win_len=500 if ds == "synthetic" else 360, shard_size=8192, out_dir=shard_dir,
                   num_windows=a.num_windows, results_dir=None, verbose=False)
    labl = train_ecg_labl.main(["--shards", os.path.join(shard_dir, "ecg_*.bin"), "--batch-sizes", *map(str, a.batch_sizes),
                                "--iters", str(a.iters), "--results-dir", res_dir]) or []
    by = {}
    for r in list(rows) + list(labl):
        by.setdefault(int(r["batch_size"]), {})[r["config"]] = {k: round(float(r[k]), 4) for k in keep if k in r}
    ratios = {bs: {"A3_over_A0": c["A3_contig_pinned_nb"]["samples_per_s"] / c["A0_baseline"]["samples_per_s"],
                   "A5_over_A0": (c["A5_b200_staged_fused"]["samples_per_s"] / c["A0_baseline"]["samples_per_s"])
                   if "A5_b200_staged_fused" in c else None}
              for bs, c in by.items() if "A0_baseline" in c and "A3_contig_pinned_nb" in c}
    doc["datasets"][ds] = {"rows": by, "ratios": ratios}
doc["clocks"] = sampler.stop()
text = json.dumps(doc, indent=1)
print(text)
if a.out:
    os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
    open(a.out, "w").write(text)
