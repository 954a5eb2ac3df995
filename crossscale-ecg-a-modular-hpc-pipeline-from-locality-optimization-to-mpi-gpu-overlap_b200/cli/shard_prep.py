"""``python -m crossscale.cli.shard_prep`` — flag-compatible with ``Module_1/shard_prep.py:41-46``."""
from __future__ import annotations

import argparse

from ..data.shard_prep import DEFAULT_OUT_DIR, prepare_shards


def build_parser() -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser(description="write ecg_%05d.bin shards")
    ap.add_argument("--dataset", choices=["mitbih", "synthetic", "mitbih_like"], default="mitbih")
    ap.add_argument("--win_len", type=int, default=500)
    ap.add_argument("--stride", type=int, default=250)
    ap.add_argument("--shard_size", type=int, default=32768, help="windows per shard")
    # extensions (defaults reproduce the reference behaviour)
    ap.add_argument("--out-dir", type=str, default=DEFAULT_OUT_DIR)
    ap.add_argument("--num-windows", type=int, default=200000, help="synthetic datasets only")
    ap.add_argument("--num-classes", type=int, default=5, help="mitbih_like only")
    ap.add_argument("--results-dir", type=str, default="results")
    ap.add_argument("--seed", type=int, default=1337)
    return ap


def main(argv=None) -> dict:
    a = build_parser().parse_args(argv)
    try:
        return prepare_shards(a.dataset, a.win_len, a.stride, a.shard_size, a.out_dir, a.num_windows,
                              a.results_dir, a.seed, a.num_classes)
    except RuntimeError as e:
        if a.dataset != "mitbih":
            raise
        print(f"[prep][WARN] MIT-BIH unavailable ({e}); falling back to MIT-BIH-shaped synthetic beats")
        return prepare_shards("mitbih_like", a.win_len, a.stride, a.shard_size, a.out_dir, a.num_windows,
                              a.results_dir, a.seed, a.num_classes)


if __name__ == "__main__":
    main()
