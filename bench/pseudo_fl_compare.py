"""Pseudo-FL driver (independent replicas, no model exchange), reference vs this repo on the same GPU and shards.

Reference arm: the UNMODIFIED `Module_3/part3_mpi_gpu_train.py` from `baseline/_ref` (`main()`, `:420-528`), run as the
package module it is written as (relative imports) under the `mpi4py` -> torch.distributed shim; its shard root is
hard-coded to `<two levels up>/PART-1/data/shards` (`:441-445`), so the script is given a scratch tree of that shape.
Our arm: `python -m crossscale.cli.part3_mpi_gpu_train` with the same flags and the same shard files.
Metric (both): `samples_per_s = batch / mean(step_ms)` per config (`part3_mpi_gpu_train.py:170-171`), G0 = fp32 baseline,
G1 = AMP config.  One rank (the drivers exchange nothing, N ranks are N independent copies).

    python bench/pseudo_fl_compare.py [--steps 200 --batch-size 256 --max-windows 20000]
"""
import argparse
import csv
import json
import os
import shutil
import subprocess
import sys
import tempfile

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)


def read_rows(path):
    with open(path) as f:
        return list(csv.DictReader(f))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--batch-size", type=int, default=256)
    ap.add_argument("--max-windows", type=int, default=20000)
    ap.add_argument("--repeats", type=int, default=3)
    a = ap.parse_args()
    import numpy as np
    from crossscale.data.shard_format import shard_name, write_shard
    from crossscale.utils.clocks import ClockSampler

    tmp = tempfile.mkdtemp(prefix="cs_pseudo_fl_")
    shard_dir = os.path.join(tmp, "PART-1", "data", "shards")
    os.makedirs(shard_dir)
    rng = np.random.default_rng(1337)
    for i in range(2):
        write_shard(os.path.join(shard_dir, shard_name(i)), rng.standard_normal((16384, 500), dtype=np.float32))
    flags = ["--batch-size", str(a.batch_size), "--steps", str(a.steps), "--max-windows", str(a.max_windows)]
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29877", RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
    out = {"steps": a.steps, "batch": a.batch_size, "max_windows": a.max_windows, "L": 500, "ranks": 1, "repeats": a.repeats}
    sampler = ClockSampler(gpu_index=0).start()

    # ---- reference ----------------------------------------------------------------------------------------------
    ref_src = os.path.join(REPO, "baseline", "_ref", "Module_3")
    if os.path.isdir(ref_src):
        pkg_root = os.path.join(tmp, "proj")                 # <tmp>/proj/Module_3/*.py ; PROJECT_ROOT resolves to <tmp>
        shutil.copytree(ref_src, os.path.join(pkg_root, "Module_3"))
        open(os.path.join(pkg_root, "Module_3", "__init__.py"), "a").close()
        code = ("import sys, runpy; sys.path.insert(0, %r); sys.path.insert(0, %r); sys.argv = ['part3_mpi_gpu_train'] + %r; "
                "runpy.run_module('Module_3.part3_mpi_gpu_train', run_name='__main__')"
                % (os.path.join(REPO, "baseline", "mpi_shim"), pkg_root, flags))
        for _ in range(a.repeats):
            r = subprocess.run([sys.executable, "-c", code], env=env, cwd=pkg_root, capture_output=True, text=True, timeout=900)
            if r.returncode != 0:
                out["reference_error"] = (r.stderr or r.stdout)[-600:]
                break
        ref_csv = os.path.join(pkg_root, "results", "part3_mpi_cuda_results.csv")
        if os.path.exists(ref_csv):
            rows = read_rows(ref_csv)
            for cfg in sorted({r["config"] for r in rows}):
                v = sorted(float(r["samples_per_s"]) for r in rows if r["config"] == cfg)
                s = sorted(float(r["step_ms"]) for r in rows if r["config"] == cfg)
                out.setdefault("reference", {})[cfg] = {"samples_per_s_median": v[len(v) // 2], "step_ms_median": s[len(s) // 2], "runs": len(v)}
    else:
        out["reference_error"] = "baseline/_ref missing (python baseline/install_reference.py)"

    # ---- ours ---------------------------------------------------------------------------------------------------
    res_dir = os.path.join(tmp, "ours_results")
    for amp in ("bf16", "fp16"):
        for _ in range(a.repeats):
            r = subprocess.run([sys.executable, "-m", "crossscale.cli.part3_mpi_gpu_train", "--data-root", shard_dir, "--amp", amp,
                                "--results-dir", os.path.join(res_dir, amp)] + flags, env=env, cwd=REPO, capture_output=True, text=True, timeout=900)
            if r.returncode != 0:
                out["ours_error"] = (r.stderr or r.stdout)[-600:]
                break
        p = os.path.join(res_dir, amp, "part3_mpi_cuda_results.csv")
        if os.path.exists(p):
            rows = read_rows(p)
            for cfg in sorted({r["config"] for r in rows}):
                v = sorted(float(r["samples_per_s"]) for r in rows if r["config"] == cfg)
                s = sorted(float(r["step_ms"]) for r in rows if r["config"] == cfg)
                out.setdefault("ours", {})[f"{cfg} (amp={amp})" if cfg.startswith("G1") else cfg] = {
                    "samples_per_s_median": v[len(v) // 2], "step_ms_median": s[len(s) // 2], "runs": len(v)}
    out["clocks"] = sampler.stop()
    shutil.rmtree(tmp, ignore_errors=True)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
