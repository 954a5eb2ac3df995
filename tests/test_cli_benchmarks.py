"""Module-1/2 benchmark drivers and the result aggregation (CPU)."""
import glob
import os

from crossscale.cli import bench_locality, benchmark_part_2, plots, train_cpu_openmp, train_ecg_labl
from crossscale.fl.stats import append_results, read_csv


def test_bench_locality_cpu(tmp_path):
    rows = bench_locality.main(["--dataset", "synthetic", "--device", "cpu", "--batch-sizes", "16", "--iters", "3",
                                "--num-workers", "0", "--results-dir", str(tmp_path), "--num-windows", "256"])
    assert [r["config"] for r in rows] == ["A0_baseline", "A1_contiguous", "A2_contig_pinned", "A3_contig_pinned_nb"]
    assert rows[0]["contiguous"] is False and rows[1]["contiguous"] is True      # A0 != A1 (reference: A0 == A1)
    got = read_csv(tmp_path / "part1_locality_results.csv")
    assert list(got[0].keys()) == bench_locality.COLUMNS and len(got) == 4
    assert all(float(r["samples_per_s"]) > 0 for r in got)
    assert os.path.exists(tmp_path / "part1_locality_summary.csv")
    a = bench_locality.build_parser().parse_args([])
    assert (a.dataset, a.device, a.batch_sizes, a.iters, a.num_workers) == ("mitbih", "cuda", [64, 128, 256, 512], 100, 4)


def test_labl_benchmark_cpu(beat_shards, tmp_path):
    rows = train_ecg_labl.main(["--shards", os.path.join(beat_shards, "ecg_*.bin"), "--batch-sizes", "32", "--iters", "4",
                                "--device", "cpu", "--results-dir", str(tmp_path)])
    assert rows[0]["config"] == "A4_LABL" and rows[0]["samples_per_s"] > 0 and rows[0]["data_ms"] >= 0
    assert list(read_csv(tmp_path / "part1_labl_results.csv")[0].keys()) == train_ecg_labl.COLUMNS
    a = train_ecg_labl.build_parser().parse_args([])
    assert a.shards == "data/shards/ecg_*.bin" and a.iters == 100


def test_part2_benchmarks(tmp_path):
    rows = benchmark_part_2.main(["--results-dir", str(tmp_path), "--trials", "2", "--batch-sizes", "64", "--kernel-sizes", "3", "7"])
    assert len(rows) == 2 and rows[0]["speedup_med"] > 0
    cols = list(read_csv(tmp_path / "part2_openmp_results.csv")[0].keys())
    for c in ("batch_size", "kernel_size", "nthreads", "torch_ms_median", "torch_ms_mean", "torch_ms_std", "torch_ms_p95",
              "omp_ms_median", "omp_ms_mean", "omp_ms_std", "omp_ms_p95", "torch_sps", "omp_sps", "speedup_med"):
        assert c in cols
    raw = read_csv(tmp_path / "part2_openmp_results_raw.csv")
    assert len(raw) == 4 and {"batch_size", "kernel_size", "trial", "torch_ms", "omp_ms"} <= set(raw[0].keys())
    rows = train_cpu_openmp.main(["--results-dir", str(tmp_path), "--iters", "3", "--threads", "1", "2", "--batches", "64"])
    assert list(rows[0].keys()) == ["threads", "batch", "compute_ms", "samples_per_s"]


def test_plot_aggregation(tmp_path):
    rd = str(tmp_path)
    append_results([{"config": "A3_contig_pinned_nb", "batch_size": 512, "pin_memory": True, "contiguous": True,
                     "non_blocking": True, "data_ms": 1.0, "h2d_ms": 0.5, "compute_ms": 2.0, "step_ms": 3.5,
                     "samples_per_s": 1000.0}], os.path.join(rd, "part1_locality_results.csv"))
    append_results([{"config": "A4_LABL", "batch_size": 512, "step_ms": 2.0, "samples_per_s": 2000.0, "data_ms": 0.5,
                     "h2d_ms": 0.5, "compute_ms": 1.0}], os.path.join(rd, "part1_labl_results.csv"))
    import json
    json.dump({"total_time_s": 10.0, "total_windows": 200000}, open(os.path.join(rd, "shard_prep_metrics.json"), "w"))
    merged = plots.plot_all_results(rd)
    a4 = [r for r in merged if r["config"] == "A4_LABL"][0]
    want = 2000.0 / (1 + 10.0 / (10 * 200000 / 2000.0))
    assert abs(a4["effective_samples_per_s"] - want) < 1e-6 and a4["shard_ms_per_step"] > 0
    assert plots.effective_throughput(0, 1, 1) == 0.0
    # FedAvg summary: mean over ranks/rounds, step = local + comm, config canonicalised
    fl = [{"config": "G1", "world_size": 2, "rank": r, "round_idx": i, "batch_size": 256, "local_steps": 50,
           "local_train_ms": 10.0 + r, "comm_ms": 1.0, "samples_per_s": 100.0, "avg_loss": 0.5} for r in (0, 1) for i in (0, 1)]
    append_results(fl, os.path.join(rd, "fedavg_results.csv"))
    t = plots.plot_part3_fl(rd)
    assert t[0]["config"] == "G1_overlap_amp" and abs(t[0]["step_ms"] - 11.5) < 1e-9
    assert plots.canonical_config("G0_baseline_GPU_CACHE") == "G0_baseline"
    assert plots.main(["part3_fl", "--results-dir", rd]) is not None
