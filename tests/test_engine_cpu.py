"""CPU plumbing config (BASELINE.json config 1), FedAvg maths, stats CSV, CLI flag surface,
and a 2-process gloo run of the FedAvg driver."""
import csv
import glob
import os
import subprocess
import sys

import pytest
import torch

from crossscale.data.shard_dataset import load_shards_to_gpu
from crossscale.fl import stats as S
from crossscale.fl.engine import FedAvgRunner, IndexStream, train_step_G0
from crossscale.models import TinyECG

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_single_client_cpu_batch64(beat_shards):
    """TinyECG 1D-CNN, single client on CPU, MIT-BIH-shaped beats (1x360, 5 classes), batch 64."""
    paths = sorted(glob.glob(os.path.join(beat_shards, "ecg_*.bin")))
    x, y = load_shards_to_gpu(paths, "cpu", max_windows=1024)
    torch.manual_seed(0)
    r = FedAvgRunner(TinyECG(5), "cpu", x, y, batch_size=64, local_steps=8, config="G0")
    rows = r.run(3)
    assert len(rows) == 3 and rows[0]["batch_size"] == 64 and rows[0]["world_size"] == 1
    assert rows[-1]["avg_loss"] < rows[0]["avg_loss"]           # it learns the beat classes
    assert all(k in rows[0] for k in S.columns_of(S.RoundStats))
    assert rows[0]["samples_per_s"] > 0 and rows[0]["local_train_ms"] > 0


def test_train_step_matches_manual_sgd():
    torch.manual_seed(1)
    m = TinyECG(2)
    ref = TinyECG(2)
    ref.load_state_dict(m.state_dict())
    opt = torch.optim.SGD(m.parameters(), lr=1e-2, momentum=0.9)
    x, y = torch.randn(16, 1, 500), torch.randint(0, 2, (16,))
    loss = train_step_G0(m, x, y, opt, torch.device("cpu"))
    l2 = torch.nn.functional.cross_entropy(ref(x), y)
    l2.backward()
    assert abs(loss - float(l2)) < 1e-6
    with torch.no_grad():
        for p, q in zip(m.parameters(), ref.parameters()):
            torch.testing.assert_close(p, q - 1e-2 * q.grad)


def test_index_stream_is_epoch_permutation():
    s = IndexStream(1000, 64, "cpu")
    out = torch.empty(64 * 15, dtype=torch.int64)
    s.fill(out)                                                  # exactly one epoch (15 full batches)
    assert out.unique().numel() == 960
    s.fill(out)                                                  # second epoch, new permutation
    assert out.unique().numel() == 960 and int(out.max()) < 1000


def test_weighted_fedavg_math_single_process():
    from crossscale.parallel.fedavg import FedAvgAggregator
    agg = FedAvgAggregator(10, "cpu", backend="gloo", server_lr=0.5)
    prev = torch.zeros(12)
    flat = torch.arange(12, dtype=torch.float32)
    agg.aggregate(flat, prev, weight=1.0)
    torch.testing.assert_close(flat[:10], 0.5 * torch.arange(10, dtype=torch.float32))
    torch.testing.assert_close(prev[:10], flat[:10])


def test_append_results_aligns_to_existing_header(tmp_path):
    p = str(tmp_path / "r" / "fedavg_results.csv")
    row = S.RoundStats("G0", 2, 0, 0, 256, 50, 10.0, 1.0, 1e6, 0.69)
    S.append_results([row], p)
    ext = S.RoundStatsExt("G1", 2, 1, 1, 256, 50, 9.0, 0.5, 2e6, 0.6, round_ms_dev=9.5)
    S.append_results([ext], p)                                   # extra columns dropped to fit the header
    rows = S.read_csv(p)
    assert list(rows[0].keys()) == S.columns_of(S.RoundStats)
    assert rows[1]["config"] == "G1" and float(rows[1]["comm_ms"]) == 0.5
    assert S.summarize(rows)["G0"] == 1e6
    assert S.columns_of(S.BenchStats) == ["config", "world_size", "rank", "batch_size", "steps", "data_ms",
                                          "h2d_ms", "compute_ms", "step_ms", "samples_per_s"]


def test_cli_flag_surface():
    from crossscale.cli import part3_fedavg_overlap_mpi_gpu as fa, part3_mpi_gpu_train as ps, shard_prep as sp
    a = fa.build_parser().parse_args(["--data-root", "d"])
    assert (a.batch_size, a.rounds, a.local_steps, a.max_windows, a.config) == (256, 10, 50, 30000, "both")
    b = ps.build_parser().parse_args([])
    assert (b.batch_size, b.steps, b.max_windows) == (256, 200, 20000)
    c = sp.build_parser().parse_args([])
    assert (c.dataset, c.win_len, c.stride, c.shard_size) == ("mitbih", 500, 250, 32768)
    with pytest.raises(SystemExit):
        fa.build_parser().parse_args([])                         # --data-root is required


def test_pseudo_fl_cli_cpu(synth_shards, tmp_path):
    from crossscale.cli import part3_mpi_gpu_train as ps
    rows = ps.main(["--data-root", synth_shards, "--batch-size", "32", "--steps", "3", "--max-windows", "256",
                    "--results-dir", str(tmp_path)])
    assert rows[0]["config"] == "G0_baseline_GPU_CACHE" and rows[0]["steps"] == 3
    assert os.path.exists(tmp_path / "part3_mpi_cuda_results.csv")


@pytest.mark.timeout(300)
def test_fedavg_two_clients_gloo(synth_shards, tmp_path):
    """torchrun, world_size 2, gloo: both clients end every round with the same averaged model."""
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", "29619",
           "-m", "crossscale.cli.part3_fedavg_overlap_mpi_gpu", "--data-root", synth_shards, "--batch-size", "32",
           "--rounds", "2", "--local-steps", "3", "--max-windows", "256", "--config", "G0",
           "--results-dir", str(tmp_path), "--save", str(tmp_path / "global.pt")]
    r = subprocess.run(cmd, cwd=REPO, env=env, capture_output=True, text=True, timeout=280)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    with open(tmp_path / "fedavg_results.csv") as f:
        rows = list(csv.DictReader(f))
    assert len(rows) == 4 and {r_["rank"] for r_ in rows} == {"0", "1"} and rows[0]["world_size"] == "2"
    sd = torch.load(tmp_path / "global.pt", weights_only=True)
    assert list(sd.keys())[0] == "net.0.weight"
