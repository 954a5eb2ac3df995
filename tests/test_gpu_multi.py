"""Real multi-process runs (one GPU per rank, CUDA-IPC peer memory, in-kernel NVLink aggregation).
Skipped when the box has fewer than 2 GPUs."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys, json
sys.path.insert(0, os.environ["CS_REPO"])
import torch, torch.distributed as dist
from crossscale.parallel import launch
from crossscale.parallel.fedavg import FedAvgAggregator
from crossscale.ops import _lib
rank, world, lr = launch.init_distributed()
dev = torch.device("cuda", lr); torch.cuda.set_device(dev)
out = {}
for n, wire in ((1458, "fp32"), (1458, "bf16"), (3000001, "fp32"), (3000001, "bf16")):
    torch.manual_seed(7)                       # same prev on every rank
    prev = torch.randn(n + (-n) % 4, device=dev)
    torch.manual_seed(100 + rank)
    local = prev + 0.01 * torch.randn_like(prev)
    counts = [1000.0 * (r + 1) for r in range(world)]
    w = counts[rank] / sum(counts)
    # oracle through NCCL in fp64
    d = ((local - prev).double() * w)
    dist.all_reduce(d)
    want = (prev.double() + 0.5 * d).float()
    agg = FedAvgAggregator(n, dev, backend="kernel", wire=wire, server_lr=0.5)
    mom = torch.ones_like(prev)
    agg.reset_momentum = True
    for rep in range(3):
        l2, p2 = local.clone(), prev.clone()
        agg.aggregate(l2, p2, w, mom)
        torch.cuda.synchronize()
    err = float((l2[:n] - want[:n]).abs().max())
    same = [torch.empty_like(l2) for _ in range(world)]
    dist.all_gather(same, l2)
    ident = all(torch.equal(same[0][:n], s[:n]) for s in same)
    out[f"{n}_{wire}"] = {"err": err, "identical": ident, "status": agg.comm.status(), "mom0": float(mom[:n].abs().max())}
    agg.close()
if rank == 0:
    print("RESULT " + json.dumps(out))
launch.shutdown()
'''


def _run(world, script, tmp_path, port):
    f = tmp_path / "worker.py"
    f.write_text(script)
    env = dict(os.environ, CS_REPO=REPO)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), str(f)]
    return subprocess.run(cmd, cwd=REPO, env=env, capture_output=True, text=True, timeout=600)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_fedavg_kernel_across_processes(world, tmp_path):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    r = _run(world, WORKER, tmp_path, 29700 + world)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")][-1]
    res = json.loads(line[7:])
    for key, v in res.items():
        tol = 1e-6 if key.endswith("fp32") else 3e-4
        assert v["err"] < tol and v["identical"] and v["status"] == 0 and v["mom0"] == 0.0, (key, v)


def test_bench_two_gpus(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29731", os.path.join(REPO, "bench.py"), "--gpus", "2", "--steps", "3", "--warmup", "3",
           "--num-windows", "8192"]
    r = subprocess.run(cmd, cwd=REPO, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    out = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert out["n_gpus"] == 2 and out["value"] > 0 and out["comm_status"] == 0 and out["e2e"]["value"] > 0
