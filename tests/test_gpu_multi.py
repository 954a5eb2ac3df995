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


@pytest.mark.parametrize("world,symm", [(2, "ipc"), (4, "ipc"), (8, "ipc"), (2, "vmm"), (8, "vmm")])
def test_fedavg_kernel_across_processes(world, symm, tmp_path, monkeypatch):
    """symm = vmm: the arenas are the repo's own fd-exported VMM allocations (parallel/vmm.py) instead of CUDA IPC handles"""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    monkeypatch.setenv("CROSSSCALE_SYMM", symm)
    r = _run(world, WORKER, tmp_path, 29700 + world + (10 if symm == "vmm" else 0))
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
           "--num-windows", "16384"]
    r = subprocess.run(cmd, cwd=REPO, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    out = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert out["n_gpus"] == 2 and out["value"] > 0 and out["comm_status"] == 0 and out["e2e"]["value"] > 0


GRAD_WORKER = r'''
import os, sys, json
sys.path.insert(0, os.environ["CS_REPO"])
import torch, torch.distributed as dist
from crossscale.parallel import launch
from crossscale.fl.engine import KernelLocalTrainer
from crossscale.models.resnet1d_wide import resnet1d_small
from crossscale.data.synth import make_mitbih_like_beats
rank, world, lr = launch.init_distributed()
dev = torch.device("cuda", lr); torch.cuda.set_device(dev)
torch.manual_seed(0)                                   # identical initial model on every rank
model = resnet1d_small(5).to(dev)
tr = KernelLocalTrainer(model, dev, lr=5e-2, sync_grads=True, bucket_bytes=8 << 10,
                        use_graph=os.environ.get("CS_TEST_GRAPH") == "1")
x, y = make_mitbih_like_beats(256, 360, 5, seed=100 + rank)   # different data per rank
x, y = torch.from_numpy(x).to(dev).unsqueeze(1), torch.from_numpy(y).to(dev)
for step in range(5):
    tr.step(x[step * 32:(step + 1) * 32], y[step * 32:(step + 1) * 32])
torch.cuda.synchronize()
flat = tr.flat.flat
allp = [torch.empty_like(flat) for _ in range(world)]
dist.all_gather(allp, flat)
ident = all(torch.equal(allp[0], p) for p in allp)      # synchronous DP keeps replicas bit-identical
# reference: same run with NCCL all-reduce of the flat gradient
torch.manual_seed(0)
m2 = resnet1d_small(5).to(dev)
tr2 = KernelLocalTrainer(m2, dev, lr=5e-2, sync_grads=False)
for step in range(5):
    m2.train(); tr2.flat.grad.zero_()
    loss = torch.nn.functional.cross_entropy(m2(x[step * 32:(step + 1) * 32]).float(), y[step * 32:(step + 1) * 32])
    loss.backward()
    dist.all_reduce(tr2.flat.grad); tr2.flat.grad.mul_(1.0 / world)
    tr2.mom.mul_(0.9).add_(tr2.flat.grad); tr2.flat.flat.add_(tr2.mom, alpha=-5e-2)
    tr2.refresh_shadow()                                 # conv layers read their weights from the bf16 shadow
err = float((flat - tr2.flat.flat).abs().max() / (tr2.flat.flat.abs().max()))
if rank == 0:
    print("RESULT " + json.dumps({"identical": ident, "rel_err_vs_nccl": err, "status": tr.sync.comm.status(),
                                  "buckets": len(tr.sync.buckets)}))
launch.shutdown()
'''


@pytest.mark.parametrize("world", [2, 8])
@pytest.mark.parametrize("graph", ["0", "1"])
def test_grad_sync_two_gpus(tmp_path, graph, world, monkeypatch):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    monkeypatch.setenv("CS_TEST_GRAPH", graph)             # "1": forward+backward+bucket all-reduces+SGD in one CUDA graph
    r = _run(world, GRAD_WORKER, tmp_path, 29741 + int(graph) + 2 * world)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    res = json.loads([l for l in r.stdout.splitlines() if l.startswith("RESULT ")][-1][7:])
    assert res["identical"] and res["status"] == 0 and res["buckets"] >= 2 and res["rel_err_vs_nccl"] < 2e-2, res


NVLS_WORKER = r'''
import os, sys, json
sys.path.insert(0, os.environ["CS_REPO"])
import torch, torch.distributed as dist
from crossscale.parallel import launch
from crossscale.parallel.fedavg import FedAvgAggregator
rank, world, lr = launch.init_distributed()
dev = torch.device("cuda", lr); torch.cuda.set_device(dev)
n = (1 << 20) + 128
try:
    agg = FedAvgAggregator(n, dev, backend="kernel", server_lr=0.5, inplace="nvls")
except Exception as e:
    if rank == 0:
        print("RESULT " + json.dumps({"skipped": str(e)[:200]}))
    launch.shutdown(); sys.exit(0)
res = {}
for equal in (True, False):
    torch.manual_seed(5)
    prev = torch.randn(n, device=dev)
    torch.manual_seed(50 + rank)
    agg.param_view.copy_(prev + 0.01 * torch.randn(n, device=dev))
    counts = [1000.0] * world if equal else [1000.0 * (r + 1) for r in range(world)]
    w = counts[rank] / sum(counts)
    agg._weights = None
    d = (agg.param_view.double() - prev.double()) * w
    dist.all_reduce(d)
    want = (prev.double() + 0.5 * d).float()
    p2 = prev.clone()
    agg.aggregate(agg.param_view, p2, w, None)
    torch.cuda.synchronize()
    dist.barrier()
    err = float((agg.param_view - want).abs().max())
    same = [torch.empty_like(agg.param_view) for _ in range(world)]
    dist.all_gather(same, agg.param_view.clone())
    res["equal" if equal else "weighted"] = {"err": err, "identical": all(torch.equal(same[0], s) for s in same),
                                             "status": agg.comm.status(),
                                             "alloc": "vmm" if isinstance(agg.nvls[0], str) else "torch"}
if rank == 0:
    print("RESULT " + json.dumps(res))
launch.shutdown()
'''


@pytest.mark.parametrize("world,alloc", [(2, "vmm"), (8, "vmm"), (2, "torch")])
def test_fedavg_nvls_multimem_two_gpus(tmp_path, world, alloc, monkeypatch):
    """alloc = vmm: the multicast mapping comes from the repo's own allocator (cuMulticastCreate / BindMem through
    parallel/vmm.py); torch: the torch symmetric-memory fallback"""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    monkeypatch.setenv("CROSSSCALE_NVLS_ALLOC", alloc)
    r = _run(world, NVLS_WORKER, tmp_path, 29751 + world + (30 if alloc == "torch" else 0))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    res = json.loads([l for l in r.stdout.splitlines() if l.startswith("RESULT ")][-1][7:])
    if "skipped" in res:
        pytest.skip("no NVLS multicast here: " + res["skipped"])
    for key, v in res.items():
        assert v["err"] < 2e-6 and v["identical"] and v["status"] == 0 and v["alloc"] == alloc, (key, v)


ROUND_WORKER = r'''
import os, sys, json
sys.path.insert(0, os.environ["CS_REPO"])
import torch, torch.distributed as dist
from crossscale.parallel import launch
from crossscale.fl.engine import FedAvgRunner
from crossscale.models import TinyECG
from crossscale.ops import tiny_ecg as tk
rank, world, lr = launch.init_distributed()
dev = torch.device("cuda", lr); torch.cuda.set_device(dev)
B, S, L, N = 64, 6, 500, 1000 + 100 * rank                 # unequal client sizes -> unequal FedAvg weights
g = torch.Generator(device="cuda").manual_seed(100 + rank)
x = torch.randn(N, L, device=dev, generator=g); y = torch.randint(0, 2, (N,), device=dev, generator=g)
torch.manual_seed(0)
model = TinyECG(2).to(dev)
gen = torch.Generator(device="cuda").manual_seed(500 + rank)
AMP = os.environ.get("CS_TEST_AMP", "bf16")                  # "none": G0 = the tf32 tensor-core round
runner = FedAvgRunner(model, dev, x, y, B, S, "G0" if AMP == "none" else "G1", amp=AMP, rank=rank, world=world, generator=gen)
assert runner.fuse_agg, "the persistent round kernel with the fused aggregation tail must be the path here"
res = {"rounds": []}
start = runner.flat.clone()
mom_prev = None
for rnd in range(3):
    before = runner.prev.clone()
    pos = runner.trainer.state.samples_drawn()
    runner.run_round_async()                                # ONE launch: S local steps + in-kernel FedAvg push
    torch.cuda.synchronize()
    runner.check_health()
    # oracle: the same batches through the single-launch step kernel, then the fp64 weighted mean over ranks (NCCL)
    rows = torch.tensor(tk.perm_batches(pos, S * B, N, B, runner.trainer.seed), device=dev)
    solo = tk.TinyTrainState(before.clone(), 2, amp=tk.AMP_BY_NAME[AMP])
    if mom_prev is not None:
        solo.mom.copy_(mom_prev)
    for s_ in range(S):
        solo.step(x, y, rows[s_ * B:(s_ + 1) * B], batch=B)
    torch.cuda.synchronize()
    mom_prev = solo.mom.clone()
    n = runner.numel
    d = (solo.params[:n].double() - before[:n].double()) * runner.weight
    dist.all_reduce(d)
    want = before[:n].double() + d
    err = float((runner.flat[:n].double() - want).norm() / want.norm())
    allp = [torch.empty_like(runner.flat) for _ in range(world)]
    dist.all_gather(allp, runner.flat)
    res["rounds"].append({"rel_err": err, "identical": all(torch.equal(allp[0][:n], p[:n]) for p in allp),
                          "prev_ok": bool(torch.equal(runner.prev[:n], runner.flat[:n])),
                          "mom_err": float((runner.mom[:n] - solo.mom[:n]).abs().max())})
res["status"] = runner.agg.comm.status()
res["moved"] = float((runner.flat - start).abs().max())
res["weight"] = runner.weight
if rank == 0:
    print("RESULT " + json.dumps(res))
runner.close()
launch.shutdown()
'''


@pytest.mark.parametrize("world", [2, 8])
@pytest.mark.parametrize("amp", ["bf16", "none"])
def test_persistent_round_with_fused_aggregation_across_processes(tmp_path, world, amp, monkeypatch):
    """The headline path on real peers: each rank's round is ONE cooperative launch whose tail pushes the weighted
    delta over NVLink; checked per round against the step kernel + an fp64 NCCL mean, replicas bit-identical.
    amp = none: the G0 round (fp32 storage, conv2 operands as TF32) against the fp32 FFMA step kernel."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    monkeypatch.setenv("CS_TEST_AMP", amp)
    r = _run(world, ROUND_WORKER, tmp_path, 29761 + world + (20 if amp == "none" else 0))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    res = json.loads([l for l in r.stdout.splitlines() if l.startswith("RESULT ")][-1][7:])
    assert res["status"] == 0 and res["moved"] > 0
    tol, mom_tol = (5e-5, 1e-3) if amp == "bf16" else (3e-4, 5e-3)        # tf32 operands vs exact fp32 products
    for rd in res["rounds"]:
        assert rd["identical"] and rd["prev_ok"] and rd["rel_err"] < tol and rd["mom_err"] < mom_tol, res


BUCKET_WORKER = r'''
import os, sys, json
sys.path.insert(0, os.environ["CS_REPO"])
import torch, torch.distributed as dist
from crossscale.parallel import launch
from crossscale.fl.engine import FedAvgRunner
from crossscale.models.resnet1d_wide import resnet1d_small
from crossscale.data.synth import make_mitbih_like_beats
rank, world, lr = launch.init_distributed()
dev = torch.device("cuda", lr); torch.cuda.set_device(dev)
x, y = make_mitbih_like_beats(512, 360, 5, seed=100 + rank)
x, y = torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev)
out = {}
finals = {}
for fused in (True, False):
    torch.manual_seed(0)
    model = resnet1d_small(5).to(dev)
    gen = torch.Generator(device="cuda").manual_seed(77 + rank)
    runner = FedAvgRunner(model, dev, x, y, 32, 3, "G1", amp="bf16", rank=rank, world=world, generator=gen,
                          model_name="resnet1d_small", inplace="on", fuse_bucket_agg=fused,
                          use_graph=(os.environ.get("CS_TEST_GRAPH") == "1"), bucket_bytes=8 << 10)
    assert runner.fuse_bucket_agg == fused
    for rnd in range(4):
        runner.run_round(rnd)
    flat = runner.flat.clone()
    allp = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(allp, flat)
    finals[fused] = flat
    out["fused" if fused else "plain"] = {"identical": all(torch.equal(allp[0], p) for p in allp), "status": runner.agg.comm.status(),
                                          "shadow_ok": bool(torch.equal(runner.trainer.shadow, flat.to(torch.bfloat16))),
                                          "loss": runner.rows[-1]["avg_loss"], "agg_launches": runner.agg.launches}
    runner.close()
out["rel_diff"] = float((finals[True] - finals[False]).norm() / finals[False].norm())
if rank == 0:
    print("RESULT " + json.dumps(out))
launch.shutdown()
'''


@pytest.mark.parametrize("world", [2, 8])
@pytest.mark.parametrize("graph", ["0", "1"])
def test_bucket_fused_fedavg_matches_separate_kernels(tmp_path, world, graph, monkeypatch):
    """ResNet-1D FedAvg: the round whose last backward feeds per-bucket {optimizer step + in-place aggregation + bf16
    shadow} launches must end where the round with separate SGD / aggregation / shadow-cast kernels ends."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    monkeypatch.setenv("CS_TEST_GRAPH", graph)
    r = _run(world, BUCKET_WORKER, tmp_path, 29781 + world + int(graph))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    res = json.loads([l for l in r.stdout.splitlines() if l.startswith("RESULT ")][-1][7:])
    for k in ("fused", "plain"):
        assert res[k]["identical"] and res[k]["status"] == 0 and res[k]["shadow_ok"], res
    assert res["fused"]["agg_launches"] > res["plain"]["agg_launches"]       # one launch per bucket vs one per round
    assert res["rel_diff"] < 2e-2, res                                       # bf16 training noise over 12 steps
