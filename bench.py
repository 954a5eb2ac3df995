#!/usr/bin/env python
"""Headline benchmark: TinyECG FedAvg, ECG samples/s per FL round (+ exposed aggregation ms/round).

Config (the reference's sweep config, ``TRUE_FL_M3/run_part3_sweep.sh:38-49``, on synthetic data):
TinyECG(num_classes=2), windows 1x500, batch 256 per client, 50 local SGD(1e-2, 0.9) steps per round,
config G1 (AMP), one client per GPU, FedAvg aggregation every round.  One bench "step" = one FL round
(50 local steps + aggregation).  ``value`` = whole-job samples/s = N * K * 50 * 256 / t, where t is the
device time (CUDA events, max over ranks) of K rounds INCLUDING aggregation.

    python bench.py --gpus 1 --steps 20 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # the unmodified reference under baseline/_ref
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import tempfile
import time

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

METRIC = "ECG samples/sec per FL round (TinyECG FedAvg, aggregation included)"
BATCH, LOCAL_STEPS, WIN_LEN, NUM_CLASSES = 256, 50, 500, 2
PREHEAT_ROUNDS = 304    # untimed spin-up rounds (~0.25 s) before the warm-up rounds (see run_ours)
# 131072 windows x 500 x 4 B = 262 MB per client: larger than the 126 MB L2, so the random-batch
# gather of every round comes from HBM (timing rule: inputs larger than L2).
NUM_WINDOWS = 131072


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20, help="timed FL rounds")
    ap.add_argument("--warmup", type=int, default=3, help="untimed FL rounds")
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--amp", choices=["bf16", "fp16", "none"], default="bf16")
    ap.add_argument("--agg-backend", default="auto")
    ap.add_argument("--num-windows", type=int, default=NUM_WINDOWS)
    ap.add_argument("--trace-e2e", action="store_true", help="per-rank timing breakdown of the staged rounds on stderr")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-resnet", action="store_true", help="skip the widened ResNet-1D FedAvg block (BASELINE.json config 5)")
    ap.add_argument("--no-arms", action="store_true", help="skip the extra fp16+scaler / fp32 arms measured after the headline")
    ap.add_argument("--windows", type=int, default=5, help="timed windows of --steps rounds each (value = median)")
    return ap.parse_args()


def make_shards(root: str, rank: int, n: int) -> str:
    """Synthetic Gaussian windows of the reference's shape, one shard directory per client."""
    import numpy as np
    from crossscale.data.shard_format import shard_name, write_shard
    d = os.path.join(root, f"client_{rank}")
    rng = np.random.default_rng(1337 + rank)
    per = 32768
    for i, lo in enumerate(range(0, n, per)):
        m = min(per, n - lo)
        write_shard(os.path.join(d, shard_name(i)), rng.standard_normal((m, WIN_LEN), dtype=np.float32))
    return d


_REAL_STDOUT = None


def quiet_stdout():
    """Libraries print banners to fd 1 (NCCL: "NCCL version ..." at communicator creation).  The contract is ONE JSON line on
    stdout, so fd 1 points at stderr until emit() writes that line to the saved descriptor."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(obj):
    line = json.dumps(obj) + "\n"
    sys.stdout.flush()
    if _REAL_STDOUT is None:
        sys.stdout.write(line)
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, line.encode())


# ------------------------------------------------------------------------------------------------
def _config(world, cfg, num_windows, K, W):
    """The structured part of the record: identical keys AND values in both arms (prose goes to `notes`)."""
    return {"model": f"TinyECG(num_classes={NUM_CLASSES})", "global_batch": BATCH * world, "seq_len": WIN_LEN,
            "parallelism": f"fedavg{world} (1 client/GPU)", "local_steps": LOCAL_STEPS, "batch_per_client": BATCH,
            "fl_config": cfg, "windows_per_client": num_windows, "rounds_timed": K, "warmup_rounds": W}


def _median(v):
    v = sorted(v)
    return v[len(v) // 2] if len(v) % 2 else 0.5 * (v[len(v) // 2 - 1] + v[len(v) // 2])


def run_ours(args):
    import torch
    import torch.distributed as dist
    from crossscale.data.shard_dataset import load_shards_to_gpu
    from crossscale.data.staging import PinnedShardStore
    from crossscale.fl.engine import FedAvgRunner
    from crossscale.models import TinyECG
    from crossscale.parallel import launch
    from crossscale.utils.clocks import ClockSampler
    from crossscale.utils.misc import set_basic_seeds, setup_device
    from crossscale.utils.timing import max_over_ranks
    from glob import glob

    rank, world, local_rank = launch.init_distributed()
    if not torch.cuda.is_available():
        emit({"metric": METRIC, "error": "no CUDA device", "n_gpus": 0})
        return
    device = setup_device(local_rank)
    set_basic_seeds(1234 + rank)
    K, W, R = args.steps, max(args.warmup, 3), max(args.windows, 1)
    tmp = tempfile.mkdtemp(prefix="cs_bench_")
    shard_dir = make_shards(tmp, rank, args.num_windows)
    paths = sorted(glob(os.path.join(shard_dir, "ecg_*.bin")))
    x_gpu, y_gpu = load_shards_to_gpu(paths, device=device, max_windows=args.num_windows)

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def timed_windows(fn, n_windows):
        """n_windows x (exactly K calls of fn), each bracketed by barrier + synchronize; device time, max over ranks"""
        out = []
        for _ in range(n_windows):
            sync_all()
            e0.record()
            for _ in range(K):
                fn()
            e1.record()
            sync_all()
            out.append(max_over_ranks(e0.elapsed_time(e1), device))
        return out

    def make_runner(amp):
        model = TinyECG(num_classes=NUM_CLASSES).to(device)
        cfg = "G0" if amp == "none" else "G1"
        return FedAvgRunner(model, device, x_gpu, y_gpu, BATCH, LOCAL_STEPS, cfg, amp=amp, rank=rank, world=world,
                            agg_backend=args.agg_backend), cfg

    def preheat(runner, n):
        # a fixed COUNT, not a wall-clock loop: every rank must run the same number of aggregation rounds
        for i in range(n):
            runner.run_round_async()
            if i % 16 == 15:
                torch.cuda.synchronize()

    runner, cfg = make_runner(args.amp)

    # one nvidia-smi poller for the job (rank 0's GPU), started BEFORE the pre-heat and given time to deliver its first
    # sample: its start-up (NVML init, driver queries) can stall kernel launches for milliseconds
    sampler = ClockSampler(gpu_index=torch.cuda.current_device())
    if rank == 0:
        sampler.start()
        t_wait = time.perf_counter()
        while sampler.proc is not None and not sampler.lines and time.perf_counter() - t_wait < 2.0:
            time.sleep(0.01)
    sync_all()
    preheat(runner, PREHEAT_ROUNDS)          # clock ramp from idle takes longer than W warm-up rounds
    for _ in range(W):
        runner.run_round_async()
    sync_all()

    # ---- device-timed rounds (GPU-resident data, like the reference): R windows of exactly K rounds ----------------
    launches0 = runner.trainer.launches + runner.agg.launches
    win_ms = timed_windows(runner.run_round_async, R)
    launches = (runner.trainer.launches + runner.agg.launches - launches0) // R
    dev_ms = _median(win_ms)
    runner.check_health()
    # the same rounds without the aggregation: exposed aggregation = round - no-communication round
    nocomm_ms = timed_windows(lambda: runner.run_round_async(aggregate=False), max(3, R // 2 + 1))
    exposed_ms = max(0.0, (dev_ms - _median(nocomm_ms)) / K)
    # ... and the stand-alone aggregation kernel back to back (its launch + barrier latency, nothing overlapping)
    agg_alone = timed_windows(lambda: runner.agg.aggregate(runner.flat, runner.prev, runner.weight, runner.mom), 1)[0] / K
    loss_dev = runner.trainer.state.loss_sum()
    rounds_run = runner.trainer.state.steps_done() // LOCAL_STEPS

    samples = world * K * LOCAL_STEPS * BATCH
    value = samples / (dev_ms / 1e3)

    # ---- multi-GPU correctness, outside the timed region -----------------------------------------------------------
    replicas_identical, agg_err = None, None
    if world > 1:
        flat = runner.flat.detach()
        n = runner.numel
        words = flat[:n].contiguous().view(torch.int32).to(torch.int64)
        h = torch.stack([words.sum(), (words * torch.arange(1, n + 1, device=device)).sum()])
        hs = [torch.zeros_like(h) for _ in range(world)]
        dist.all_gather(hs, h)
        replicas_identical = bool(all(torch.equal(hs[0], t) for t in hs))
        # one extra aggregation with UNEQUAL weights against the fp64 oracle sum_i w_i theta_i
        g = torch.Generator(device=device).manual_seed(99 + rank)
        prev_t = runner.prev.clone()
        loc_t = prev_t + 0.01 * torch.randn(prev_t.shape, device=device, generator=g)
        w_r = (rank + 1) / (world * (world + 1) / 2)
        locs = [torch.zeros_like(loc_t) for _ in range(world)]
        dist.all_gather(locs, loc_t)
        want = prev_t[:n].double() + sum(((r + 1) / (world * (world + 1) / 2)) * (locs[r][:n].double() - prev_t[:n].double())
                                         for r in range(world))
        runner.agg.aggregate(loc_t, prev_t, w_r, None)
        torch.cuda.synchronize()
        err = torch.tensor([float((loc_t[:n].double() - want).abs().max())], device=device, dtype=torch.float64)
        dist.all_reduce(err, op=dist.ReduceOp.MAX)
        agg_err = float(err.item())

    # ---- end-to-end rounds: pinned host -> H2D every round, loss D2H every round -------------------
    e2e = None
    if not args.no_e2e:
        dt = torch.bfloat16 if args.amp == "bf16" else torch.float32
        store = PinnedShardStore(paths, max_windows=args.num_windows, dtype=dt)
        runner.enable_e2e(store, seed=rank)
        for _ in range(max(W, 4)):
            runner.run_round_e2e(pipelined=True)
        runner.drain_e2e()
        if args.trace_e2e:
            runner.e2e_trace, runner._stager.trace = [], []
        e2e_win, e2e_dev_win, e2e_wall_win, last = [], [], [], 0.0
        for _ in range(R):
            sync_all()
            t0 = time.perf_counter()
            e0.record()
            for _ in range(K):
                last = runner.run_round_e2e(pipelined=True)      # reads an earlier round's loss while this one runs
            last = runner.drain_e2e()                            # ... and the last rounds' losses, inside the timed region
            e1.record()
            sync_all()
            wall = (time.perf_counter() - t0) * 1e3
            devt = e0.elapsed_time(e1)
            e2e_wall_win.append(wall)
            e2e_dev_win.append(devt)
            e2e_win.append(max_over_ranks(max(devt, wall), device))
        e2e_ms = _median(e2e_win)
        if args.trace_e2e:                                   # per-rank diagnostics of the staged rounds (stderr)
            tr_, cp_ = runner.e2e_trace, runner._stager.trace
            mean = lambda v: sum(v) / max(1, len(v))
            steps_ms = mean([t["ev"][0].elapsed_time(t["ev"][1]) for t in tr_])
            agg_ms_ = mean([t["ev"][1].elapsed_time(t["ev"][2]) for t in tr_])
            period = mean([a["ev"][0].elapsed_time(b["ev"][0]) for a, b in zip(tr_[:-1], tr_[1:])])
            gap = mean([a["ev"][2].elapsed_time(b["ev"][0]) for a, b in zip(tr_[:-1], tr_[1:])])
            copy_ms = mean([a.elapsed_time(b) for a, b in cp_])
            row = {"rank": rank, "period_ms": round(period, 4), "steps_ms": round(steps_ms, 4), "agg_incl_barrier_wait_ms": round(agg_ms_, 4),
                   "gap_ms": round(gap, 4), "h2d_copy_ms": round(copy_ms, 4),
                   "host_enqueue_ms": round(mean([t["host_enqueue_ms"] for t in tr_]), 4),
                   "host_total_ms": round(mean([t["host_total_ms"] for t in tr_]), 4)}
            rows_ = sorted(launch.gather_rows([row]), key=lambda r: r["rank"])
            if rank == 0:
                sys.stderr.write("[e2e trace] " + json.dumps(rows_) + "\n")
                sys.stderr.flush()
        e2e = {"value": samples / (e2e_ms / 1e3), "unit": "samples/s",
               "h2d_bytes_per_step": int(runner._stager.h2d_bytes), "d2h_bytes_per_step": int(runner.d2h_bytes),
               "ms_per_step": e2e_ms / K, "windows_ms_per_step": [round(t / K, 5) for t in e2e_win],
               "device_ms_per_step": _median(e2e_dev_win) / K, "wall_ms_per_step": _median(e2e_wall_win) / K,
               "stage_dtype": str(dt).replace("torch.", ""), "last_loss": last,
               "sampler": "contiguous blocks of 12800 windows from the pinned store, block order reshuffled per epoch",
               "loss_readback": "every round's loss is copied device->host and read by the host in order, up to two rounds "
                                "behind the enqueue (drained inside the timed region)"}
    status = runner.agg.comm.status() if getattr(runner.agg, "comm", None) is not None else 0
    fused = bool(getattr(runner, "fuse_agg", False))
    backend = runner.agg.backend
    runner.close()

    # ---- like-for-like arms in the same job: fp16 + GradScaler (the reference's G1) and fp32 (its G0) ------------
    arms = {}
    if not args.no_arms:
        for amp in ("fp16", "none", "bf16"):
            if amp == args.amp:
                continue
            name = {"fp16": "fp16_scaler", "none": "fp32", "bf16": "bf16"}[amp]
            try:                                             # the headline must survive a failure of a side measurement
                r2, cfg2 = make_runner(amp)
                preheat(r2, 48)
                ms2 = timed_windows(r2.run_round_async, 3)
                r2.check_health()
            except Exception as e:
                arms[name] = {"error": f"{type(e).__name__}: {e}"[:300]}
                continue
            arms[name] = {
                "fl_config": cfg2, "value": samples / (_median(ms2) / 1e3), "ms_per_step": _median(ms2) / K,
                "windows_ms_per_step": [round(t / K, 5) for t in ms2],
                "loss_scale": r2.trainer.state.loss_scale(), "skipped_steps": r2.trainer.state.skipped_steps(),
                "one_launch_per_round": bool(r2.trainer.use_round)}
            if amp == "none":                                # G0: fp32 storage / accumulation / conv1 / head / SGD; conv2 operands as TF32
                arms["fp32"]["conv2_operands"] = "tf32 (tcgen05.mma kind::tf32)" if r2.trainer.use_round else "fp32 FFMA"
            r2.close()
    clocks = sampler.stop()

    # ---- BASELINE.json config 5: widened ResNet-1D FedAvg (aggregation-bandwidth stress), a few rounds ------------------
    resnet = None
    if not args.no_resnet:
        import importlib.util
        spec = importlib.util.spec_from_file_location("cs_resnet_fedavg", os.path.join(REPO, "bench", "resnet_fedavg.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        try:
            resnet = mod.measure(rank, world, device, rounds=5, local_steps=2, batch=64)
        except Exception as e:                               # the headline must survive a failure of the side measurement
            resnet = {"error": f"{type(e).__name__}: {e}"[:300]}

    if rank == 0:
        base = None
        try:
            with open(os.path.join(REPO, "BASELINE.json")) as f:
                pub = json.load(f).get("published", {})
            base = pub.get("samples_per_s") if isinstance(pub, dict) else None
        except Exception:
            pass
        out = {
            "metric": METRIC, "value": value, "unit": "samples/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": (value / base) if base else None,
            "dtype": {"bf16": "bf16", "fp16": "fp16", "none": "fp32"}[args.amp], "data": "synthetic",
            "impl": "ours",
            "config": _config(world, cfg, args.num_windows, K, W),
            "windows_ms_per_step": [round(t / K, 5) for t in win_ms],
            "ms_per_step_min": min(win_ms) / K, "ms_per_step_max": max(win_ms) / K,
            "exposed_agg_ms_per_round": exposed_ms, "agg_alone_ms_per_round": agg_alone,
            "no_comm_ms_per_step": _median(nocomm_ms) / K,
            "avg_loss": loss_dev / max(1, rounds_run * LOCAL_STEPS),
            "clocks": clocks, "gpu_launches": int(launches), "comm_status": status,
            "notes": {
                "step": "one FL round = 50 local steps + FedAvg aggregation"
                        + (" in ONE cooperative kernel launch (persistent round kernel, aggregation push in its tail)" if fused else ""),
                "value": f"median of {R} windows of exactly {K} rounds; each window bracketed by barrier + cuda synchronize, "
                         "CUDA events on the launching stream, max over ranks",
                "l2": "inputs larger than L2 (262 MB resident dataset per client, random gather)",
                "sampler_resident": "per-epoch permutation of the client's windows, generated in-kernel (counter-based bijection), "
                                    "ragged tail dropped" if fused else "per-epoch torch.randperm, ragged tail dropped",
                "aggregation": backend,
                "exposed_agg": "median round with aggregation minus median round without (same kernel, comm=NULL)",
                "preheat": f"{PREHEAT_ROUNDS} untimed rounds before the {W} warm-up rounds (clock ramp from idle)"},
        }
        if replicas_identical is not None:
            out["replicas_identical"] = replicas_identical
            out["agg_max_abs_err"] = agg_err
        if arms:
            out["arms"] = arms
        if resnet is not None:
            out["resnet"] = resnet
        if e2e is not None:
            out["e2e"] = e2e
        emit(out)
    launch.shutdown()


# ------------------------------------------------------------------------------------------------
def run_reference(args):
    ref_root = os.path.join(REPO, "baseline", "_ref")
    script_dir = os.path.join(ref_root, "Module_3", "TRUE_FL_M3")
    script = os.path.join(script_dir, "part3_fedavg_overlap_mpi_gpu.py")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not os.path.exists(script):
        if rank == 0:
            emit({"impl": "reference", "unavailable": "baseline/_ref missing (run python baseline/install_reference.py; "
                  "pip cannot install the reference: no setup.py/pyproject.toml)"})
        return
    # the reference puts every rank on cuda:0 -> give each rank its own device
    os.environ["CUDA_VISIBLE_DEVICES"] = os.environ.get("LOCAL_RANK", "0")
    import runpy
    import warnings
    import torch
    if not torch.cuda.is_available():
        if rank == 0:
            emit({"impl": "reference", "unavailable": "no CUDA device"})
        return
    sys.path.insert(0, os.path.join(REPO, "baseline", "mpi_shim"))   # mpi4py is not in the image
    sys.path.insert(0, script_dir)
    K, W = args.steps, max(args.warmup, 3)
    ref_cfg = "G0" if args.amp == "none" else "G1"          # the reference's own --config: G0 = fp32, G1 = fp16 autocast + GradScaler
    # one directory for the whole job: every rank is a child of the same torchrun agent
    tmp = (tempfile.mkdtemp(prefix="cs_bench_ref_") if world == 1
           else os.path.join(tempfile.gettempdir(), f"cs_bench_ref_{os.getppid()}"))
    os.makedirs(tmp, exist_ok=True)
    # all ranks see one shard directory; the reference stripes shards over ranks itself
    import numpy as np
    from crossscale.data.shard_format import shard_name, write_shard
    shard_dir = os.path.join(tmp, "shards")
    if rank == 0:
        per = args.num_windows
        for r in range(world):
            rng = np.random.default_rng(1337 + r)
            write_shard(os.path.join(shard_dir, shard_name(r)), rng.standard_normal((per, WIN_LEN), dtype=np.float32))
        open(os.path.join(tmp, "ready"), "w").close()
    else:
        while not os.path.exists(os.path.join(tmp, "ready")):
            time.sleep(0.05)
    csv = os.path.join(ref_root, "Module_3", "results", "fedavg_results.csv")
    if rank == 0 and os.path.exists(csv):
        os.remove(csv)
    from crossscale.utils.clocks import ClockSampler
    sampler = ClockSampler(gpu_index=int(os.environ.get("LOCAL_RANK", "0"))).start()
    argv = [script, "--data-root", shard_dir, "--batch-size", str(BATCH), "--rounds", str(W + K),
            "--local-steps", str(LOCAL_STEPS), "--max-windows", str(args.num_windows), "--config", ref_cfg]
    old = sys.argv
    sys.argv = argv
    try:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            runpy.run_path(script, run_name="__main__")
    finally:
        sys.argv = old
    clocks = sampler.stop()
    if rank != 0:
        return
    import csv as _csv
    with open(csv) as f:
        rows = [r for r in _csv.DictReader(f) if r["config"] == ref_cfg and int(r["round_idx"]) >= W]
    per_rank = {}
    for r in rows:
        per_rank.setdefault(int(r["rank"]), []).append(float(r["local_train_ms"]) + float(r["comm_ms"]))
    total_ms = max(sum(v) for v in per_rank.values())
    comm_ms = max(float(r["comm_ms"]) for r in rows)
    samples = world * K * LOCAL_STEPS * BATCH
    value = samples / (total_ms / 1e3)
    emit({"metric": METRIC, "value": value, "unit": "samples/s", "n_gpus": world, "steps": K, "warmup": W,
          "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
          "dtype": "fp32" if ref_cfg == "G0" else "fp16", "data": "synthetic", "impl": "reference",
          "config": _config(world, ref_cfg, args.num_windows, K, W),
          "notes": {"step": "one FL round = 50 local steps + Bcast + Allreduce",
                    "fl_config": "G0 fp32" if ref_cfg == "G0" else "G1 (torch.cuda.amp.autocast fp16 + GradScaler)",
                    "timing": "reference's own perf_counter brackets (every step ends in cuda.synchronize), "
                              "local_train_ms + comm_ms summed over timed rounds, max over ranks",
                    "mpi": "mpi4py shim over torch.distributed gloo (host buffers), torchrun launch"},
          "exposed_agg_ms_per_round": comm_ms, "clocks": clocks,
          "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0,
                  "d2h_bytes_per_step": 4 * LOCAL_STEPS + 2 * 5832,
                  "note": "reference keeps its dataset GPU-resident; per round it reads loss.item() every step and "
                          "stages all parameters D2H+H2D twice"},
          "gpu_launches": None})


def main():
    args = parse()
    quiet_stdout()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
