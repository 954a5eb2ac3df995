"""Widened ResNet-1D (~24 M parameters) FedAvg: the aggregation-bandwidth stress config of BASELINE.json.

Per rank: K rounds of `local_steps` local steps (tcgen05 conv kernels) + one aggregation of the 97 MB fp32
model.  Reports device-timed local ms/round, aggregation ms/round and the achieved NVLink bandwidth of the
fused two-shot kernel against a flattened NCCL all-reduce (+ scale) of the same vector and against the
roofline (P-1)/P*bytes per direction / 770 GB/s (measured peer copy bandwidth, B200_PROFILING.md).

    torchrun --nproc-per-node N bench/resnet_fedavg.py [--rounds 3 --local-steps 2 --batch 64 --wire bf16]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from crossscale.data.synth import make_mitbih_like_beats
from crossscale.fl.engine import FedAvgRunner
from crossscale.models import build_model
from crossscale.parallel import launch
from crossscale.utils.misc import set_basic_seeds, setup_device
from crossscale.utils.timing import max_over_ranks


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rounds", type=int, default=3)
    ap.add_argument("--local-steps", type=int, default=2)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--model", default="resnet1d_wide")
    ap.add_argument("--wire", default="fp32")
    ap.add_argument("--sync-grads", action="store_true")
    ap.add_argument("--inplace", default="auto", choices=["auto", "on", "off", "nvls"])
    ap.add_argument("--agg-iters", type=int, default=10)
    a = ap.parse_args()
    rank, world, lr = launch.init_distributed()
    dev = setup_device(lr)
    set_basic_seeds(1234 + rank)
    x, y = make_mitbih_like_beats(2048, 360, 5, seed=rank)
    xg, yg = torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev)
    model = build_model(a.model, 5).to(dev)
    r = FedAvgRunner(model, dev, xg, yg, a.batch, a.local_steps, "G1", rank=rank, world=world, wire=a.wire,
                     model_name=a.model, sync_grads=a.sync_grads, inplace=a.inplace)
    r.run_round(0, record=False)                                           # warm-up
    rows = [r.run_round(i + 1) for i in range(a.rounds)]
    n = r.numel
    # aggregation alone, device-timed: fused kernel vs NCCL all-reduce of the same flat vector
    def timed(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.agg_iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return max_over_ranks(e0.elapsed_time(e1) / a.agg_iters, dev)

    t_kernel = timed(lambda: r.agg.aggregate(r.flat, r.prev, r.weight, r.mom))
    buf = torch.empty_like(r.flat)

    def nccl():
        torch.sub(r.flat, r.prev, out=buf)
        if world > 1:
            dist.all_reduce(buf)
        r.flat.copy_(r.prev.add(buf, alpha=1.0 / world))
        r.prev.copy_(r.flat)

    t_nccl = timed(nccl)
    if rank == 0:
        es = 2 if a.wire == "bf16" else 4
        # bytes that cross NVLink PER DIRECTION per GPU (reduce-scatter in, all-gather out run concurrently)
        wire_bytes = (world - 1) / world * n * es if world > 1 else 0.0
        roof_ms = wire_bytes / 770e9 * 1e3
        print(json.dumps({
            "model": a.model, "params": n, "world": world, "wire": a.wire, "batch": a.batch, "local_steps": a.local_steps,
            "local_ms_per_round": sum(w["local_train_ms"] for w in rows) / len(rows),
            "samples_per_s_per_client": sum(w["samples_per_s"] for w in rows) / len(rows),
            "agg_ms_in_round": sum(w["comm_ms"] for w in rows) / len(rows),
            "agg_kernel_ms": t_kernel, "agg_nccl_ms": t_nccl, "speedup_vs_nccl": t_nccl / t_kernel,
            "nvlink_bytes_per_direction_per_gpu": wire_bytes, "inplace_symmetric": bool(r.agg.inplace), "nvls": r.agg.nvls is not None, "roofline_ms_at_770GBs": roof_ms,
            "fraction_of_nvlink_roofline": (roof_ms / t_kernel) if t_kernel > 0 and world > 1 else None,
            "avg_loss": rows[-1]["avg_loss"], "comm_status": r.agg.comm.status()}))
    r.close()
    launch.shutdown()


if __name__ == "__main__":
    main()
