"""Aggregation kernels alone on the 24.15 M-parameter vector (96.6 MB fp32): in-place P2P two-shot and NVLS multimem, swept
over their tuning knobs (loads in flight per thread, CTA count), next to a bare NCCL all-reduce of the same vector.
Device time per call (CUDA events, max over ranks).  NVLink floor per GPU and direction at 770 GB/s: P2P schedule
2 (P-1)/P n 4 B (own stores + the data served to the peers' loads leave on the same links); NVLS n 4 B (the slices served to
the switch's reductions + one multicast store of the own slice).

    torchrun --nproc-per-node 8 bench/agg_sweep.py [--n 24150000]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist


def main():
    from crossscale.parallel import launch
    from crossscale.parallel.fedavg import FedAvgAggregator
    from crossscale.utils.clocks import ClockSampler
    from crossscale.utils.misc import setup_device
    from crossscale.utils.timing import max_over_ranks
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=24_150_000)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--blocks", default="0,96,64", help="CTA caps to try (0 = all SMs)")
    ap.add_argument("--modes", default="on,nvls")
    a = ap.parse_args()
    rank, world, lr = launch.init_distributed()
    dev = setup_device(lr)
    n = a.n - a.n % 4
    sampler = ClockSampler(gpu_index=torch.cuda.current_device()).start() if rank == 0 else None

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn):
        for _ in range(3):
            fn()
        sync_all()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.iters):
            fn()
        e1.record()
        sync_all()
        return max_over_ranks(e0.elapsed_time(e1) / a.iters, dev)

    wire = (world - 1) / world * n * 4.0
    out = {"world": world, "n": n, "mb": n * 4 / 1e6, "iters": a.iters,
           "p2p_floor_ms": 2.0 * wire / 770e9 * 1e3, "nvls_floor_ms": (wire + n * 4.0 / world) / 770e9 * 1e3, "rows": []}
    prev = torch.zeros(n, device=dev)
    blocks = [int(b) for b in a.blocks.split(",")]
    modes = [m for m in a.modes.split(",") if m == "on" or world > 1]
    for mode in modes:
        try:
            agg = FedAvgAggregator(n, dev, inplace=mode, channel=0)
        except Exception as e:                               # no multicast mapping on this system
            out["rows"].append({"mode": mode, "error": f"{type(e).__name__}: {e}"[:200]})
            continue
        flat = agg.param_view
        flat.normal_()
        knobs = ([("CS_INPLACE_UNROLL", u, "CS_AGG_BLOCKS", b) for u in (2, 1, 3) for b in blocks] if mode == "on"
                 else [("CS_NVLS_UNROLL", u, "CS_AGG_BLOCKS", b) for u in (4, 2, 8) for b in blocks])
        for k1, v1, k2, v2 in knobs:
            os.environ[k1], os.environ[k2] = str(v1), str(v2)
            ms = timed(lambda: agg.aggregate(flat, prev, 1.0 / world))
            floor = out["p2p_floor_ms"] if mode == "on" else out["nvls_floor_ms"]
            out["rows"].append({"mode": "p2p in-place" if mode == "on" else "nvls multimem", "unroll": v1, "max_blocks": v2 or "all SMs",
                                "ms": round(ms, 4), "fraction_of_floor": round(floor / ms, 3) if world > 1 else None,
                                "status": agg.comm.status()})
            os.environ.pop(k1), os.environ.pop(k2)
        agg.close()
        del agg
    if world > 1:
        buf = torch.randn(n, device=dev)
        out["nccl_allreduce_ms"] = timed(lambda: dist.all_reduce(buf))
    if sampler is not None:
        out["clocks"] = sampler.stop()
    if rank == 0:
        print(json.dumps(out))
    launch.shutdown()


if __name__ == "__main__":
    main()
