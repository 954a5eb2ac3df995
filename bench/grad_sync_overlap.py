"""K8: how much of the bucketed gradient all-reduce hides under backward?  (north-star extension; the reference has no
intra-round gradient exchange)

Per rank, widened ResNet-1D, one training step = forward + backward + flat SGD:
  * sync off                      (no communication)
  * sync on, in-kernel NVLink two-shot per bucket on side streams (`cs_allreduce_bucket`, parallel/grad_sync.py)
  * sync on, NCCL all-reduce per bucket on the same side streams
For each: median device ms per step (max over ranks) and, for the two sync arms, the per-bucket CUDA-event timeline of one
eager step: when the bucket's gradients were ready, when its all-reduce started and ended, all relative to the end of
backward.  exposed_ms = step(on) - step(off); hidden fraction = 1 - exposed / sum of the buckets' all-reduce durations.

    torchrun --nproc-per-node N bench/grad_sync_overlap.py [--batch 256 --bucket-mb 4]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist


def main():
    from crossscale.fl.engine import KernelLocalTrainer
    from crossscale.models import build_model
    from crossscale.parallel import launch
    from crossscale.parallel.grad_sync import GradSync
    from crossscale.utils.clocks import ClockSampler
    from crossscale.utils.misc import setup_device
    from crossscale.utils.timing import max_over_ranks
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--bucket-mb", type=float, default=4.0)
    ap.add_argument("--model", default="resnet1d_wide")
    a = ap.parse_args()
    rank, world, lr = launch.init_distributed()
    dev = setup_device(lr)
    torch.manual_seed(1 + rank)
    x = torch.randn(a.batch, 1, 360, device=dev)
    y = torch.randint(0, 5, (a.batch,), device=dev)
    sampler = ClockSampler(gpu_index=torch.cuda.current_device()).start() if rank == 0 else None

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def make(mode):
        torch.manual_seed(0)
        m = build_model(a.model, 5).to(dev)
        tr = KernelLocalTrainer(m, dev, sync_grads=False, use_graph=False, bucket_bytes=int(a.bucket_mb * (1 << 20)))
        if mode != "off":
            tr.sync = GradSync(tr.flat, bucket_bytes=int(a.bucket_mb * (1 << 20)), backend=mode)
            out["kernel_max_blocks"] = getattr(tr.sync, "max_blocks", None) if mode == "kernel" else out.get("kernel_max_blocks")
        return tr

    def step_ms(tr):
        for _ in range(4):
            tr.step(x, y)
        ts = []
        for _ in range(a.steps):
            sync_all()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            tr.step(x, y)
            e1.record()
            sync_all()
            ts.append(max_over_ranks(e0.elapsed_time(e1), dev))
        ts.sort()
        return ts[len(ts) // 2]

    out = {"model": a.model, "world": world, "batch": a.batch, "bucket_mb": a.bucket_mb, "note": "eager launches (no whole-step CUDA graph) in all three arms"}
    off = make("off")
    out["step_ms_sync_off"] = step_ms(off)
    del off
    for mode in (("kernel", "nccl") if world > 1 else ("kernel",)):
        tr = make(mode)
        ms = step_ms(tr)
        # one more step with the per-bucket timeline switched on
        tr.sync.timeline = []
        sync_all()
        bwd_end = torch.cuda.Event(enable_timing=True)
        tr.model.train(); tr.flat.grad.zero_()
        from crossscale.ops.ce_scaler import ScaledCrossEntropy
        loss = ScaledCrossEntropy.apply(tr.model(x), y, None)
        loss.backward()
        bwd_end.record()
        tr.sync.finish()
        torch.cuda.synchronize()
        rows, total = [], 0.0
        for j, ready, t0, t1 in tr.sync.timeline:
            b = tr.sync.buckets[j]
            dur = t0.elapsed_time(t1)
            total += dur
            rows.append({"bucket": j, "mb": round((b["hi_al"] - b["lo_al"]) * 4 / 2 ** 20, 2),
                         "ready_ms_before_bwd_end": round(ready.elapsed_time(bwd_end), 3),
                         "start_ms_before_bwd_end": round(t0.elapsed_time(bwd_end), 3),
                         "end_ms_after_bwd_end": round(bwd_end.elapsed_time(t1), 3), "allreduce_ms": round(dur, 3)})
        tr.sync.timeline = None
        exposed = ms - out["step_ms_sync_off"]
        out[mode] = {"step_ms": ms, "exposed_ms": exposed, "sum_bucket_allreduce_ms": total,
                     "exposed_fraction_of_allreduce_time": (exposed / total) if total > 0 else None,
                     "buckets": len(tr.sync.buckets), "timeline_rank0": rows,
                     "comm_status": tr.sync.comm.status() if mode == "kernel" else 0}
        del tr
    if sampler is not None:
        out["clocks"] = sampler.stop()
    if rank == 0:
        print(json.dumps(out))
    launch.shutdown()


if __name__ == "__main__":
    main()
