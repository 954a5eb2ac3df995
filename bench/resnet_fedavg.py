"""Widened ResNet-1D (~24 M parameters) FedAvg: the aggregation-bandwidth stress config of BASELINE.json (config 5).

What replaces what: in the reference every byte of aggregation is exposed (`comm_ms`,
`Module_3/TRUE_FL_M3/part3_fedavg_overlap_mpi_gpu.py:188-190,209-214`: Bcast + Allreduce after the last local step).
Here three round schedules are timed on the same model, data and batches:

  * ``no_comm``    S local steps, nothing else (the floor)
  * ``separate``   S local steps, then ONE in-place aggregation kernel, then the bf16 shadow cast
  * ``fused``      the last step's backward launches, per gradient bucket, ONE kernel = optimizer step of the
                   bucket + in-place two-shot aggregation over NVLink + bf16 shadow (`cs_fedavg_bucket`), on side
                   streams, so only the last bucket's collective is exposed

plus the aggregation kernel alone, a BARE ``dist.all_reduce`` of the same fp32 vector (collective vs collective) and the
NVLink roofline of the in-place schedule: per GPU (P-1)/P*n*4 B come in (P2P loads) and the same go out (P2P stores);
``roofline_ms`` counts both (2*(P-1)/P*n*4 B / 770 GB/s, the per-direction peer-copy figure of B200_PROFILING.md),
``roofline_full_duplex_ms`` assumes the two directions overlap perfectly.

    torchrun --nproc-per-node N bench/resnet_fedavg.py [--rounds 5 --local-steps 2 --batch 64]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist


def conv_flops_per_sample(model, L: int) -> float:
    """forward + data gradient + weight gradient FLOPs of the conv layers for one window (2 * MACs * 3)"""
    from crossscale.ops.conv1d import Conv1dB200
    lens, hooks = {}, []

    def mk():
        def hook(m, inp, out):
            lens[m] = int(out.numel() // (out.shape[0] * m.cout))
        return hook
    for m in model.modules():
        if isinstance(m, Conv1dB200):
            hooks.append(m.register_forward_hook(mk()))
    dev = next(model.parameters()).device
    was = model.training
    with torch.no_grad():
        model.eval()
        model(torch.zeros(2, 1, L, device=dev))
    model.train(was)
    for h in hooks:
        h.remove()
    return sum(2.0 * lo * m.cin * m.cout * m.k * 3.0 for m, lo in lens.items())


def measure(rank, world, dev, rounds=5, local_steps=2, batch=64, model_name="resnet1d_wide", inplace="auto", agg_iters=10):
    from crossscale.data.synth import make_mitbih_like_beats
    from crossscale.fl.engine import FedAvgRunner
    from crossscale.models import build_model
    from crossscale.utils.clocks import ClockSampler
    from crossscale.utils.timing import max_over_ranks
    x, y = make_mitbih_like_beats(4096, 360, 5, seed=rank)
    xg, yg = torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev)

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_rounds(runner, fuse, aggregate=True):
        """median device ms of `rounds` rounds (max over ranks), one round = local_steps steps (+ aggregation)"""
        tr, ts = runner.trainer, []
        for i in range(rounds + 2):
            sync_all()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            tr.run_steps(runner.batch_iter, local_steps, fuse_last=fuse)
            if aggregate and not fuse:
                runner.agg.aggregate(runner.flat, runner.prev, runner.weight, runner.mom)
                tr.load_flat(runner.flat)
            e1.record()
            sync_all()
            if i >= 2:
                ts.append(max_over_ranks(e0.elapsed_time(e1), dev))
        ts.sort()
        return ts[len(ts) // 2]

    sampler = ClockSampler(gpu_index=torch.cuda.current_device()).start() if rank == 0 else None
    out = {"model": model_name, "world": world, "batch": batch, "local_steps": local_steps, "L": 360}
    model = build_model(model_name, 5).to(dev)
    out["params"] = sum(p.numel() for p in model.parameters())
    flops = conv_flops_per_sample(model, 360) * batch
    r = FedAvgRunner(model, dev, xg, yg, batch, local_steps, "G1", rank=rank, world=world, model_name=model_name,
                     inplace=inplace, use_graph=True, fuse_bucket_agg=True)
    n = r.numel
    for _ in range(6):                                      # eager warm-ups + both graph captures (plain step, fused last step)
        r.trainer.run_steps(r.batch_iter, 2, fuse_last=r.fuse_bucket_agg)
    sync_all()
    t_nocomm = timed_rounds(r, fuse=False, aggregate=False)
    t_sep = timed_rounds(r, fuse=False, aggregate=True)
    t_fused = timed_rounds(r, fuse=True) if r.fuse_bucket_agg else None
    step_ms = t_nocomm / local_steps
    out.update({"step_ms": step_ms, "step_tflops": flops / (step_ms * 1e-3) / 1e12,
                "round_ms_no_comm": t_nocomm, "round_ms_separate": t_sep, "round_ms_fused": t_fused,
                "exposed_agg_ms_separate": t_sep - t_nocomm,
                "exposed_agg_ms_fused": (t_fused - t_nocomm) if t_fused is not None else None,
                "fused_bucket_path": bool(r.fuse_bucket_agg), "inplace_symmetric": bool(r.agg.inplace),
                "buckets": len(r.trainer.sync.buckets) if r.trainer.sync is not None else 0})

    def timed(fn):
        for _ in range(3):
            fn()
        sync_all()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(agg_iters):
            fn()
        e1.record()
        sync_all()
        return max_over_ranks(e0.elapsed_time(e1) / agg_iters, dev)

    out["agg_kernel_ms"] = timed(lambda: r.agg.aggregate(r.flat, r.prev, r.weight, r.mom))
    buf = r.flat.clone()
    if world > 1:
        out["nccl_allreduce_bare_ms"] = timed(lambda: dist.all_reduce(buf))          # collective vs collective

        def nccl_round():                                                          # what a torch user writes around it
            torch.sub(r.flat, r.prev, out=buf)
            dist.all_reduce(buf)
            buf.mul_(1.0 / world)
        out["nccl_allreduce_with_delta_and_scale_ms"] = timed(nccl_round)
        wire = (world - 1) / world * n * 4.0
        out["nvlink_bytes_in_per_gpu"] = wire
        out["nvlink_bytes_out_per_gpu"] = wire
        out["roofline_ms"] = 2.0 * wire / 770e9 * 1e3
        out["roofline_full_duplex_ms"] = wire / 770e9 * 1e3
        out["fraction_of_nvlink_roofline"] = out["roofline_ms"] / out["agg_kernel_ms"]
        out["speedup_vs_bare_nccl"] = out["nccl_allreduce_bare_ms"] / out["agg_kernel_ms"]
    out["comm_status"] = r.agg.comm.status()
    r.close()
    if sampler is not None:
        out["clocks"] = sampler.stop()
    try:
        with open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")) as f:
            peak = json.load(f).get("bf16_tflops_sustained")
        if peak:
            out["step_fraction_of_measured_bf16_sustained"] = out["step_tflops"] / peak
    except Exception:
        pass
    return out


def main():
    from crossscale.parallel import launch
    from crossscale.utils.misc import set_basic_seeds, setup_device
    ap = argparse.ArgumentParser()
    ap.add_argument("--rounds", type=int, default=5)
    ap.add_argument("--local-steps", type=int, default=2)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--model", default="resnet1d_wide")
    ap.add_argument("--inplace", default="auto", choices=["auto", "on", "off", "nvls"])
    ap.add_argument("--agg-iters", type=int, default=10)
    a = ap.parse_args()
    rank, world, lr = launch.init_distributed()
    dev = setup_device(lr)
    set_basic_seeds(1234 + rank)
    res = measure(rank, world, dev, a.rounds, a.local_steps, a.batch, a.model, a.inplace, a.agg_iters)
    if rank == 0:
        print(json.dumps(res))
    launch.shutdown()


if __name__ == "__main__":
    main()
