"""K8: bucketed gradient all-reduce overlapped with backward (north-star extension; the reference has no
intra-round gradient exchange).

The model's gradients live in one flat fp32 arena (``models.tiny_ecg.FlatParams``); a bucket is a
contiguous slice of it.  A post-accumulate-grad hook counts the parameters of each bucket; when the
last one arrives the bucket's all-reduce is enqueued on a SIDE stream (after an event on the compute
stream), so the collective of bucket *i* runs while autograd is still producing bucket *i+1*.  Each
in-flight bucket has its own barrier channel and arena region.  The collective itself is the in-kernel
NVLink two-shot of ``csrc/cuda/fedavg.cu`` (``cs_allreduce_bucket``) — no NCCL on this path.
``finish()`` makes the compute stream wait for every bucket before the optimizer step.

The same bucket pipeline also carries the FedAvg aggregation of large models: with ``fused_op`` set for one
backward pass, a finished bucket launches ``cs_fedavg_bucket`` (optimizer step of the bucket + in-place two-shot
aggregation + bf16 shadow, ``csrc/cuda/fedavg.cu``) instead of an all-reduce, so the aggregation of bucket *i* is on
the wire while backward still computes bucket *i+1* (``backend="off"``: hooks only, no per-step all-reduce).
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.distributed as dist

from ..models.tiny_ecg import FlatParams
from ..ops import _lib
from .symm import SymmetricComm

MAX_CHANNELS = 8


class GradSync:
    def __init__(self, flat: FlatParams, bucket_bytes: int = 4 << 20, wire: str = "fp32",
                 comm: Optional[SymmetricComm] = None, backend: str = "kernel", max_blocks: Optional[int] = None,
                 inplace: bool = True):
        self.flat, self.wire_bf16 = flat, int(wire == "bf16")
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.backend = backend if self.world > 1 or backend in ("kernel", "off") else "none"
        self.fused_op = None          # callable(bucket_index, bucket, channel, side_stream): replaces the all-reduce for ONE backward
        self.timeline = None          # set to a list: (bucket, ready, start, end) timing events of every launch
        dev = flat.flat.device
        # buckets in REVERSE parameter order (gradients arrive roughly last layer first); the arena must have been
        # built with ``align=4`` so every bucket boundary is 16-byte aligned and buckets are disjoint
        if any(o % 4 for o in flat.offsets):
            raise ValueError("GradSync needs FlatParams(model, align=4)")
        self.buckets: List[dict] = []
        hi = (flat.numel + 3) // 4 * 4
        cur, members = 0, []
        for i in reversed(range(len(flat.params))):
            members.append(i)
            cur += flat.params[i].numel() * 4
            if cur >= bucket_bytes or i == 0:
                self.buckets.append({"lo_al": flat.offsets[i], "hi_al": hi, "members": list(members), "pending": len(members)})
                hi, cur, members = flat.offsets[i], 0, []
        lib = _lib.cuda_lib()
        self.channels = min(MAX_CHANNELS, max(1, len(self.buckets)))
        self.region = 0 if self.backend == "off" else max(int(lib.cs_fedavg_arena_bytes(b["hi_al"] - b["lo_al"], self.wire_bf16, self.world)) for b in self.buckets)
        self.region = (self.region + 255) // 256 * 256
        self.comm = comm
        # fp32 wire: the gradient arena itself moves into peer-mapped memory and every bucket is reduced IN PLACE (rank r
        # averages slice r of all replicas over NVLink and stores it back into every replica): one pass over HBM instead
        # of the staged kernel's encode / reduce / apply passes, which compete with backward for memory bandwidth
        self.inplace = bool(self.backend == "kernel" and not self.wire_bf16 and inplace)
        if self.backend == "kernel":
            need = flat.grad.numel() * 4 if self.inplace else self.region * self.channels
            if self.comm is None:
                self.comm = SymmetricComm(need)
            elif self.comm.arena_bytes < need:
                raise ValueError("grad-sync arena too small")
            # these collectives run under backward and their CTAs spin at the cross-GPU barriers: keep them few
            import os
            self.max_blocks = int(os.environ.get("CS_GRADSYNC_BLOCKS", "16")) if max_blocks is None else int(max_blocks)
            _lib.check(lib.cs_comm_set_max_blocks(self.comm.handle, self.max_blocks), "cs_comm_set_max_blocks")
            if self.inplace:
                import ctypes as C
                flat.rebind_grad(self.comm.arena_tensor(flat.grad.numel()))
                self._mean_w = (C.c_float * self.world)(*[1.0 / self.world] * self.world)
        self.streams = [torch.cuda.Stream(device=dev) for _ in range(self.channels)]
        self.done_events = [torch.cuda.Event() for _ in self.buckets]
        self.launches = 0
        self._owner = {}
        for j, b in enumerate(self.buckets):
            for i in b["members"]:
                self._owner[i] = j
        for i, p in enumerate(flat.params):
            p.register_post_accumulate_grad_hook(self._make_hook(i))

    def _make_hook(self, i: int):
        def hook(_p):
            j = self._owner[i]
            b = self.buckets[j]
            b["pending"] -= 1
            if b["pending"] == 0 and not b.get("launched") and (self.fused_op is not None or self.backend != "off"):
                self._launch(j)
        return hook

    def _launch(self, j: int) -> None:
        b = self.buckets[j]
        ch = j % self.channels
        side = self.streams[ch]
        timed = self.timeline is not None
        ready = torch.cuda.Event(enable_timing=timed)
        ready.record()                                     # gradients of this bucket are complete on the compute stream
        side.wait_event(ready)
        if timed:
            t0 = torch.cuda.Event(enable_timing=True)
            t0.record(side)
        g = self.flat.grad[b["lo_al"]:b["hi_al"]]
        n = b["hi_al"] - b["lo_al"]
        if self.fused_op is not None:
            self.fused_op(j, b, ch, side)
        elif self.backend == "kernel" and self.inplace:
            _lib.check(_lib.cuda_lib().cs_fedavg_aggregate_inplace(self.comm.handle, b["lo_al"] * 4, None, n, self._mean_w, 1.0, None,
                                                                   ch, side.cuda_stream), "cs_fedavg_aggregate_inplace")
        elif self.backend == "kernel":
            _lib.check(_lib.cuda_lib().cs_allreduce_bucket(self.comm.handle, g.data_ptr(), n, 1.0 / self.world, self.wire_bf16,
                                                           ch * self.region, ch, side.cuda_stream), "cs_allreduce_bucket")
        elif self.backend == "nccl" and self.world > 1:
            with torch.cuda.stream(side):
                dist.all_reduce(g)
                g.mul_(1.0 / self.world)
        if timed:
            t1 = torch.cuda.Event(enable_timing=True)
            t1.record(side)
            self.timeline.append((j, ready, t0, t1))
        self.done_events[j].record(side)
        b["launched"] = True
        self.launches += 1

    def finish(self) -> None:
        """Compute stream waits for every bucket; re-arms the hooks for the next backward."""
        cur = torch.cuda.current_stream()
        for j, b in enumerate(self.buckets):
            if b.get("launched"):
                cur.wait_event(self.done_events[j])
            elif b["pending"] not in ((0, len(b["members"])) if self.backend == "off" else (len(b["members"]),)):
                raise RuntimeError(f"gradient bucket {j}: {b['pending']} of {len(b['members'])} gradients never arrived")
            b["pending"], b["launched"] = len(b["members"]), False
