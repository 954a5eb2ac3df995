"""Phase timers: CUDA events on the launching stream (device time) with a wall-clock fallback.

The reference brackets phases with ``time.perf_counter`` + ``cuda.synchronize``; here
device phases are timed with events and multi-GPU numbers are reduced as max over ranks.
"""
from __future__ import annotations

import time
from contextlib import contextmanager
from typing import Dict, List, Optional

import torch


class DeviceTimer:
    """Accumulates named phases; ``ms(name)`` synchronises once and sums event pairs."""

    def __init__(self, device: Optional[torch.device] = None):
        self.cuda = bool(device is not None and torch.device(device).type == "cuda")
        self.device = device
        self._pairs: Dict[str, List] = {}
        self._wall: Dict[str, float] = {}

    @contextmanager
    def phase(self, name: str):
        if self.cuda:
            a = torch.cuda.Event(enable_timing=True)
            b = torch.cuda.Event(enable_timing=True)
            a.record()
            try:
                yield
            finally:
                b.record()
                self._pairs.setdefault(name, []).append((a, b))
        else:
            t0 = time.perf_counter()
            try:
                yield
            finally:
                self._wall[name] = self._wall.get(name, 0.0) + (time.perf_counter() - t0) * 1e3

    def ms(self, name: str) -> float:
        if self.cuda:
            torch.cuda.synchronize(self.device)
            return float(sum(a.elapsed_time(b) for a, b in self._pairs.get(name, [])))
        return float(self._wall.get(name, 0.0))

    def reset(self):
        self._pairs.clear()
        self._wall.clear()


def max_over_ranks(value: float, device=None) -> float:
    """MAX all-reduce of a python float (no-op without an initialised process group)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    dev = device if (device is not None and dist.get_backend() == "nccl") else "cpu"
    t = torch.tensor([value], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
