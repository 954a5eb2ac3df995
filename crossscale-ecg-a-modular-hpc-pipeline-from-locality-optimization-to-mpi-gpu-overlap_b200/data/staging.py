"""K10: pinned-host shard store + double-buffered host->device staging on a copy stream.

The reference designed this overlap but left it dead (``Module_3/part3_mpi_gpu_train.py:187-305``:
``h2d_stream``, ``non_blocking`` copies of batch n+1, ``wait_stream`` before compute on batch n); its
live paths keep the whole dataset on the GPU.  Here it is real: shards are read once into
**pinned** host memory (optionally as bf16, which is exactly what autocast feeds conv1 in G1), and
each round's block of ``steps*B`` windows goes host->device with one ``cudaMemcpyAsync`` per tensor on a
dedicated copy stream while the previous block trains; events hand slots back and forth.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from ..ops import _lib
from . import shard_format as _fmt


class PinnedShardStore:
    """This client's windows in page-locked host memory: ``x [N, L]`` (fp32/bf16), ``y [N]``."""

    def __init__(self, shard_paths: Sequence[str], max_windows: Optional[int] = None,
                 dtype: torch.dtype = torch.float32, x: Optional[torch.Tensor] = None,
                 y: Optional[torch.Tensor] = None):
        pin = torch.cuda.is_available()
        if x is None:
            from .shard_dataset import ShardDataset
            ds = ShardDataset(shard_paths, max_windows=max_windows)
            x, y = ds.x, ds.y
        self.x = torch.empty(x.shape, dtype=dtype, pin_memory=pin)
        self.x.copy_(x)
        self.y = torch.empty(y.shape, dtype=torch.int64, pin_memory=pin)
        self.y.copy_(y)
        self.n, self.l = self.x.shape

    def block(self, start: int, count: int) -> Tuple[torch.Tensor, torch.Tensor]:
        return self.x[start:start + count], self.y[start:start + count]


class BlockSampler:
    """Contiguous sampling (Module-1 configs A1-A3): blocks of ``count`` windows, block order
    reshuffled every epoch, ragged tail dropped."""

    def __init__(self, n: int, count: int, seed: int = 0):
        if n < count:
            raise ValueError(f"store has {n} windows < block of {count}")
        self.nblocks, self.count = n // count, count
        self.rng = np.random.default_rng(seed)
        self.order: List[int] = []

    def next_start(self) -> int:
        if not self.order:
            self.order = list(self.rng.permutation(self.nblocks))
        return int(self.order.pop()) * self.count


class H2DStager:
    """Two device slots; ``stage`` runs on the copy stream, ``acquire``/``release`` on the compute stream."""

    def __init__(self, rows: int, l: int, dtype: torch.dtype, device, slots: int = 2):
        self.device = torch.device(device)
        self.xs = [torch.empty((rows, l), dtype=dtype, device=self.device) for _ in range(slots)]
        self.ys = [torch.empty((rows,), dtype=torch.int64, device=self.device) for _ in range(slots)]
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self.ready = [torch.cuda.Event() for _ in range(slots)]
        self.free = [torch.cuda.Event() for _ in range(slots)]
        for ev in self.free:
            ev.record()
        self.h2d_bytes = 0
        self.trace = None                # set to a list to collect (start, end) timing events of every H2D block copy

    def stage(self, slot: int, hx: torch.Tensor, hy: torch.Tensor) -> None:
        """Enqueue pinned ``hx``/``hy`` -> device slot on the copy stream (after the slot was released)."""
        lib, cs = _lib.cuda_lib(), self.copy_stream
        assert hx.dtype == self.xs[slot].dtype and hx.is_contiguous()
        cs.wait_event(self.free[slot])
        nbx, nby = hx.numel() * hx.element_size(), hy.numel() * 8
        if self.trace is not None:
            t0 = torch.cuda.Event(enable_timing=True)
            t0.record(cs)
        _lib.check(lib.cs_stage_h2d(self.xs[slot].data_ptr(), hx.data_ptr(), nbx, cs.cuda_stream), "cs_stage_h2d")
        _lib.check(lib.cs_stage_h2d(self.ys[slot].data_ptr(), hy.data_ptr(), nby, cs.cuda_stream), "cs_stage_h2d")
        if self.trace is not None:
            t1 = torch.cuda.Event(enable_timing=True)
            t1.record(cs)
            self.trace.append((t0, t1))
        self.ready[slot].record(cs)
        self.h2d_bytes = nbx + nby

    def acquire(self, slot: int) -> Tuple[torch.Tensor, torch.Tensor]:
        torch.cuda.current_stream(self.device).wait_event(self.ready[slot])
        return self.xs[slot], self.ys[slot]

    def release(self, slot: int) -> None:
        self.free[slot].record(torch.cuda.current_stream(self.device))
