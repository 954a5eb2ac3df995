"""LABL-style loader: mmap shard reader + pinned ring + native prefetch thread.

API parity with ``Module_1/labl_loader(EXPERIMENTAL).py``: ``LABLShardedReader.open_shard`` (:7-28),
``PinnedRing`` (:30-36), ``LABLPrefetcher.{start, shutdown, next_batch_cpu, recycle}`` (:38-136).
The fill loop runs in ``libcrossscale_rt.so`` (C++ ``std::thread``; no GIL, optional bf16 output,
labels from sidecars) instead of a Python thread copying window by window.
"""
from __future__ import annotations

import ctypes as C
import mmap
import os
import struct
from contextlib import contextmanager
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from ..ops import _lib
from .shard_format import HEADER_BYTES


class LABLShardedReader:
    def __init__(self, shard_paths: Sequence[str]):
        self.paths: List[str] = list(shard_paths)
        if not self.paths:
            raise RuntimeError("no shards given")

    @contextmanager
    def open_shard(self, path: str):
        """Yields ``(mm, base, N, L)`` — the whole shard mapped read-only."""
        fd = os.open(path, os.O_RDONLY)
        mm = None
        try:
            mm = mmap.mmap(fd, length=os.path.getsize(path), access=mmap.ACCESS_READ)
            n, l = struct.unpack_from("<qq", mm, 0)
            yield mm, HEADER_BYTES, n, l
        finally:
            if mm is not None:
                try:
                    mm.close()
                except BufferError:      # a caller still holds a numpy view: the GC unmaps it later
                    pass
            os.close(fd)

    def window_len(self) -> int:
        with self.open_shard(self.paths[0]) as (_, _, _, l):
            return l


class PinnedRing:
    """``num_slots`` preallocated page-locked batches ``[B,1,L]`` (+ label slots ``[B]``)."""

    def __init__(self, num_slots: int, batch_shape: Tuple[int, int, int], dtype=torch.float32):
        pin = torch.cuda.is_available()
        self.slots = [torch.empty(batch_shape, dtype=dtype, pin_memory=pin) for _ in range(num_slots)]
        self.labels = [torch.zeros(batch_shape[0], dtype=torch.int64, pin_memory=pin)
                       for _ in range(num_slots)]


class LABLPrefetcher:
    def __init__(self, reader: LABLShardedReader, batch_size: int, num_slots: int = 4,
                 normalize: bool = True, dtype: torch.dtype = torch.float32, loop: bool = False):
        assert dtype in (torch.float32, torch.bfloat16)
        self.reader, self.B = reader, int(batch_size)
        self.L = reader.window_len()
        self.ring = PinnedRing(num_slots, (self.B, 1, self.L), dtype=dtype)
        self.normalize = bool(normalize)
        self._lib = _lib.rt_lib()
        paths = (C.c_char_p * len(reader.paths))(*[p.encode() for p in reader.paths])
        slots = (C.c_void_p * num_slots)(*[t.data_ptr() for t in self.ring.slots])
        labels = (C.c_void_p * num_slots)(*[t.data_ptr() for t in self.ring.labels])
        self._h = self._lib.cs_prefetcher_create(paths, len(reader.paths), self.B, num_slots, slots, labels,
                                                 int(self.normalize), int(dtype == torch.bfloat16), int(loop))
        if not self._h:
            raise RuntimeError(f"prefetcher: {self._lib.cs_rt_last_error().decode()}")
        self._started = False

    def start(self) -> None:
        if self._lib.cs_prefetcher_start(self._h) != 0:
            raise RuntimeError(self._lib.cs_rt_last_error().decode())
        self._started = True

    def shutdown(self) -> None:
        if self._h:
            self._lib.cs_prefetcher_destroy(self._h)
            self._h = None

    def next_batch_cpu(self) -> Optional[Tuple[int, torch.Tensor, float]]:
        """``(slot, batch_view [n,1,L], fill_ms)`` or ``None`` at end of data."""
        slot, n, ms = C.c_int(), C.c_int(), C.c_double()
        rc = self._lib.cs_prefetcher_next(self._h, C.byref(slot), C.byref(n), C.byref(ms))
        if rc < 0:
            raise RuntimeError(f"prefetcher: {self._lib.cs_rt_last_error().decode()}")
        if rc == 0:
            return None
        return slot.value, self.ring.slots[slot.value][: n.value], float(ms.value)

    def labels_of(self, slot: int, n: int) -> torch.Tensor:
        return self.ring.labels[slot][:n]

    def recycle(self, slot: int) -> None:
        self._lib.cs_prefetcher_recycle(self._h, int(slot))

    def __del__(self):  # pragma: no cover
        try:
            self.shutdown()
        except Exception:
            pass
