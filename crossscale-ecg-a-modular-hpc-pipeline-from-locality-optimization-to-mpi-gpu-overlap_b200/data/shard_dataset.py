"""Shard assignment + dataset + GPU-resident batching.

API parity with ``Module_3/shard_dataset.py`` (``assign_shards_evenly:9-27``, ``load_shard:30-47``,
``ShardDataset:50-77``, ``make_dataloader:82-98``, ``load_shards_to_gpu:103-115``,
``make_gpu_batch_iter:118-136``).  Same call signatures and return shapes; the GPU iterator uses
the framework's permutation-gather kernel (``ops.gather``) on CUDA and stages the one-shot upload
through pinned memory on a copy stream.
"""
from __future__ import annotations

from typing import Iterator, List, Optional, Sequence, Tuple

import numpy as np
import torch
from torch.utils.data import DataLoader, Dataset

from . import shard_format as _fmt

load_shard = _fmt.load_shard


def assign_shards_evenly(shard_paths: Sequence[str], world_size: int, rank: int) -> List[str]:
    """Sorted round-robin stripe; a rank left empty wraps to ``shards[rank % len]``."""
    if len(shard_paths) == 0:
        raise RuntimeError("No shards exist on disk.")
    if not 0 <= rank < world_size:
        raise ValueError(f"rank {rank} outside world of {world_size}")
    ordered = sorted(shard_paths)
    mine = ordered[rank::world_size]
    return mine if mine else [ordered[rank % len(ordered)]]


def get_shards_for_rank(shard_paths: Sequence[str], rank: int, world_size: int) -> List[str]:
    """Older helper kept for parity (``part3_mpi_gpu_train.py:89-95``): plain stripe, may be empty."""
    return sorted(shard_paths)[rank::world_size]


class ShardDataset(Dataset):
    """Concatenation of shards: ``x [N, L] float32``, ``y [N] int64``.

    Labels come from the ``.lbl`` sidecars when every shard has one, else they are all zero
    (the reference's dummy labels, ``shard_dataset.py:70``).
    """

    def __init__(self, shard_paths: Sequence[str], max_windows: Optional[int] = None,
                 use_labels: bool = True):
        if not shard_paths:
            raise RuntimeError("No shards assigned to this rank.")
        xs, ys, have_labels, total = [], [], use_labels, 0
        self.num_classes = 0
        for p in shard_paths:
            x = _fmt.load_shard(p)
            if xs and x.shape[1] != xs[0].shape[1]:
                raise RuntimeError(f"shard {p}: window length {x.shape[1]} != {xs[0].shape[1]}")
            xs.append(x)
            lab = _fmt.load_labels(p) if have_labels else None
            if lab is None:
                have_labels = False
            else:
                ys.append(lab[0])
                self.num_classes = max(self.num_classes, lab[1])
            total += x.shape[0]
            if max_windows is not None and total >= max_windows:
                break
        arr = xs[0] if len(xs) == 1 else np.concatenate(xs, axis=0)
        arr = np.ascontiguousarray(arr[:max_windows] if max_windows is not None else arr,
                                   dtype=np.float32)
        self.x = torch.from_numpy(arr)
        if have_labels and ys:
            lab = np.concatenate(ys)[: arr.shape[0]]
            self.y = torch.from_numpy(np.ascontiguousarray(lab, dtype=np.int64))
        else:
            self.y = torch.zeros(arr.shape[0], dtype=torch.long)
        self.has_labels = bool(have_labels and ys)
        if self.y.shape[0] != self.x.shape[0]:
            raise RuntimeError(f"labels ({self.y.shape[0]}) and windows ({self.x.shape[0]}) disagree")

    def __len__(self) -> int:
        return self.x.shape[0]

    def __getitem__(self, idx):
        return self.x[idx].unsqueeze(0), self.y[idx]


def make_dataloader(shard_paths, batch_size: int, max_windows: Optional[int] = None,
                    num_workers: int = 2, pin_memory: bool = True):
    ds = ShardDataset(shard_paths, max_windows=max_windows)
    pin = bool(pin_memory and torch.cuda.is_available())
    dl = DataLoader(ds, batch_size=batch_size, shuffle=True, num_workers=num_workers,
                    pin_memory=pin, drop_last=True)
    return dl, len(ds)


def load_shards_to_gpu(shard_paths, device, max_windows: Optional[int] = None
                       ) -> Tuple[torch.Tensor, torch.Tensor]:
    """Load this rank's shards and place ``x [N, L]`` / ``y [N]`` on ``device`` once.

    On CUDA the upload goes pinned-host -> device with ``non_blocking`` copies on a side stream
    (the reference copies from pageable memory, ``shard_dataset.py:113-114``).
    """
    ds = ShardDataset(shard_paths, max_windows=max_windows)
    device = torch.device(device)
    if device.type != "cuda":
        return ds.x.to(device), ds.y.to(device)
    copy_stream = torch.cuda.Stream(device=device)
    xh, yh = ds.x.pin_memory(), ds.y.pin_memory()
    with torch.cuda.stream(copy_stream):
        xg = xh.to(device, non_blocking=True)
        yg = yh.to(device, non_blocking=True)
    torch.cuda.current_stream(device).wait_stream(copy_stream)
    xg.record_stream(torch.cuda.current_stream(device))
    yg.record_stream(torch.cuda.current_stream(device))
    copy_stream.synchronize()  # pinned staging buffers are released right after
    return xg, yg


def make_gpu_batch_iter(x_gpu: torch.Tensor, y_gpu: torch.Tensor, batch_size: int,
                        generator: Optional[torch.Generator] = None
                        ) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
    """Infinite iterator of random mini-batches on the device: ``([B,1,L], [B])``.

    A fresh permutation per epoch, ragged tail dropped (``shard_dataset.py:132-136``).
    """
    n = x_gpu.size(0)
    if n < batch_size:
        raise ValueError(f"dataset has {n} windows < batch_size {batch_size}")
    dev = x_gpu.device
    gather = None
    if dev.type == "cuda":
        from ..ops import gather as _g
        gather = _g.gather_rows
    while True:
        perm = torch.randperm(n, device=dev, generator=generator)
        for lo in range(0, n - batch_size + 1, batch_size):
            sel = perm[lo:lo + batch_size]
            if gather is not None:
                xb, yb = gather(x_gpu, y_gpu, sel)
                yield xb.unsqueeze(1), yb
            else:
                yield x_gpu[sel].unsqueeze(1), y_gpu[sel]
