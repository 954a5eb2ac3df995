"""Dataset factories for the Module-1 locality benchmark.

The reference imports ``datasets.mitbih.make_mitbih_loader`` / ``datasets.synth.make_synth_loader``
(``Module_1/bench_locality.py:97-106``) but does not ship them; they are designed here
from the call signature: ``(batch_size, num_workers, pin_memory, contiguous) -> DataLoader`` yielding
``(x [B,1,L] float32, y [B] int64)``; ``contiguous`` selects a sequential vs random sampler.
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch
from torch.utils.data import DataLoader, Dataset, RandomSampler, SequentialSampler

from .synth import make_mitbih_like_beats, make_synth_windows


class WindowDataset(Dataset):
    def __init__(self, x: np.ndarray, y: Optional[np.ndarray] = None):
        self.x = torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32))
        self.y = (torch.zeros(len(self.x), dtype=torch.long) if y is None
                  else torch.from_numpy(np.ascontiguousarray(y, dtype=np.int64)))

    def __len__(self):
        return self.x.shape[0]

    def __getitem__(self, i):
        return self.x[i].unsqueeze(0), self.y[i]


def _loader(ds, batch_size, num_workers, pin_memory, contiguous):
    sampler = SequentialSampler(ds) if contiguous else RandomSampler(ds)
    pin = bool(pin_memory and torch.cuda.is_available())
    return DataLoader(ds, batch_size=batch_size, sampler=sampler, num_workers=num_workers,
                      pin_memory=pin, drop_last=True,
                      persistent_workers=bool(num_workers > 0))


def make_synth_loader(batch_size: int, num_workers: int = 4, pin_memory: bool = False,
                      contiguous: bool = True, num_windows: int = 20000, win_len: int = 500,
                      seed: int = 1337) -> DataLoader:
    return _loader(WindowDataset(make_synth_windows(num_windows, win_len, seed)),
                   batch_size, num_workers, pin_memory, contiguous)


def make_mitbih_loader(batch_size: int, num_workers: int = 4, pin_memory: bool = False,
                       contiguous: bool = True, num_windows: int = 20000, win_len: int = 360,
                       num_classes: int = 5, seed: int = 1337, real: bool = False) -> DataLoader:
    """MIT-BIH loader.  ``real=True`` needs wfdb + network; otherwise MIT-BIH-shaped synthetic beats."""
    if real:
        from .shard_prep import make_mitbih_windows
        return _loader(WindowDataset(make_mitbih_windows(win_len=win_len)),
                       batch_size, num_workers, pin_memory, contiguous)
    x, y = make_mitbih_like_beats(num_windows, win_len, num_classes, seed)
    return _loader(WindowDataset(x, y), batch_size, num_workers, pin_memory, contiguous)
