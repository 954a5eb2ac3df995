"""assign_shards_evenly / ShardDataset / iterators (Module_3/shard_dataset.py parity)."""
import glob
import os

import numpy as np
import pytest
import torch
from hypothesis import given, settings, strategies as st

from crossscale.data import shard_dataset as sd
from crossscale.data.shard_format import load_shard


@settings(max_examples=200, deadline=None)
@given(n=st.integers(1, 40), world=st.integers(1, 16))
def test_assign_shards_properties(n, world):
    paths = [f"/d/ecg_{i:05d}.bin" for i in np.random.default_rng(n).permutation(n)]
    ordered = sorted(paths)
    seen = []
    for r in range(world):
        mine = sd.assign_shards_evenly(paths, world, r)
        assert len(mine) >= 1                                   # every rank gets work
        stripe = ordered[r::world]
        if stripe:
            assert mine == stripe                               # sorted round-robin stripe
            seen += mine
        else:
            assert mine == [ordered[r % n]]                     # wrap-around when shards < ranks
    assert sorted(seen) == ordered                              # stripes partition the shards


def test_assign_raises_without_shards():
    with pytest.raises(RuntimeError):
        sd.assign_shards_evenly([], 2, 0)


def test_shard_dataset_concat_truncate_dummy_labels(synth_shards):
    paths = sorted(glob.glob(os.path.join(synth_shards, "ecg_*.bin")))
    ds = sd.ShardDataset(paths)
    assert len(ds) == 2000 and ds.x.shape == (2000, 500) and ds.x.dtype == torch.float32
    assert ds.y.dtype == torch.long and int(ds.y.abs().sum()) == 0      # dummy labels, all class 0
    np.testing.assert_array_equal(ds.x[:700].numpy(), load_shard(paths[0]))
    x0, y0 = ds[5]
    assert x0.shape == (1, 500) and y0.shape == ()
    assert len(sd.ShardDataset(paths, max_windows=900)) == 900


def test_labels_from_sidecar(beat_shards):
    paths = sorted(glob.glob(os.path.join(beat_shards, "ecg_*.bin")))
    ds = sd.ShardDataset(paths)
    assert ds.has_labels and ds.num_classes == 5 and ds.x.shape[1] == 360
    assert int(ds.y.max()) <= 4 and int((ds.y > 0).sum()) > 0


def test_gpu_batch_iter_on_cpu_device(synth_shards):
    paths = sorted(glob.glob(os.path.join(synth_shards, "ecg_*.bin")))
    x, y = sd.load_shards_to_gpu(paths, torch.device("cpu"), max_windows=1000)
    it = sd.make_gpu_batch_iter(x, y, 64)
    seen = set()
    for _ in range(1000 // 64):                                  # exactly one epoch, ragged tail dropped
        xb, yb = next(it)
        assert xb.shape == (64, 1, 500) and yb.shape == (64,)
        for row in xb[:, 0, :3].tolist():
            seen.add(tuple(row))
    assert len(seen) == (1000 // 64) * 64                        # a permutation: no repeats within the epoch
    xb, _ = next(it)                                             # next epoch keeps going
    assert xb.shape == (64, 1, 500)


def test_make_dataloader(synth_shards):
    paths = sorted(glob.glob(os.path.join(synth_shards, "ecg_*.bin")))
    dl, n = sd.make_dataloader(paths, batch_size=32, max_windows=200, num_workers=0, pin_memory=False)
    assert n == 200
    xb, yb = next(iter(dl))
    assert xb.shape == (32, 1, 500) and yb.dtype == torch.long
    assert len(dl) == 200 // 32                                  # drop_last


def test_loader_factories_signature():
    from crossscale.data.loaders import make_mitbih_loader, make_synth_loader
    dl = make_synth_loader(batch_size=16, num_workers=0, pin_memory=False, contiguous=True, num_windows=64)
    xb, yb = next(iter(dl))
    assert xb.shape == (16, 1, 500) and yb.shape == (16,)
    a = next(iter(make_synth_loader(16, 0, False, True, num_windows=64)))[0]
    np.testing.assert_array_equal(a.numpy(), xb.numpy())        # contiguous = sequential order
    dl2 = make_mitbih_loader(batch_size=8, num_workers=0, pin_memory=False, contiguous=False, num_windows=64)
    xb2, yb2 = next(iter(dl2))
    assert xb2.shape == (8, 1, 360) and int(yb2.max()) <= 4
