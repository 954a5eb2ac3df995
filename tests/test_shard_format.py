"""Shard container: round trip and byte compatibility with the reference reader/writer semantics
(Module_1/shard_prep.py:10-19, Module_3/shard_dataset.py:30-47)."""
import glob
import json
import os

import numpy as np
import pytest

from crossscale.data import shard_format as sf
from crossscale.data.shard_prep import prepare_shards
from crossscale.data.synth import make_mitbih_like_beats, make_synth_windows


def _reference_style_write(path, arr):
    # what the reference writer emits: two int64 then the float32 payload, native little endian
    with open(path, "wb") as f:
        f.write(np.asarray([arr.shape[0]], dtype=np.int64).tobytes())
        f.write(np.asarray([arr.shape[1]], dtype=np.int64).tobytes())
        arr.astype(np.float32).tofile(f)


def _reference_style_read(path):
    with open(path, "rb") as f:
        n, l = (int(v) for v in np.fromfile(f, dtype=np.int64, count=2))
        return np.fromfile(f, dtype=np.float32, count=n * l).reshape(n, l)


def test_round_trip(tmp_path):
    x = np.random.default_rng(0).normal(size=(37, 500)).astype(np.float32)
    p = str(tmp_path / sf.shard_name(3))
    hdr = sf.write_shard(p, x)
    assert os.path.basename(p) == "ecg_00003.bin"
    assert (hdr.num_windows, hdr.window_len) == (37, 500)
    assert os.path.getsize(p) == 16 + 37 * 500 * 4
    np.testing.assert_array_equal(sf.load_shard(p), x)
    np.testing.assert_array_equal(np.asarray(sf.load_shard(p, mmap=True)), x)


def test_byte_compat_both_directions(tmp_path):
    x = np.random.default_rng(1).normal(size=(11, 360)).astype(np.float32)
    ours, theirs = str(tmp_path / "a.bin"), str(tmp_path / "b.bin")
    sf.write_shard(ours, x)
    _reference_style_write(theirs, x)
    assert open(ours, "rb").read() == open(theirs, "rb").read()
    np.testing.assert_array_equal(_reference_style_read(ours), x)
    np.testing.assert_array_equal(sf.load_shard(theirs), x)


def test_truncated_shard_rejected(tmp_path):
    x = np.zeros((5, 8), np.float32)
    p = str(tmp_path / "t.bin")
    sf.write_shard(p, x)
    with open(p, "r+b") as f:
        f.truncate(16 + 5 * 8 * 4 - 4)
    with pytest.raises(RuntimeError):
        sf.load_shard(p)
    with open(p, "r+b") as f:
        f.truncate(7)
    with pytest.raises(RuntimeError):
        sf.read_header(p)


def test_label_sidecar(tmp_path):
    x, y = make_mitbih_like_beats(64, 360, 5, seed=3)
    p = str(tmp_path / sf.shard_name(0))
    sf.write_shard(p, x, y, num_classes=5)
    lab, nc = sf.load_labels(p)
    assert nc == 5 and lab.dtype == np.int64
    np.testing.assert_array_equal(lab, y)
    assert sf.load_labels(str(tmp_path / "missing.bin")) is None


def test_synth_generators_match_reference_distribution():
    a = make_synth_windows(100, 500, seed=1337)
    b = np.random.default_rng(1337).normal(0, 1, size=(100, 500)).astype(np.float32)
    assert a.shape == (100, 500) and a.dtype == np.float32
    # same generator family and seed: moments agree (the exact stream differs by draw dtype)
    assert abs(a.mean() - b.mean()) < 0.02 and abs(a.std() - b.std()) < 0.02
    x, y = make_mitbih_like_beats(2000, 360, 5, seed=1)
    assert x.shape == (2000, 360) and set(np.unique(y)) <= set(range(5))
    assert np.bincount(y, minlength=5)[0] > 1400          # class N dominates like MIT-BIH
    assert np.isfinite(x).all()


def test_prepare_shards_cli_outputs(tmp_path):
    m = prepare_shards("synthetic", win_len=128, shard_size=300, out_dir=str(tmp_path / "s"),
                       num_windows=1000, results_dir=str(tmp_path / "r"), verbose=False)
    files = sorted(glob.glob(str(tmp_path / "s" / "ecg_*.bin")))
    assert [os.path.basename(f) for f in files] == [f"ecg_{i:05d}.bin" for i in range(4)]
    assert [sf.read_header(f).num_windows for f in files] == [300, 300, 300, 100]
    meta = json.load(open(tmp_path / "r" / "shard_prep_metrics.json"))
    for k in ("dataset", "total_windows", "window_len", "shard_size_windows", "num_shards",
              "load_time_s", "write_time_s", "total_time_s", "timestamp"):
        assert k in meta
    assert meta["num_shards"] == 4 and meta["total_windows"] == 1000 == m["total_windows"]


def test_shard_prep_module_imports_cleanly():
    # the reference module raises NameError on import; ours must not
    import importlib
    importlib.import_module("crossscale.data.shard_prep")
    importlib.import_module("crossscale.cli.shard_prep")


def test_stale_label_sidecar_is_never_paired_with_new_windows(tmp_path):
    """re-preparing an unlabelled data set into a directory that held a labelled one must not keep the
    old .lbl files, and a sidecar whose count disagrees with the shard is ignored."""
    import numpy as np
    from crossscale.data import shard_format as fmt
    from crossscale.data.shard_dataset import ShardDataset
    p = str(tmp_path / fmt.shard_name(0))
    fmt.write_shard(p, np.zeros((100, 64), np.float32), labels=np.arange(100) % 5, num_classes=5)
    assert fmt.load_labels(p) is not None
    fmt.write_shard(p, np.ones((300, 64), np.float32))               # same name, no labels
    assert not os.path.exists(fmt.label_path_for(p))
    ds = ShardDataset([p])
    assert ds.x.shape == (300, 64) and ds.y.shape == (300,) and not ds.has_labels
    # a sidecar of the wrong length (left by another tool) is ignored with a warning
    fmt.write_shard(p, np.zeros((100, 64), np.float32), labels=np.arange(100) % 5, num_classes=5)
    lbl = open(fmt.label_path_for(p), "rb").read()
    fmt.write_shard(p, np.ones((300, 64), np.float32))
    open(fmt.label_path_for(p), "wb").write(lbl)
    with pytest.warns(UserWarning):
        assert fmt.load_labels(p) is None
