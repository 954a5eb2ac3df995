import os
import sys

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if REPO not in sys.path:
    sys.path.insert(0, REPO)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:  # pragma: no cover
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def synth_shards(tmp_path_factory):
    """Three small synthetic shards (reference-shaped: L=500) + metrics JSON."""
    from crossscale.data.shard_prep import prepare_shards
    root = tmp_path_factory.mktemp("shards")
    prepare_shards("synthetic", win_len=500, shard_size=700, out_dir=str(root / "data"),
                   num_windows=2000, results_dir=str(root / "results"), verbose=False)
    return str(root / "data")


@pytest.fixture(scope="session")
def beat_shards(tmp_path_factory):
    """MIT-BIH-shaped labelled beats (L=360, 5 classes)."""
    from crossscale.data.shard_prep import prepare_shards
    root = tmp_path_factory.mktemp("beats")
    prepare_shards("mitbih_like", win_len=360, shard_size=512, out_dir=str(root / "data"),
                   num_windows=1500, results_dir=None, verbose=False)
    return str(root / "data")
