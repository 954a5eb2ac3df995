"""Native CPU pieces: OpenMP/SIMD conv1d kernel (C13 ABI) and the C++ pinned-ring prefetcher."""
import ctypes
import glob
import os
import subprocess

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from crossscale.ops import cpu_conv
from crossscale.data.pinned_ring import LABLPrefetcher, LABLShardedReader
from crossscale.data.shard_format import load_shard

REF_C = "/root/reference/Module_2/conv1d_openmp_simd.c"


@pytest.mark.parametrize("B,L,K", [(64, 500, 3), (64, 500, 5), (128, 500, 7), (8, 500, 32), (3, 37, 9), (1, 8, 8)])
def test_conv1d_matches_torch(B, L, K):
    rng = np.random.default_rng(B * 1000 + K)
    x = rng.normal(size=(B, L)).astype(np.float32)
    w = rng.normal(size=(K,)).astype(np.float32)
    y = cpu_conv.conv1d_batch(x, w, nthreads=4)
    ref = F.conv1d(torch.from_numpy(x).unsqueeze(1), torch.from_numpy(w).view(1, 1, K)).squeeze(1).numpy()
    assert y.shape == (B, L - K + 1)
    np.testing.assert_allclose(y, ref, rtol=1e-5, atol=1e-5)
    assert cpu_conv.isa() in (0, 2, 5)


@pytest.mark.skipif(not os.path.exists(REF_C), reason="reference not mounted")
def test_conv1d_matches_unmodified_reference_kernel(tmp_path):
    so = str(tmp_path / "ref_conv1d.so")
    subprocess.check_call(["/usr/bin/gcc", "-O2", "-fopenmp", "-mavx2", "-mfma", "-shared", "-fPIC", REF_C, "-o", so])
    ref = ctypes.CDLL(so)
    ref.conv1d_batch_omp_simd.argtypes = [ctypes.c_void_p] * 3 + [ctypes.c_int] * 4
    rng = np.random.default_rng(7)
    for B, L, K in [(64, 500, 3), (32, 500, 7), (16, 500, 32)]:
        x = rng.normal(size=(B, L)).astype(np.float32)
        w = rng.normal(size=(K,)).astype(np.float32)
        y_ref = np.zeros((B, L - K + 1), np.float32)
        ref.conv1d_batch_omp_simd(x.ctypes.data, w.ctypes.data, y_ref.ctypes.data, B, L, K, 4)
        np.testing.assert_allclose(cpu_conv.conv1d_batch(x, w, 4), y_ref, rtol=1e-5, atol=1e-5)


def _drain(pf):
    out, labels, n_batches = [], [], 0
    pf.start()
    while True:
        r = pf.next_batch_cpu()
        if r is None:
            break
        slot, b, ms = r
        assert ms >= 0.0
        out.append(b.clone()); labels.append(pf.labels_of(slot, b.shape[0]).clone())
        n_batches += 1
        pf.recycle(slot)
    pf.shutdown()
    return torch.cat(out), torch.cat(labels), n_batches


def test_prefetcher_streams_all_windows_in_order(synth_shards):
    paths = sorted(glob.glob(os.path.join(synth_shards, "ecg_*.bin")))        # 700 + 700 + 600 windows
    x, _, nb = _drain(LABLPrefetcher(LABLShardedReader(paths), 256, num_slots=3, normalize=False))
    want = np.concatenate([load_shard(p) for p in paths])
    assert x.shape == (2000, 1, 500)
    np.testing.assert_array_equal(x[:, 0].numpy(), want)
    assert nb == 3 + 3 + 3                                                      # batches never span shards


def test_prefetcher_exact_multiple_does_not_stop_early(synth_shards):
    # shard sizes 700/700/600 with B=100: every shard is an exact multiple (reference bug C10)
    paths = sorted(glob.glob(os.path.join(synth_shards, "ecg_*.bin")))
    x, _, nb = _drain(LABLPrefetcher(LABLShardedReader(paths), 100, num_slots=2, normalize=False))
    assert x.shape[0] == 2000 and nb == 20


def test_prefetcher_zscore_bf16_and_labels(beat_shards):
    paths = sorted(glob.glob(os.path.join(beat_shards, "ecg_*.bin")))
    x, y, _ = _drain(LABLPrefetcher(LABLShardedReader(paths), 128, num_slots=4, normalize=True))
    assert abs(float(x.mean())) < 1e-3 and abs(float(x[:, 0].std(dim=1, unbiased=False).mean()) - 1.0) < 1e-3
    assert int(y.max()) <= 4 and int((y > 0).sum()) > 0
    xb, _, _ = _drain(LABLPrefetcher(LABLShardedReader(paths), 128, num_slots=4, normalize=True, dtype=torch.bfloat16))
    assert xb.dtype == torch.bfloat16
    torch.testing.assert_close(xb.float(), x, rtol=1e-2, atol=1e-2)


def test_reader_open_shard(synth_shards):
    paths = sorted(glob.glob(os.path.join(synth_shards, "ecg_*.bin")))
    rd = LABLShardedReader(paths)
    with rd.open_shard(paths[0]) as (mm, base, n, l):
        assert (base, n, l) == (16, 700, 500)
        first = np.frombuffer(mm, dtype=np.float32, count=l, offset=base)
        np.testing.assert_array_equal(first, load_shard(paths[0])[0])


def test_file_rendezvous_allgather(tmp_path):
    """native bootstrap: 4 'ranks' (threads) exchange 64-byte blobs through a directory"""
    import threading
    from crossscale.parallel.launch import file_allgather
    world, out = 4, {}

    def worker(r):
        out[r] = file_allgather(str(tmp_path / "rdv"), "ipc", r, world, bytes([r]) * 64, timeout_s=20)

    ts = [threading.Thread(target=worker, args=(r,)) for r in range(world)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    for r in range(world):
        assert out[r] == [bytes([q]) * 64 for q in range(world)]
    with pytest.raises(RuntimeError):
        file_allgather(str(tmp_path / "rdv2"), "x", 0, 2, b"a" * 8, timeout_s=0.2)      # peer never shows up


def test_mpi4py_shim_single_process():
    import importlib, sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline", "mpi_shim"))
    try:
        MPI = importlib.import_module("mpi4py").MPI
        comm = MPI.COMM_WORLD
        assert comm.Get_rank() == 0 and comm.Get_size() == 1
        a = np.arange(5, dtype=np.float32); b = np.empty_like(a)
        comm.Bcast(a, root=0); comm.Allreduce(a, b, op=MPI.SUM); comm.Barrier()
        np.testing.assert_array_equal(a, b)
        assert comm.gather({"x": 1}, root=0) == [{"x": 1}] and comm.allreduce(2.5, op=MPI.SUM) == 2.5
    finally:
        sys.path.pop(0)
        for k in [k for k in sys.modules if k.startswith("mpi4py")]:
            del sys.modules[k]
