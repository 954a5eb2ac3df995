"""ctypes front end of the native CPU conv1d kernel (``csrc/cpu/conv1d_omp_simd.cpp``).

Same call shape as the reference's ``run_omp_conv`` (``Module_2/benchmark_part_2.py:48-59``):
``y[b, i] = sum_k x[b, i+k] * w[k]``, valid mode, fp32, single channel.
"""
from __future__ import annotations

import os
from typing import Optional

import numpy as np

from . import _lib


def conv1d_batch(x: np.ndarray, w: np.ndarray, nthreads: Optional[int] = None,
                 out: Optional[np.ndarray] = None) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32)
    w = np.ascontiguousarray(w, dtype=np.float32)
    b, l = x.shape
    k = w.shape[0]
    if k > l:
        raise ValueError("kernel longer than the signal")
    y = out if out is not None else np.zeros((b, l - k + 1), dtype=np.float32)
    _lib.cpu_lib().conv1d_batch_omp_simd(x.ctypes.data, w.ctypes.data, y.ctypes.data, b, l, k,
                                         int(nthreads or os.cpu_count() or 1))
    return y


def isa() -> int:
    """0 scalar, 2 AVX2+FMA, 5 AVX-512."""
    return int(_lib.cpu_lib().cs_cpu_conv1d_isa())
