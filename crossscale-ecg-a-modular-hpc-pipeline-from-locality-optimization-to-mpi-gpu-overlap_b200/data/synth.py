"""Synthetic window generators (there is no network for PhysioNet in this environment).

* :func:`make_synth_windows` — Gaussian windows, same distribution and seed as the reference's
  ``--dataset synthetic`` path (``Module_1/shard_prep.py:35-37``).
* :func:`make_mitbih_like_beats` — MIT-BIH-shaped beat windows (1 x 360 samples at 360 Hz, 5 AAMI
  classes N/S/V/F/Q) named by ``BASELINE.json`` configs; a P-QRS-T sum-of-Gaussians morphology
  whose widths/amplitudes depend on the class, plus baseline wander and noise.
"""
from __future__ import annotations

from typing import Tuple

import numpy as np

AAMI_CLASSES = ("N", "S", "V", "F", "Q")
AAMI_PRIORS = (0.83, 0.03, 0.07, 0.01, 0.06)


def make_synth_windows(N: int = 20000, L: int = 500, seed: int = 1337) -> np.ndarray:
    rng = np.random.default_rng(seed)
    return rng.standard_normal(size=(N, L), dtype=np.float32)


def _gauss(t, mu, sigma, amp):
    return amp[:, None] * np.exp(-0.5 * ((t[None, :] - mu[:, None]) / sigma[:, None]) ** 2)


def make_mitbih_like_beats(N: int = 20000, L: int = 360, num_classes: int = 5,
                           seed: int = 1337) -> Tuple[np.ndarray, np.ndarray]:
    """Return ``(x [N, L] float32, y [N] int64)``."""
    if not 1 <= num_classes <= len(AAMI_CLASSES):
        raise ValueError("num_classes must be in 1..5")
    rng = np.random.default_rng(seed)
    pri = np.asarray(AAMI_PRIORS[:num_classes], dtype=np.float64)
    pri /= pri.sum()
    y = rng.choice(num_classes, size=N, p=pri).astype(np.int64)
    t = np.linspace(0.0, 1.0, L, dtype=np.float32)
    jit = lambda s: rng.normal(0.0, s, size=N).astype(np.float32)
    # per-class morphology: (P amp, QRS width, QRS amp, T amp, R position)
    table = np.array([
        [0.15, 0.012, 1.00, 0.30, 0.50],   # N  normal
        [0.02, 0.012, 0.90, 0.25, 0.44],   # S  supraventricular: early, P absent
        [0.00, 0.035, 1.30, -0.35, 0.48],  # V  ventricular: wide QRS, inverted T
        [0.08, 0.024, 1.10, 0.05, 0.49],   # F  fusion
        [0.00, 0.020, 0.40, 0.00, 0.50],   # Q  unknown / paced
    ], dtype=np.float32)[:num_classes]
    par = table[y]
    r_pos = par[:, 4] + jit(0.01)
    x = _gauss(t, r_pos - 0.18 + jit(0.01), np.full(N, 0.025, np.float32), par[:, 0] * (1 + jit(0.1)))
    x += _gauss(t, r_pos - 0.025, np.full(N, 0.010, np.float32), -0.12 * par[:, 2])
    x += _gauss(t, r_pos, par[:, 1] * (1 + np.abs(jit(0.1))), par[:, 2] * (1 + jit(0.1)))
    x += _gauss(t, r_pos + 0.03, np.full(N, 0.012, np.float32), -0.20 * par[:, 2])
    x += _gauss(t, r_pos + 0.25 + jit(0.015), np.full(N, 0.05, np.float32), par[:, 3] * (1 + jit(0.15)))
    phase = rng.uniform(0, 2 * np.pi, size=N).astype(np.float32)
    x += 0.05 * np.sin(2 * np.pi * 0.7 * t[None, :] + phase[:, None])
    x += rng.normal(0.0, 0.03, size=(N, L)).astype(np.float32)
    return np.ascontiguousarray(x, dtype=np.float32), y
