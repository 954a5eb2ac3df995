"""Binary shard container.

On-disk layout (byte-compatible with the reference writer/reader,
``Module_1/shard_prep.py:10-19`` and ``Module_3/shard_dataset.py:30-47``)::

    offset 0   int64 little-endian   N   (number of windows)
    offset 8   int64 little-endian   L   (samples per window)
    offset 16  N*L float32, row-major

Labels are NOT part of the reference format.  We keep the ``.bin`` payload untouched and put
labels in an optional sidecar ``<stem>.lbl`` (``[int64 N][int64 num_classes][N uint8]``), so a
reference reader can still consume every shard this framework writes.
"""
from __future__ import annotations

import os
import struct
from dataclasses import dataclass
from typing import Optional, Tuple

import numpy as np

HEADER_BYTES = 16
_HDR = struct.Struct("<qq")
SHARD_GLOB = "ecg_*.bin"


def shard_name(shard_id: int) -> str:
    """File name used by the reference (``shard_prep.py:63``)."""
    return f"ecg_{shard_id:05d}.bin"


def label_path_for(path: str) -> str:
    stem, _ = os.path.splitext(path)
    return stem + ".lbl"


@dataclass(frozen=True)
class ShardHeader:
    num_windows: int
    window_len: int

    @property
    def payload_bytes(self) -> int:
        return self.num_windows * self.window_len * 4


def write_shard(path: str, windows: np.ndarray, labels: Optional[np.ndarray] = None,
                num_classes: int = 0) -> ShardHeader:
    """Write ``windows`` ([N, L], any float dtype) as one shard; atomic via rename."""
    arr = np.ascontiguousarray(windows, dtype="<f4")
    if arr.ndim != 2:
        raise ValueError(f"windows must be [N, L], got shape {arr.shape}")
    n, l = arr.shape
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    tmp = path + ".tmp"
    with open(tmp, "wb") as f:
        f.write(_HDR.pack(n, l))
        f.write(memoryview(arr).cast("B"))
    os.replace(tmp, path)
    if labels is not None:
        lab = np.ascontiguousarray(labels, dtype=np.uint8)
        if lab.shape != (n,):
            raise ValueError(f"labels must be [N]={n}, got {lab.shape}")
        nc = int(num_classes) if num_classes else int(lab.max(initial=0)) + 1
        ltmp = label_path_for(path) + ".tmp"
        with open(ltmp, "wb") as f:
            f.write(_HDR.pack(n, nc))
            f.write(lab.tobytes())
        os.replace(ltmp, label_path_for(path))
    else:
        # an unlabelled shard must not inherit the sidecar of whatever was written under this name before
        try:
            os.unlink(label_path_for(path))
        except FileNotFoundError:
            pass
    return ShardHeader(n, l)


def read_header(path: str) -> ShardHeader:
    with open(path, "rb") as f:
        raw = f.read(HEADER_BYTES)
    if len(raw) != HEADER_BYTES:
        raise RuntimeError(f"shard {path}: truncated header ({len(raw)} bytes)")
    n, l = _HDR.unpack(raw)
    if n < 0 or l <= 0:
        raise RuntimeError(f"shard {path}: invalid header N={n} L={l}")
    return ShardHeader(n, l)


def load_shard(path: str, mmap: bool = False) -> np.ndarray:
    """Return the ``[N, L] float32`` payload.  ``mmap=True`` maps it read-only (zero copy)."""
    hdr = read_header(path)
    have = os.path.getsize(path) - HEADER_BYTES
    if have < hdr.payload_bytes:
        raise RuntimeError(
            f"shard {path}: payload has {have} bytes, header promises {hdr.payload_bytes}")
    if mmap:
        return np.memmap(path, dtype="<f4", mode="r", offset=HEADER_BYTES,
                         shape=(hdr.num_windows, hdr.window_len))
    with open(path, "rb") as f:
        f.seek(HEADER_BYTES)
        data = np.fromfile(f, dtype="<f4", count=hdr.num_windows * hdr.window_len)
    return data.reshape(hdr.num_windows, hdr.window_len)


def load_labels(path: str) -> Optional[Tuple[np.ndarray, int]]:
    """Labels from the sidecar of shard ``path`` or ``None`` when there is none."""
    lp = label_path_for(path)
    if not os.path.exists(lp):
        return None
    with open(lp, "rb") as f:
        n, nc = _HDR.unpack(f.read(HEADER_BYTES))
        lab = np.frombuffer(f.read(n), dtype=np.uint8)
    if lab.size != n:
        raise RuntimeError(f"label sidecar {lp}: truncated")
    shard_n = read_header(path).num_windows
    if n != shard_n:
        # a sidecar left behind by an earlier data set of a different size: labels that belong to other windows
        import warnings
        warnings.warn(f"label sidecar {lp} holds {n} labels but shard {path} has {shard_n} windows; ignoring the sidecar")
        return None
    return lab.astype(np.int64), int(nc)
