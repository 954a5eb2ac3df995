"""K11: permutation gather of a mini-batch from the GPU-resident dataset.

Replaces ``x_gpu[sel]`` / ``y_gpu[sel]`` (``Module_3/shard_dataset.py:135-136``) with one kernel that
moves both tensors and can emit bf16.  The fused training-step kernel takes the index vector directly
(no materialised batch); this op serves the iterator API and the unfused paths.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import _lib


def gather_rows(x: torch.Tensor, y: Optional[torch.Tensor], idx: torch.Tensor,
                out_bf16: bool = False) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    if not x.is_cuda:
        xb = x[idx]
        return (xb.to(torch.bfloat16) if out_bf16 else xb), (None if y is None else y[idx])
    assert x.dtype == torch.float32 and x.is_contiguous() and idx.dtype == torch.int64
    b, l = idx.numel(), x.shape[1]
    xo = torch.empty((b, l), dtype=torch.bfloat16 if out_bf16 else torch.float32, device=x.device)
    yo = None if y is None else torch.empty((b,), dtype=torch.int64, device=x.device)
    _lib.check(_lib.cuda_lib().cs_gather_rows(x.data_ptr(), _lib.ptr(y), idx.data_ptr(), xo.data_ptr(),
                                              _lib.ptr(yo), b, l, int(out_bf16), _lib.stream_ptr()),
               "cs_gather_rows")
    return xo, yo
