"""FedAvg aggregation (K7) — replaces ``broadcast_model`` + ``fedavg_allreduce``
(``TRUE_FL_M3/part3_fedavg_overlap_mpi_gpu.py:75-98``).

Three back-ends with identical semantics ``theta <- prev + server_lr * sum_i w_i (theta_i - prev)``,
``w_i = n_i / sum_j n_j`` (equal ``n_i`` gives the reference's unweighted mean):

* ``kernel``   the fused in-kernel NVLink collective (``csrc/cuda/fedavg.cu``) — the product path
* ``nccl``     one flattened ``dist.all_reduce`` — the "only calls NCCL" baseline
* ``gloo``     host tensors through gloo — CPU plumbing and the semantics oracle for tests

Because every rank leaves an aggregation with bit-identical parameters the reference's start-of-round
``Bcast`` is redundant; an explicit :func:`broadcast_model` is kept for round 0 /
checkpoint loads.
"""
from __future__ import annotations

import os
from typing import Optional, Sequence

import torch
import torch.distributed as dist

from ..ops import _lib
from .symm import LocalGroup, SymmetricComm, ptr_array


def _world() -> int:
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def client_weight(n_local: int, device=None) -> float:
    """n_i / sum_j n_j via a one-time scalar all-reduce (setup, not hot path)."""
    if _world() == 1:
        return 1.0
    dev = device if (device is not None and dist.get_backend() == "nccl") else "cpu"
    t = torch.tensor([float(n_local)], dtype=torch.float64, device=dev)
    dist.all_reduce(t)
    return float(n_local) / float(t.item())


def broadcast_model(flat: torch.Tensor, root: int = 0) -> None:
    """Push rank ``root``'s parameters to everyone (round 0 only)."""
    if _world() > 1:
        if flat.is_cuda and dist.get_backend() != "nccl":
            h = flat.detach().cpu()
            dist.broadcast(h, src=root)
            flat.copy_(h)
        else:
            dist.broadcast(flat, src=root)


class FedAvgAggregator:
    def __init__(self, numel: int, device, backend: str = "auto", wire: str = "fp32",
                 server_lr: float = 1.0, reset_momentum: bool = False, channel: int = 0,
                 comm: Optional[SymmetricComm] = None, inplace: str = "auto"):
        self.numel, self.device = int(numel), torch.device(device)
        self.wire_bf16 = int(wire == "bf16")
        self.server_lr, self.reset_momentum, self.channel = float(server_lr), bool(reset_momentum), channel
        if backend == "auto":
            backend = "kernel" if self.device.type == "cuda" else "gloo"
        self.backend = backend
        self.comm = comm
        self.launches = 0
        # large fp32 models: keep the parameters IN the arena and aggregate in place (no staging passes)
        self.inplace = (backend == "kernel" and not self.wire_bf16 and self.numel % 4 == 0
                        and ((self.numel > 65536 and inplace == "auto") or inplace in ("on", "nvls")))
        self.nvls = None
        self.param_view: Optional[torch.Tensor] = None
        self._weights = None
        self._mc_ptr = 0
        if self.inplace and inplace == "nvls":
            if _world() < 2:
                raise ValueError("NVLS aggregation needs at least 2 ranks")
            from . import vmm
            if self.comm is None and (vmm.support() & 2) and os.environ.get("CROSSSCALE_NVLS_ALLOC", "vmm") == "vmm":
                # the repo's own allocator: the arena is a VMM allocation bound to a multicast object (parallel/vmm.py)
                self.comm = SymmetricComm(self.numel * 4, backend="vmm", multicast=True)
                self._mc_ptr = int(self.comm.multicast_ptr)
                if not self._mc_ptr:
                    raise RuntimeError("this system offers no NVLS multicast mapping")
                self.param_view = self.comm.arena_tensor(self.numel)
                self.nvls = ("vmm", None)
            else:
                # fallback: a torch symmetric-memory buffer that is also mapped at a multicast address; torch is used for
                # that allocation / rendezvous only, the collective is our kernel
                import torch.distributed._symmetric_memory as symm_mem
                t = symm_mem.empty(self.numel, dtype=torch.float32, device=self.device)
                h = symm_mem.rendezvous(t, dist.group.WORLD)
                if not int(h.multicast_ptr):
                    raise RuntimeError("this system offers no NVLS multicast mapping")
                self.nvls = (t, h)
                self._mc_ptr = int(h.multicast_ptr)
                self.param_view = t
        if backend == "kernel":
            need = (4096 if (self.nvls is not None and self.nvls[0] != "vmm") else self.numel * 4 if self.inplace
                    else int(_lib.cuda_lib().cs_fedavg_arena_bytes(self.numel, self.wire_bf16, _world())))
            if self.comm is None:
                self.comm = SymmetricComm(need)
            elif self.comm.arena_bytes < need:
                raise ValueError(f"arena too small: {self.comm.arena_bytes} < {need}")
            if self.inplace and self.nvls is None:
                self.param_view = self.comm.arena_tensor(self.numel)

    def set_weights(self, weight: float) -> None:
        """in-place mode needs every client's weight on every rank (one host all-gather, set-up time)"""
        import ctypes as C
        ws = [float(weight)]
        if _world() > 1:
            ws = [None] * _world()
            dist.all_gather_object(ws, float(weight))
        self._weights = (C.c_float * len(ws))(*ws)
        # NVLS: the switch computes a plain sum; unequal weights are applied by pre-scaling the local replica
        self._prescale = 1.0 if max(ws) - min(ws) < 1e-12 * max(ws) else float(weight) * len(ws)

    def aggregate(self, flat: torch.Tensor, prev: torch.Tensor, weight: float,
                  mom: Optional[torch.Tensor] = None) -> None:
        """In place on ``flat`` and ``prev`` (both end up holding the new global model)."""
        n = self.numel
        if self.backend == "kernel" and self.inplace:
            if flat.data_ptr() != self.param_view.data_ptr():
                raise ValueError("in-place aggregation: the parameters must live in aggregator.param_view (FlatParams.rebind)")
            if self._weights is None:
                self.set_weights(weight)
            use_prev = self.server_lr != 1.0
            if self.nvls is not None:
                _lib.check(_lib.cuda_lib().cs_fedavg_aggregate_nvls(
                    self.comm.handle, flat.data_ptr(), self._mc_ptr, prev.data_ptr() if use_prev else None,
                    n, self._prescale, 1.0 / _world(), self.server_lr, _lib.ptr(mom) if self.reset_momentum else None,
                    self.channel, _lib.stream_ptr()), "cs_fedavg_aggregate_nvls")
                self.launches += 1
                return
            _lib.check(_lib.cuda_lib().cs_fedavg_aggregate_inplace(
                self.comm.handle, 0, prev.data_ptr() if use_prev else None, n, self._weights, self.server_lr,
                _lib.ptr(mom) if self.reset_momentum else None, self.channel, _lib.stream_ptr()), "cs_fedavg_aggregate_inplace")
            self.launches += 1
            return
        if self.backend == "kernel":
            _lib.check(_lib.cuda_lib().cs_fedavg_aggregate(
                self.comm.handle, flat.data_ptr(), prev.data_ptr(), n, float(weight), self.wire_bf16,
                self.server_lr, _lib.ptr(mom) if self.reset_momentum else None, 0, self.channel,
                _lib.stream_ptr()), "cs_fedavg_aggregate")
            self.launches += 1
            return
        delta = (flat[:n] - prev[:n]) * weight
        if _world() > 1:
            if self.backend == "gloo" and delta.is_cuda:
                h = delta.cpu()
                dist.all_reduce(h)
                delta = h.to(flat.device)
            else:
                dist.all_reduce(delta)
        new = prev[:n] + self.server_lr * delta
        flat[:n].copy_(new)
        prev[:n].copy_(new)
        if self.reset_momentum and mom is not None:
            mom.zero_()

    def close(self):
        if self.comm is not None and self.backend == "kernel":
            self.comm.close()


def aggregate_local_group(group: LocalGroup, flats: Sequence[torch.Tensor], prevs: Sequence[torch.Tensor],
                          weights: Sequence[float], numel: int, wire_bf16: bool = False,
                          server_lr: float = 1.0, moms: Optional[Sequence[torch.Tensor]] = None,
                          channel: int = 0) -> None:
    """All emulated clients of a :class:`LocalGroup` in one cooperative launch."""
    import ctypes as C
    w = (C.c_float * group.world)(*[float(x) for x in weights])
    _lib.check(_lib.cuda_lib().cs_fedavg_aggregate_local(
        group.handles, group.world, ptr_array(flats), ptr_array(prevs) if prevs is not None else None,
        int(numel), w, int(wire_bf16), float(server_lr), ptr_array(moms) if moms is not None else None,
        0, channel, _lib.stream_ptr()), "cs_fedavg_aggregate_local")
