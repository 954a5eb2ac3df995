"""Symmetric peer memory: one arena + signal pad per rank, mapped into every peer (CUDA IPC, or the repo's own
VMM allocator with ``backend="vmm"``, which can also map the arena at an NVSwitch multicast address).

This is the bootstrap layer under the in-kernel collectives: the only host-side
communication is the one-time exchange of two 64-byte IPC handles per rank through the
``torch.distributed`` store/gloo/nccl object collectives — the hot paths never call MPI or NCCL.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import torch
import torch.distributed as dist

from ..ops import _lib

HANDLE_BYTES = 64


class SymmetricComm:
    """A communicator over peer-mapped arenas (real multi-process, one GPU per rank)."""

    def __init__(self, arena_bytes: int, group: Optional[dist.ProcessGroup] = None, backend: Optional[str] = None,
                 multicast: bool = False):
        """``backend``: "ipc" (cudaMalloc + CUDA IPC handles, the default) or "vmm" (the repo's own allocator,
        ``parallel/vmm.py``: fd-exported VMM allocations; ``multicast=True`` additionally maps the arena at an NVSwitch
        multicast address, ``self.multicast_ptr``).  Environment default: ``CROSSSCALE_SYMM``."""
        import os
        lib = _lib.cuda_lib()
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self._h = C.c_void_p()
        self.backend = backend or os.environ.get("CROSSSCALE_SYMM", "ipc")
        self.multicast_ptr = 0
        self._vmm = None
        if self.backend == "vmm":
            from .vmm import VmmBuffer
            arena = (int(arena_bytes) + 4095) // 4096 * 4096
            sig = int(lib.cs_comm_signal_bytes())
            self._vmm = VmmBuffer(arena + sig, multicast=multicast, group=group)
            ap = (C.c_void_p * self.world)(*[p for p in self._vmm.peer_ptrs])
            sp = (C.c_void_p * self.world)(*[p + arena for p in self._vmm.peer_ptrs])
            _lib.check(lib.cs_comm_create_external(self.rank, self.world, arena, ap, sp, C.byref(self._h)), "cs_comm_create_external")
            self.multicast_ptr = self._vmm.multicast_ptr
            if self.world > 1:
                dist.barrier(group=group)
        else:
            ah = (C.c_char * HANDLE_BYTES)()
            sh = (C.c_char * HANDLE_BYTES)()
            _lib.check(lib.cs_comm_create(self.rank, self.world, int(arena_bytes), C.byref(self._h), ah, sh),
                       "cs_comm_create")
            if self.world > 1:
                gathered: List = [None] * self.world
                dist.all_gather_object(gathered, (bytes(ah.raw), bytes(sh.raw)), group=group)
                ahs = b"".join(g[0] for g in gathered)
                shs = b"".join(g[1] for g in gathered)
                _lib.check(lib.cs_comm_connect(self._h, ahs, shs), "cs_comm_connect")
                dist.barrier(group=group)  # every rank has mapped every peer before first use
        self.arena_bytes = int(lib.cs_comm_arena_bytes(self._h))
        self.emulated = False

    @property
    def handle(self):
        return self._h

    def arena_tensor(self, numel: int, offset_bytes: int = 0, dtype=torch.float32) -> torch.Tensor:
        """A torch view of this rank's arena (peer-mapped memory): tensors placed here are readable and
        writable by every peer's kernels at the same offset."""
        return _arena_view(int(_lib.cuda_lib().cs_comm_arena(self._h)), self.arena_bytes, numel, offset_bytes, dtype)

    def barrier(self, channel: int = 0) -> None:
        _lib.check(_lib.cuda_lib().cs_comm_barrier(self._h, channel, _lib.stream_ptr()), "cs_comm_barrier")

    def status(self) -> int:
        out = C.c_int(0)
        _lib.check(_lib.cuda_lib().cs_comm_status(self._h, C.byref(out)), "cs_comm_status")
        return out.value

    def set_timeout_ms(self, ms: int) -> None:
        """how long an in-kernel cross-GPU flag wait may spin before it gives up (default 4000 ms)"""
        _lib.check(_lib.cuda_lib().cs_comm_set_timeout_ms(self._h, int(ms)), "cs_comm_set_timeout_ms")

    def close(self) -> None:
        if self._h:
            torch.cuda.synchronize()
            _lib.cuda_lib().cs_comm_destroy(self._h)
            self._h = C.c_void_p()
            if self._vmm is not None:
                self._vmm.close()
                self._vmm = None

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass


class LocalGroup:
    """``world`` emulated ranks on ONE device; every collective is one cooperative launch.

    Used by single-GPU tests and the 1-GPU multi-client simulation: the same kernels run with
    ``blockIdx.y`` as the rank, so the peer-pointer / flag protocol is exercised without peers.
    """

    def __init__(self, world: int, arena_bytes: int):
        lib = _lib.cuda_lib()
        self.world = world
        self._arr = (C.c_void_p * world)()
        _lib.check(lib.cs_comm_create_local(world, int(arena_bytes), self._arr), "cs_comm_create_local")
        self.arena_bytes = int(lib.cs_comm_arena_bytes(self._arr[0]))
        self.emulated = True

    @property
    def handles(self):
        return self._arr

    def arena_tensor(self, rank: int, numel: int, offset_bytes: int = 0, dtype=torch.float32) -> torch.Tensor:
        return _arena_view(int(_lib.cuda_lib().cs_comm_arena(self._arr[rank])), self.arena_bytes, numel, offset_bytes, dtype)

    def barrier(self, channel: int = 0) -> None:
        _lib.check(_lib.cuda_lib().cs_comm_barrier(self._arr[0], channel, _lib.stream_ptr()), "cs_comm_barrier")

    def status(self) -> int:
        out = C.c_int(0)
        worst = 0
        for r in range(self.world):
            _lib.check(_lib.cuda_lib().cs_comm_status(self._arr[r], C.byref(out)), "cs_comm_status")
            worst = max(worst, out.value)
        return worst

    def set_timeout_ms(self, ms: int) -> None:
        _lib.check(_lib.cuda_lib().cs_comm_set_timeout_ms(self._arr[0], int(ms)), "cs_comm_set_timeout_ms")

    def close(self) -> None:
        if self._arr is not None and self._arr[0]:
            torch.cuda.synchronize()
            _lib.cuda_lib().cs_comm_destroy(self._arr[0])
            self._arr = None


class _RawCuda:
    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3}


def _arena_view(base: int, arena_bytes: int, numel: int, offset_bytes: int, dtype) -> torch.Tensor:
    nbytes = numel * torch.empty((), dtype=dtype).element_size()
    if offset_bytes % 16 or offset_bytes + nbytes > arena_bytes:
        raise ValueError("arena view out of range / misaligned")
    raw = torch.as_tensor(_RawCuda(base + offset_bytes, nbytes), device=torch.device("cuda", torch.cuda.current_device()))
    return raw.view(dtype)


def ptr_array(tensors) -> "C.Array":
    arr = (C.c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = None if t is None else t.data_ptr()
    return arr
