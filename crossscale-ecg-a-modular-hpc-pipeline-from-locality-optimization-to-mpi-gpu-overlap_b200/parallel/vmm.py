"""Own symmetric-memory allocator: CUDA VMM allocations shared between the ranks of a node, optionally multicast (NVLS).

The native half is ``csrc/cuda/symm_vmm.cu`` (create / export / import / map / multicast bind).  This module does what needs
a host-side transport: the exported POSIX file descriptors travel over AF_UNIX sockets (``SCM_RIGHTS``), one listening
socket per rank in the abstract namespace, and ``torch.distributed`` object collectives carry only the socket token, the
allocation size and the barriers between the collective steps.  After set-up nothing here is on a hot path: kernels load
from / store to ``peer_ptrs`` directly and issue ``multimem`` instructions on ``multicast_ptr``.

Replaces ``cudaIpc*`` handles and ``torch.distributed._symmetric_memory`` of round 1; the reference's
counterpart is the MPI library under ``comm.Bcast`` / ``comm.Allreduce``
(``Module_3/TRUE_FL_M3/part3_fedavg_overlap_mpi_gpu.py:74-98``).
"""
from __future__ import annotations

import ctypes as C
import os
import socket
import threading
from typing import List, Optional

import torch
import torch.distributed as dist

from ..ops import _lib


def support() -> int:
    """bit 0: fd-exportable VMM allocations on the current device, bit 1: multicast objects"""
    if not _lib.have_gpu():
        return 0
    return int(_lib.cuda_lib().cs_vmm_support())


def _exchange_fds(my_fds: Optional[List[int]], servers: List[int], rank: int, world: int, token: str, group=None) -> List[List[int]]:
    """ranks in ``servers`` offer ``my_fds`` to every other rank; returns, per serving rank, the descriptors as received
    in THIS process (empty for non-servers and for the own rank)"""
    name = lambda r: "\0" + f"cs_vmm_{token}_{r}"      # abstract namespace: no file to clean up
    err: List[BaseException] = []
    srv = th = None
    if rank in servers:
        srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        srv.bind(name(rank))
        srv.listen(world)

        def serve():
            try:
                for _ in range(world - 1):
                    conn, _ = srv.accept()
                    with conn:
                        conn.recv(4)
                        socket.send_fds(conn, [b"f"], my_fds)
            except BaseException as e:  # pragma: no cover
                err.append(e)

        th = threading.Thread(target=serve, daemon=True)
        th.start()
    dist.barrier(group=group)                                # every server listens
    out: List[List[int]] = [[] for _ in range(world)]
    for r in servers:
        if r == rank:
            continue
        with socket.socket(socket.AF_UNIX, socket.SOCK_STREAM) as c:
            c.connect(name(r))
            c.sendall(rank.to_bytes(4, "little"))
            _, fds, _, _ = socket.recv_fds(c, 16, 8)
            if not fds:
                raise RuntimeError(f"vmm: no descriptor received from rank {r}")
            out[r] = list(fds)
    if th is not None:
        th.join(timeout=60)
        srv.close()
    if err:
        raise err[0]
    dist.barrier(group=group)
    return out


class VmmBuffer:
    """One symmetric buffer of ``nbytes`` per rank: ``peer_ptrs[r]`` is this process's mapping of rank r's copy
    (``peer_ptrs[rank]`` the local one); ``multicast_ptr`` (0 if unavailable / not requested) addresses all copies at once."""

    def __init__(self, nbytes: int, multicast: bool = False, group: Optional[dist.ProcessGroup] = None):
        lib = _lib.cuda_lib()
        self.lib = lib
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        sup = support()
        if not sup & 1:
            raise RuntimeError("this device / driver offers no fd-exportable VMM allocations")
        self.multicast = bool(multicast) and bool(sup & 2) and self.world >= 2      # (a multicast object needs >= 2 devices)
        ptr, size, handle, fd = C.c_void_p(), C.c_size_t(), C.c_ulonglong(), C.c_int(-1)
        _lib.check(lib.cs_vmm_alloc(int(nbytes), int(self.multicast), self.world, C.byref(ptr), C.byref(size), C.byref(handle), C.byref(fd)),
                   "cs_vmm_alloc")
        self.alloc_bytes, self.nbytes = int(size.value), int(nbytes)
        self._maps = {self.rank: (int(ptr.value), int(handle.value))}         # rank -> (ptr, handle)
        self.multicast_ptr = 0
        self._mc = None                                                       # (mc_handle, mc_ptr)
        if self.world > 1:
            tok = [os.urandom(8).hex() if self.rank == 0 else None]
            dist.broadcast_object_list(tok, src=0, group=group)
            sizes = [None] * self.world
            dist.all_gather_object(sizes, self.alloc_bytes, group=group)
            if len(set(sizes)) != 1:
                raise RuntimeError(f"vmm: ranks allocated different sizes {sizes}")
            got = _exchange_fds([fd.value], list(range(self.world)), self.rank, self.world, tok[0] + "a", group)
            for r in range(self.world):
                if r == self.rank:
                    continue
                p, h = C.c_void_p(), C.c_ulonglong()
                _lib.check(lib.cs_vmm_import(got[r][0], self.alloc_bytes, C.byref(p), C.byref(h)), "cs_vmm_import")
                os.close(got[r][0])
                self._maps[r] = (int(p.value), int(h.value))
        lib.cs_vmm_close_fd(fd.value)
        if self.multicast:
            self._setup_multicast(group)
        self.peer_ptrs = [self._maps[r][0] for r in range(self.world)]

    def _setup_multicast(self, group) -> None:
        lib = self.lib
        mc, mfd = C.c_ulonglong(), C.c_int(-1)
        if self.rank == 0:
            _lib.check(lib.cs_vmm_mc_create(self.world, self.alloc_bytes, C.byref(mc), C.byref(mfd)), "cs_vmm_mc_create")
        if self.world > 1:
            tok = [os.urandom(8).hex() if self.rank == 0 else None]
            dist.broadcast_object_list(tok, src=0, group=group)
            got = _exchange_fds([mfd.value] if self.rank == 0 else None, [0], self.rank, self.world, tok[0] + "m", group)
            if self.rank != 0:
                _lib.check(lib.cs_vmm_mc_import(got[0][0], C.byref(mc)), "cs_vmm_mc_import")
                os.close(got[0][0])
        if self.rank == 0:
            lib.cs_vmm_close_fd(mfd.value)
        _lib.check(lib.cs_vmm_mc_add_device(mc.value), "cs_vmm_mc_add_device")
        if self.world > 1:
            dist.barrier(group=group)                        # every device is in the group before anyone binds
        mptr = C.c_void_p()
        _lib.check(lib.cs_vmm_mc_bind_map(mc.value, self._maps[self.rank][1], self.alloc_bytes, C.byref(mptr)), "cs_vmm_mc_bind_map")
        if self.world > 1:
            dist.barrier(group=group)
        self._mc = (int(mc.value), int(mptr.value))
        self.multicast_ptr = int(mptr.value)

    @property
    def local_ptr(self) -> int:
        return self._maps[self.rank][0]

    def tensor(self, numel: int, offset_bytes: int = 0, dtype=torch.float32) -> torch.Tensor:
        from .symm import _arena_view
        return _arena_view(self.local_ptr, self.alloc_bytes, numel, offset_bytes, dtype)

    def close(self) -> None:
        if getattr(self, "_maps", None) is None:
            return
        torch.cuda.synchronize()
        if dist.is_initialized() and self.world > 1:
            try:
                dist.barrier()                               # nobody unmaps while a peer's kernel may still touch the memory
            except Exception:  # pragma: no cover
                pass
        if self._mc is not None:
            self.lib.cs_vmm_free(self._mc[1], self.alloc_bytes, self._mc[0])
            self._mc = None
        for r, (p, h) in self._maps.items():
            self.lib.cs_vmm_free(p, self.alloc_bytes, h)
        self._maps = None
