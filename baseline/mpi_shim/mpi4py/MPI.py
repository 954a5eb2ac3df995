from __future__ import annotations

import datetime
import os

import numpy as np
import torch
import torch.distributed as dist

SUM = "sum"
MAX = "max"
MIN = "min"

_OPS = {SUM: dist.ReduceOp.SUM, MAX: dist.ReduceOp.MAX, MIN: dist.ReduceOp.MIN}


class _Comm:
    def __init__(self):
        self._world = int(os.environ.get("WORLD_SIZE", "1"))
        self._rank = int(os.environ.get("RANK", "0"))
        self._pg = None

    def _group(self):
        if self._world == 1:
            return None
        if self._pg is None:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29533")
            if not dist.is_initialized():
                dist.init_process_group("gloo", rank=self._rank, world_size=self._world,
                                        timeout=datetime.timedelta(seconds=600))
                self._pg = dist.group.WORLD
            else:
                self._pg = dist.new_group(backend="gloo")
        return self._pg

    def Get_rank(self):
        return self._rank

    def Get_size(self):
        return self._world

    def Barrier(self):
        g = self._group()
        if g is not None:
            dist.barrier(group=g)

    barrier = Barrier

    def Bcast(self, buf, root=0):
        g = self._group()
        if g is None:
            return
        arr = np.asarray(buf)
        t = torch.from_numpy(arr) if arr.flags.writeable else torch.from_numpy(arr.copy())
        dist.broadcast(t, src=root, group=g)
        if not arr.flags.writeable:  # pragma: no cover
            np.copyto(buf, t.numpy())

    def Allreduce(self, sendbuf, recvbuf, op=SUM):
        src = np.asarray(sendbuf)
        t = torch.from_numpy(np.ascontiguousarray(src).copy())
        g = self._group()
        if g is not None:
            dist.all_reduce(t, op=_OPS[op], group=g)
        np.copyto(np.asarray(recvbuf), t.numpy().reshape(src.shape))

    def allreduce(self, value, op=SUM):
        g = self._group()
        if g is None:
            return value
        t = torch.tensor([float(value)], dtype=torch.float64)
        dist.all_reduce(t, op=_OPS[op], group=g)
        return float(t.item())

    def gather(self, obj, root=0):
        g = self._group()
        if g is None:
            return [obj]
        out = [None] * self._world if self._rank == root else None
        dist.gather_object(obj, out, dst=root, group=g)
        return out


COMM_WORLD = _Comm()
