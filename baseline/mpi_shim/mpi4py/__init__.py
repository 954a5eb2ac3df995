"""Minimal stand-in for ``mpi4py`` so the UNMODIFIED reference scripts run where no MPI exists.

Only what the reference uses: ``MPI.COMM_WORLD.{Get_rank, Get_size, Bcast,
Allreduce, allreduce, gather, Barrier}`` and ``MPI.SUM``.  Collectives run over ``torch.distributed``
**gloo on host buffers** — the same data path as mpi4py's buffer-protocol collectives on numpy arrays
(the reference stages every tensor through the host anyway).  Launched by ``torchrun`` instead of
``mpiexec``; a single process needs no rendezvous at all.
"""
from . import MPI  # noqa: F401
