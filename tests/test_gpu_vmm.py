"""Own symmetric-memory allocator (csrc/cuda/symm_vmm.cu, parallel/vmm.py) on one GPU: allocation, export / import of the
descriptor inside one process, a communicator over the mappings, multicast object with a single device.  The multi-process
paths (descriptors over AF_UNIX sockets, P2P aggregation and NVLS over the mappings) are in tests/test_gpu_multi.py."""
import ctypes as C
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


def _need():
    from crossscale.parallel import vmm
    s = vmm.support()
    if not s & 1:
        pytest.skip("no fd-exportable VMM allocations on this device")
    return s


def test_vmm_alloc_export_import_same_process():
    _need()
    from crossscale.ops import _lib
    from crossscale.parallel.symm import _arena_view
    lib = _lib.cuda_lib()
    ptr, size, handle, fd = C.c_void_p(), C.c_size_t(), C.c_ulonglong(), C.c_int(-1)
    _lib.check(lib.cs_vmm_alloc(3 << 20, 0, 1, C.byref(ptr), C.byref(size), C.byref(handle), C.byref(fd)), "cs_vmm_alloc")
    assert size.value >= 3 << 20 and fd.value >= 0 and ptr.value
    a = _arena_view(ptr.value, size.value, 1000, 0, torch.float32)
    assert float(a.abs().sum()) == 0.0                      # zero-filled
    a.copy_(torch.arange(1000, dtype=torch.float32, device=a.device))
    p2, h2 = C.c_void_p(), C.c_ulonglong()
    _lib.check(lib.cs_vmm_import(os.dup(fd.value), size.value, C.byref(p2), C.byref(h2)), "cs_vmm_import")
    assert p2.value != ptr.value                            # a second mapping of the same physical memory
    b = _arena_view(p2.value, size.value, 1000, 0, torch.float32)
    torch.cuda.synchronize()
    assert torch.equal(a, b)
    b[5] = -7.0
    torch.cuda.synchronize()
    assert float(a[5]) == -7.0
    del a, b
    lib.cs_vmm_close_fd(fd.value)
    lib.cs_vmm_free(p2, size.value, h2.value)
    lib.cs_vmm_free(ptr, size.value, handle.value)


def test_comm_over_vmm_arena_single_rank():
    _need()
    from crossscale.parallel.fedavg import FedAvgAggregator
    from crossscale.parallel.symm import SymmetricComm
    comm = SymmetricComm(1 << 20, backend="vmm")
    assert comm.arena_bytes >= 1 << 20
    n = 70000
    agg = FedAvgAggregator(n, torch.device("cuda", 0), comm=comm, inplace="on")
    flat = agg.param_view
    flat.copy_(torch.randn(n, device=flat.device))
    want = flat.clone()
    prev = torch.zeros(n, device=flat.device)
    agg.aggregate(flat, prev, 1.0)                          # world 1: the weighted mean of one model is the model
    torch.cuda.synchronize()
    assert comm.status() == 0 and torch.allclose(flat, want)
    comm.barrier()
    torch.cuda.synchronize()
    comm.close()


def test_multicast_needs_two_devices():
    """a multicast object cannot be made for one device (cuMulticastCreate rejects it): the buffer then simply has no
    multicast mapping; the two- and eight-process NVLS tests cover the real thing"""
    _need()
    from crossscale.parallel.vmm import VmmBuffer
    buf = VmmBuffer(1 << 21, multicast=True)
    assert buf.multicast_ptr == 0
    t = buf.tensor(1024)
    t.fill_(2.0)
    torch.cuda.synchronize()
    assert float(t.sum()) == 2048.0
    del t
    buf.close()
