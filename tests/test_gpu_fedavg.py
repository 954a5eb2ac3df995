"""In-kernel FedAvg / all-reduce collectives, emulated ranks on one GPU (one cooperative launch),
vs the fp64 CPU oracle sum_i(n_i theta_i)/sum_i(n_i)  (SURVEY 6.3 (c))."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _oracle(thetas, prev, weights, server_lr):
    d = sum(w * (t.double().cpu() - prev.double().cpu()) for t, w in zip(thetas, weights))
    return (prev.double().cpu() + server_lr * d).float()


@pytest.mark.parametrize("world", [1, 2, 4, 8])
@pytest.mark.parametrize("n", [1458, 70001, 1 << 20])
@pytest.mark.parametrize("wire_bf16", [False, True])
def test_fedavg_local_group(world, n, wire_bf16):
    from crossscale.ops import _lib
    from crossscale.parallel.fedavg import aggregate_local_group
    from crossscale.parallel.symm import LocalGroup
    dev = torch.device("cuda", 0)
    torch.manual_seed(world * 7 + n % 13)
    need = int(_lib.cuda_lib().cs_fedavg_arena_bytes(n, int(wire_bf16), world))
    grp = LocalGroup(world, need)
    counts = torch.arange(1, world + 1, dtype=torch.float64) * 1000
    weights = (counts / counts.sum()).tolist()
    prev0 = torch.randn(n, device=dev)
    for rnd in range(3):                                        # repeated rounds: flags/epochs/parity reuse
        thetas = [prev0 + 0.01 * torch.randn(n, device=dev) for _ in range(world)]
        prevs = [prev0.clone() for _ in range(world)]
        moms = [torch.ones(n, device=dev) for _ in range(world)]
        want = _oracle(thetas, prev0, weights, 0.5)
        aggregate_local_group(grp, thetas, prevs, weights, n, wire_bf16=wire_bf16, server_lr=0.5, moms=moms)
        torch.cuda.synchronize()
        assert grp.status() == 0
        tol = dict(rtol=0, atol=1e-6) if not wire_bf16 else dict(rtol=0, atol=2e-4)
        for r in range(world):
            torch.testing.assert_close(thetas[r].cpu(), want, **tol)
            torch.testing.assert_close(prevs[r], thetas[r], rtol=0, atol=0)
            assert float(moms[r].abs().max()) == 0.0
            assert torch.equal(thetas[r], thetas[0])            # every client holds the same bits
        prev0 = thetas[0].clone()
    grp.close()


@pytest.mark.parametrize("world", [2, 8])
def test_allreduce_bucket_mean(world):
    import ctypes as C
    from crossscale.ops import _lib
    from crossscale.parallel.symm import LocalGroup, ptr_array
    dev, n = torch.device("cuda", 0), 300000
    grp = LocalGroup(world, int(_lib.cuda_lib().cs_fedavg_arena_bytes(n, 0, world)))
    grads = [torch.randn(n, device=dev) for _ in range(world)]
    want = torch.stack(grads).double().mean(0).float()
    _lib.check(_lib.cuda_lib().cs_allreduce_bucket_local(grp.handles, world, ptr_array(grads), n, 1.0 / world, 0, 0, 1,
                                                         _lib.stream_ptr()))
    torch.cuda.synchronize()
    for g in grads:
        torch.testing.assert_close(g, want, rtol=1e-5, atol=1e-6)
    grp.barrier(channel=2)
    torch.cuda.synchronize()
    assert grp.status() == 0
    grp.close()


@pytest.mark.parametrize("world", [2, 8])
@pytest.mark.parametrize("server_lr", [1.0, 0.5])
def test_fedavg_inplace_symmetric_parameters(world, server_lr):
    """Zero-staging variant: parameters live in the peer-mapped arena; rank r reduces slice r and pushes it
    into every replica."""
    import ctypes as C
    from crossscale.ops import _lib
    from crossscale.parallel.symm import LocalGroup, ptr_array
    dev, n = torch.device("cuda", 0), (1 << 20) + 64
    grp = LocalGroup(world, n * 4)
    thetas = [grp.arena_tensor(r, n) for r in range(world)]
    torch.manual_seed(world)
    prev0 = torch.randn(n, device=dev)
    counts = torch.arange(1, world + 1, dtype=torch.float64)
    weights = (counts / counts.sum()).tolist()
    wa = (C.c_float * world)(*weights)
    for rnd in range(2):
        for t in thetas:
            t.copy_(prev0 + 0.01 * torch.randn(n, device=dev))
        prevs = [prev0.clone() for _ in range(world)]
        moms = [torch.ones(n, device=dev) for _ in range(world)]
        avg = sum(w * t.double() for w, t in zip(weights, thetas))
        want = (prev0.double() + server_lr * (avg - prev0.double())).float()
        _lib.check(_lib.cuda_lib().cs_fedavg_aggregate_inplace_local(
            grp.handles, world, 0, ptr_array(prevs) if server_lr != 1.0 else None, n, wa, server_lr, ptr_array(moms), 3,
            _lib.stream_ptr()))
        torch.cuda.synchronize()
        assert grp.status() == 0
        for r in range(world):
            torch.testing.assert_close(thetas[r], want, rtol=0, atol=2e-6)
            assert torch.equal(thetas[r], thetas[0]) and float(moms[r].abs().max()) == 0.0
        prev0 = thetas[0].clone()
    del thetas
    grp.close()
