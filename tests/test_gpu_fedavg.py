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


def test_barrier_timeout_is_loud_and_leaves_the_model_untouched():
    """One emulated rank aggregates while its peer never arrives (ADVICE r1 / VERDICT r1 #10): the kernel must give up
    after the configured time-out, raise the communicator's status word, write NOTHING into the parameters or the
    previous-global copy, and the round loop's health check must turn that into an exception."""
    import types
    from crossscale.fl.engine import FedAvgRunner
    from crossscale.ops import _lib
    from crossscale.parallel.symm import LocalGroup
    lib = _lib.cuda_lib()
    dev = torch.device("cuda", 0)
    n = 1458
    group = LocalGroup(2, int(lib.cs_fedavg_arena_bytes(n, 0, 2)))
    group.set_timeout_ms(40)
    torch.manual_seed(0)
    flat, prev = torch.randn(n + 2, device=dev), torch.randn(n + 2, device=dev)
    flat0, prev0 = flat.clone(), prev.clone()
    _lib.check(lib.cs_debug_fedavg_aggregate_alone(group.handles[0], flat.data_ptr(), prev.data_ptr(), n, 0.5, None,
                                                   _lib.stream_ptr()), "cs_debug_fedavg_aggregate_alone")
    torch.cuda.synchronize()
    assert group.status() != 0
    assert torch.equal(flat, flat0) and torch.equal(prev, prev0)
    fake = types.SimpleNamespace(agg=types.SimpleNamespace(comm=group, backend="kernel"), fused=False)
    with pytest.raises(RuntimeError, match="did not reach the cross-GPU barrier"):
        FedAvgRunner.check_health(fake)
    group.close()


@pytest.mark.parametrize("world,server_lr", [(2, 1.0), (4, 1.0), (8, 0.5)])
def test_bucket_fused_sgd_aggregate_shadow_emulated(world, server_lr):
    """cs_fedavg_bucket: optimizer step of the bucket + in-place two-shot weighted mean + bf16 shadow in ONE launch
    (emulated ranks, one cooperative launch).  Oracle in fp64: theta_i' = theta_i - lr*(mu*m_i + g_i), then
    prev + server_lr * (sum_i w_i theta_i' - prev)."""
    import ctypes as C
    from crossscale.ops import _lib
    from crossscale.parallel.symm import LocalGroup, ptr_array
    lib = _lib.cuda_lib()
    dev = torch.device("cuda", 0)
    n_total, lo, n = 40000, 8192, 24000                       # a bucket in the middle of the parameter vector
    group = LocalGroup(world, n_total * 4)
    torch.manual_seed(1)
    params = [group.arena_tensor(r, n_total) for r in range(world)]
    grads, moms, prevs, shadows = [], [], [], []
    prev0 = torch.randn(n_total, device=dev)
    for r in range(world):
        params[r].copy_(prev0 + 0.05 * torch.randn(n_total, device=dev))
        grads.append(torch.randn(n_total, device=dev)); moms.append(torch.randn(n_total, device=dev) * 0.1)
        prevs.append(prev0.clone()); shadows.append(torch.zeros(n_total, dtype=torch.bfloat16, device=dev))
    counts = [100 + 37 * r for r in range(world)]
    wts = [c / sum(counts) for c in counts]
    lr, mu = 0.01, 0.9
    p64 = [p.double().clone() for p in params]
    m64 = [mu * moms[r].double() + grads[r].double() for r in range(world)]
    stepped = [p64[r] - lr * m64[r] for r in range(world)]
    mean = sum(wts[r] * stepped[r] for r in range(world))
    want = prev0.double() + server_lr * (mean - prev0.double())
    w = (C.c_float * world)(*wts)
    sl = slice(lo, lo + n)
    use_prev = server_lr != 1.0
    _lib.check(lib.cs_fedavg_bucket_local(
        group.handles, world, lo * 4, ptr_array([g[sl] for g in grads]), ptr_array([m[sl] for m in moms]),
        ptr_array([p[sl] for p in prevs]) if use_prev else None, ptr_array([s[sl] for s in shadows]), n, w, lr, mu, 0.0, 1.0,
        server_lr, 1, 3, _lib.stream_ptr()), "cs_fedavg_bucket_local")
    torch.cuda.synchronize()
    assert group.status() == 0
    for r in range(world):
        assert float((params[r][sl].double() - want[sl]).abs().max()) < 2e-6
        assert torch.equal(params[r][sl], params[0][sl])                      # bit-identical replicas
        assert torch.equal(shadows[r][sl], params[r][sl].to(torch.bfloat16))  # shadow = bf16 of the aggregated bucket
        assert float((moms[r][sl].double() - m64[r][sl]).abs().max()) < 1e-6  # momentum stepped, stays local
        # everything outside the bucket is untouched
        assert torch.equal(params[r][:lo].double(), p64[r][:lo]) and torch.equal(params[r][lo + n:].double(), p64[r][lo + n:])
        assert float(shadows[r][:lo].abs().max()) == 0.0
        if use_prev:                                   # a rank keeps the previous global model of the slice it owns
            g4 = n // 4
            gps = (g4 + world - 1) // world
            own = slice(lo + 4 * r * gps, lo + min(4 * (r + 1) * gps, n))
            assert torch.equal(prevs[r][own], params[r][own])
    group.close()
