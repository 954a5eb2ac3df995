"""tcgen05 implicit-GEMM conv1d (fwd / dgrad / wgrad), Cin=1 FFMA kernels, BN/ReLU/pool kernels and the
ResNet-1D built from them, against plain PyTorch fp32 references."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _rel(a, b):
    return float((a.double() - b.double()).norm() / (b.double().norm() + 1e-30))


def _bf(t):
    return t.to(torch.bfloat16).float()


CASES = [(4, 500, 16, 16, 5, 2), (2, 360, 64, 128, 7, 3), (3, 130, 256, 256, 3, 1), (2, 500, 32, 48, 31, 15),
         (5, 127, 128, 64, 1, 0), (1, 300, 768, 256, 7, 3), (2, 64, 16, 16, 3, 0),
         # 128-byte-row (swizzled) operand path: wide taps, Cout tile of 32, ragged last tile
         (2, 500, 64, 64, 31, 15), (2, 200, 128, 96, 9, 4), (1, 513, 64, 256, 5, 2),
         # short windows: several samples packed into one 128-row tile (slot 64 / 32 / 16)
         (6, 45, 64, 128, 7, 3), (3, 20, 32, 32, 5, 2), (19, 10, 16, 16, 3, 1), (5, 58, 256, 256, 7, 3), (64, 45, 768, 768, 7, 3),
         (9, 24, 64, 64, 5, 2), (7, 30, 128, 256, 3, 1),
         # narrow layers, persistent second-generation kernels (conv1d_narrow.cu): more work items than SMs (ring wrap-around),
         # two item groups per sample with a ragged last one, 48/32-channel operands, tap groups of 16
         (300, 500, 16, 16, 3, 1), (40, 1100, 32, 16, 7, 3), (7, 700, 48, 32, 16, 8), (3, 333, 16, 48, 31, 15), (600, 130, 16, 32, 5, 2)]


@pytest.mark.parametrize("B,L,Cin,Cout,K,pad", CASES)
def test_conv1d_fwd_dgrad_wgrad(B, L, Cin, Cout, K, pad):
    _check_conv(B, L, Cin, Cout, K, pad)


@pytest.mark.parametrize("B,L,Cin,Cout,K,pad", [(6, 500, 16, 16, 3, 1), (5, 360, 16, 32, 31, 15), (3, 512, 32, 16, 7, 3), (4, 128, 32, 64, 5, 0),
                                                (2, 1000, 16, 48, 9, 4), (3, 400, 16, 16, 4, 3)])
def test_conv1d_channel_packed_rows(monkeypatch, B, L, Cin, Cout, K, pad):
    """narrow layers as 64-channel layers on packed positions (conv1d_narrow.cu: 4 or 2 positions per 128-byte row, packed taps
    with zero blocks) forced on small shapes, forward + data gradient + the packed weight gradient folded back"""
    monkeypatch.setenv("CS_CONV_PACK_MIN", "0")
    monkeypatch.setenv("CS_CONV_PACK_WGRAD", "1")
    _check_conv(B, L, Cin, Cout, K, pad)


def _check_conv(B, L, Cin, Cout, K, pad):
    from crossscale.ops.conv1d import Conv1dB200
    torch.manual_seed(B * 100 + K)
    m = Conv1dB200(Cin, Cout, K, padding=pad, bias=True).to(DEV)
    with torch.no_grad():
        m.weight.copy_(_bf(m.weight)); m.bias.copy_(torch.randn(Cout, device=DEV))
    x = _bf(torch.randn(B, L, Cin, device=DEV)).to(torch.bfloat16).requires_grad_(True)     # channels-last
    y = m(x)
    Lout = L + 2 * pad - K + 1
    assert y.shape == (B, Lout, Cout) and y.dtype == torch.bfloat16
    xr = x.detach().float().permute(0, 2, 1).contiguous().requires_grad_(True)              # [B,C,L] fp32 oracle
    assert tuple(m.weight.shape) == (K, Cout, Cin)                                          # tap-major parameter
    wr = m.weight_oik.detach().clone().contiguous().requires_grad_(True)                   # PyTorch [Cout,Cin,K] oracle
    br = m.bias.detach().clone().requires_grad_(True)
    yr = F.conv1d(xr, wr, br, padding=pad)
    assert _rel(y.float().permute(0, 2, 1), yr) < 1e-2
    dy = _bf(torch.randn(B, Lout, Cout, device=DEV))
    y.backward(dy.to(torch.bfloat16))
    yr.backward(dy.permute(0, 2, 1))
    assert _rel(x.grad.float().permute(0, 2, 1), xr.grad) < 1e-2           # dgrad (bf16 output rounding)
    assert _rel(m.weight.grad.permute(1, 2, 0), wr.grad) < 5e-3             # wgrad (fp32 output)
    assert _rel(m.bias.grad, br.grad) < 5e-3


@pytest.mark.parametrize("B,L,Cin,Cout,K,pad", [(5, 300, 128, 256, 7, 3), (3, 200, 64, 128, 5, 2), (9, 45, 256, 512, 7, 3), (2, 130, 64, 64, 3, 1)])
def test_conv1d_cta_pair_kernel_matches_single_cta(monkeypatch, B, L, Cin, Cout, K, pad):
    """cta_group::2 path (two CTAs of a cluster issue one M256 MMA, each staging half of the weight tile) forced on
    small shapes — odd tile counts, packed short windows, both weight layouts — against the single-CTA kernel and
    the fp32 oracle"""
    from crossscale.ops import _lib
    lib, st = _lib.cuda_lib(), _lib.stream_ptr()
    torch.manual_seed(K + B)
    x = _bf(torch.randn(B, L, Cin, device=DEV)).to(torch.bfloat16)
    w = _bf(torch.randn(K, Cout, Cin, device=DEV) * 0.1).to(torch.bfloat16)
    bias = torch.randn(Cout, device=DEV)
    Lout = L + 2 * pad - K + 1
    dy = _bf(torch.randn(B, Lout, Cout, device=DEV)).to(torch.bfloat16)
    out = {}
    for mode in ("0", "2"):
        monkeypatch.setenv("CS_CONV_PAIR", mode)
        y = torch.zeros(B, Lout, Cout, dtype=torch.bfloat16, device=DEV)
        dx = torch.zeros(B, L, Cin, dtype=torch.bfloat16, device=DEV)
        _lib.check(lib.cs_conv1d_fwd(x.data_ptr(), w.data_ptr(), bias.data_ptr(), y.data_ptr(), B, L, Cin, Cout, K, pad, 1, 1, st))
        _lib.check(lib.cs_conv1d_dgrad_direct(dy.data_ptr(), w.data_ptr(), dx.data_ptr(), B, L, Cin, Cout, K, pad, 1, st))
        torch.cuda.synchronize()
        out[mode] = (y.float(), dx.float())
    yr = F.relu(F.conv1d(x.float().permute(0, 2, 1), w.float().permute(1, 2, 0), bias, padding=pad)).permute(0, 2, 1)
    assert _rel(out["2"][0], yr) < 1e-2
    assert torch.equal(out["0"][0], out["2"][0])            # same MMA order per output element: bit-identical
    assert _rel(out["2"][1], out["0"][1]) < 1e-6


def test_conv1d_relu_epilogue_and_cin1():
    from crossscale.ops import _lib
    from crossscale.ops.conv1d import Conv1dB200, conv1d_fwd_nlc
    torch.manual_seed(1)
    x = _bf(torch.randn(2, 200, 32, device=DEV)).to(torch.bfloat16)
    w = _bf(torch.randn(5, 64, 32, device=DEV) * 0.1).to(torch.bfloat16)
    bias = torch.randn(64, device=DEV)
    y = conv1d_fwd_nlc(x, w, bias, 2, relu=True)
    ref = F.relu(F.conv1d(x.float().permute(0, 2, 1), w.float().permute(1, 2, 0).contiguous(), bias, padding=2))
    assert _rel(y.float().permute(0, 2, 1), ref) < 1e-2 and float(y.min()) >= 0
    # first layer: [B,1,L] fp32 -> [B,L,C] bf16, and its weight gradient
    m = Conv1dB200(1, 64, 7, padding=3, bias=True).to(DEV)
    xin = torch.randn(6, 1, 500, device=DEV)
    out = m(xin)
    wr, br = m.weight_oik.detach().clone().contiguous().requires_grad_(True), m.bias.detach().clone().requires_grad_(True)
    ref = F.conv1d(xin, wr, br, padding=3)
    assert out.shape == (6, 500, 64) and _rel(out.float().permute(0, 2, 1), ref) < 1e-2
    dy = _bf(torch.randn(6, 500, 64, device=DEV))
    out.backward(dy.to(torch.bfloat16)); ref.backward(dy.permute(0, 2, 1))
    assert _rel(m.weight.grad.permute(1, 2, 0), wr.grad) < 1e-3 and _rel(m.bias.grad, br.grad) < 1e-3


@pytest.mark.parametrize("relu,with_res,C", [(True, False, 64), (True, True, 64), (False, False, 64), (True, True, 768), (True, False, 16)])
def test_bn_relu_residual_kernels(relu, with_res, C):
    from crossscale.ops.norm_act import BNReLU1d
    torch.manual_seed(3)
    B, L = 8, 250
    bn = BNReLU1d(C, relu=relu).to(DEV)
    with torch.no_grad():
        bn.weight.copy_(torch.rand(C, device=DEV) + 0.5); bn.bias.copy_(torch.randn(C, device=DEV) * 0.1)
    x = _bf(torch.randn(B, L, C, device=DEV) * 2 + 0.3).to(torch.bfloat16).requires_grad_(True)
    res = _bf(torch.randn(B, L, C, device=DEV)).to(torch.bfloat16).requires_grad_(True) if with_res else None
    y = bn(x, res)
    xr = x.detach().float().permute(0, 2, 1).contiguous().requires_grad_(True)
    rr = res.detach().float().permute(0, 2, 1).contiguous().requires_grad_(True) if with_res else None
    g, bta = bn.weight.detach().clone().requires_grad_(True), bn.bias.detach().clone().requires_grad_(True)
    yr = F.batch_norm(xr, None, None, g, bta, True, 0.1, 1e-5)
    if with_res:
        yr = yr + rr
    if relu:
        yr = F.relu(yr)
    assert _rel(y.float().permute(0, 2, 1), yr) < 1e-2
    dy = _bf(torch.randn(B, L, C, device=DEV))
    y.backward(dy.to(torch.bfloat16)); yr.backward(dy.permute(0, 2, 1))
    assert _rel(x.grad.float().permute(0, 2, 1), xr.grad) < 3e-2
    assert _rel(bn.weight.grad, g.grad) < 2e-2 and _rel(bn.bias.grad, bta.grad) < 2e-2
    if with_res:
        assert _rel(res.grad.float().permute(0, 2, 1), rr.grad) < 1e-2
    ref_bn = torch.nn.BatchNorm1d(C).to(DEV)
    ref_bn(xr.detach())
    assert _rel(bn.running_mean, ref_bn.running_mean) < 1e-2 and _rel(bn.running_var, ref_bn.running_var) < 1e-2


def test_bn_workspace_is_self_cleaning_and_grads_can_go_straight_to_the_arena():
    """two steps through the same layer: the persistent accumulators / tickets are zero again after every kernel, the
    second step matches the first, and with ``fused_grad`` the reduction kernel adds dgamma / dbeta into .grad itself"""
    from crossscale.ops.norm_act import BNReLU1d
    torch.manual_seed(5)
    B, L, C = 6, 200, 128
    x = _bf(torch.randn(B, L, C, device=DEV)).to(torch.bfloat16)
    dy = _bf(torch.randn(B, L, C, device=DEV)).to(torch.bfloat16)
    bn = BNReLU1d(C, relu=True).to(DEV)
    grads = []
    for fused in (False, True, True):
        bn.fused_grad = fused
        bn.weight.grad = torch.zeros(C, device=DEV) if fused else None
        bn.bias.grad = torch.zeros(C, device=DEV) if fused else None
        xi = x.clone().requires_grad_(True)
        bn(xi).backward(dy)
        torch.cuda.synchronize()
        acc, ticket = bn._ws
        assert float(acc.abs().max()) == 0.0 and int(ticket.abs().max()) == 0
        grads.append((xi.grad.float().clone(), bn.weight.grad.clone(), bn.bias.grad.clone()))
    for a, b in zip(grads[0], grads[1]):
        assert _rel(a, b) < 1e-5
    for a, b in zip(grads[1], grads[2]):
        assert _rel(a, b) < 1e-5


@pytest.mark.parametrize("B,L,Cin,Cout,K", [(6, 360, 64, 128, 7), (9, 45, 128, 256, 7), (3, 130, 256, 512, 3), (5, 200, 64, 64, 5)])
def test_conv_epilogue_produces_the_bn_statistics(B, L, Cin, Cout, K):
    """conv_bn (statistics summed in the convolution's epilogue, finished by its last CTA) against conv followed by the
    separate statistics kernel: same outputs, running statistics and gradients; the workspace is left zero"""
    import copy
    from crossscale.ops.conv1d import Conv1dB200
    from crossscale.ops.norm_act import BNReLU1d, conv_bn
    torch.manual_seed(B + K)
    conv = Conv1dB200(Cin, Cout, K, padding=K // 2, bias=False).to(DEV)
    with torch.no_grad():
        conv.weight.copy_(_bf(conv.weight))
    bns = [BNReLU1d(Cout, relu=True).to(DEV) for _ in range(2)]
    with torch.no_grad():
        bns[0].weight.copy_(torch.rand(Cout, device=DEV) + 0.5); bns[0].bias.copy_(torch.randn(Cout, device=DEV) * 0.1)
    bns[1].load_state_dict(copy.deepcopy(bns[0].state_dict()))
    x = _bf(torch.randn(B, L, Cin, device=DEV)).to(torch.bfloat16)
    res = _bf(torch.randn(B, L, Cout, device=DEV)).to(torch.bfloat16)
    dy = _bf(torch.randn(B, L, Cout, device=DEV)).to(torch.bfloat16)
    outs = []
    for fused, bn in zip((False, True), bns):
        conv.weight.grad = None
        xi = x.clone().requires_grad_(True)
        y = conv_bn(conv, bn, xi, residual=res) if fused else bn(conv(xi), res)
        y.backward(dy)
        torch.cuda.synchronize()
        acc, ticket = bn._ws
        assert float(acc.abs().max()) == 0.0 and int(ticket.abs().max()) == 0
        outs.append((y.float(), xi.grad.float(), conv.weight.grad.clone(), bn.weight.grad.clone(), bn.bias.grad.clone(),
                     bn.running_mean.clone(), bn.running_var.clone()))
    for a, b in zip(*outs):
        assert _rel(a, b) < 2e-3, (_rel(a, b))
    yr = F.batch_norm(F.conv1d(x.float().permute(0, 2, 1), conv.weight_oik.detach(), padding=K // 2), None, None, bns[0].weight.detach(),
                      bns[0].bias.detach(), True, 0.1, 1e-5)
    assert _rel(outs[1][0].permute(0, 2, 1), F.relu(yr + res.float().permute(0, 2, 1))) < 1.5e-2


def test_pool_and_layout_kernels():
    from crossscale.ops import _lib
    from crossscale.ops.norm_act import global_avg_pool1d
    x = _bf(torch.randn(5, 123, 48, device=DEV)).to(torch.bfloat16).requires_grad_(True)
    p = global_avg_pool1d(x)
    torch.testing.assert_close(p, x.detach().float().mean(1), rtol=1e-4, atol=1e-5)
    p.backward(torch.ones_like(p))
    torch.testing.assert_close(x.grad.float(), torch.full_like(x.grad.float(), _bf(torch.tensor(1 / 123)).item()), rtol=1e-2, atol=1e-4)
    from crossscale.ops.norm_act import avg_pool2
    z = _bf(torch.randn(3, 91, 32, device=DEV)).to(torch.bfloat16).requires_grad_(True)
    q = avg_pool2(z)
    ref = F.avg_pool1d(z.detach().float().permute(0, 2, 1), 2).permute(0, 2, 1)
    assert q.shape == (3, 45, 32) and _rel(q.float(), ref) < 1e-2
    q.backward(torch.ones_like(q))
    assert float(z.grad[:, :90].float().mean()) == 0.5 and float(z.grad[:, 90].float().abs().max()) == 0.0
    lib = _lib.cuda_lib()
    a = torch.randn(3, 40, 77, device=DEV)
    nlc = torch.empty(3, 77, 40, dtype=torch.bfloat16, device=DEV)
    _lib.check(lib.cs_ncl_f32_to_nlc_bf16(a.data_ptr(), nlc.data_ptr(), 3, 40, 77, _lib.stream_ptr()))
    torch.testing.assert_close(nlc.float(), _bf(a).permute(0, 2, 1))
    back = torch.empty_like(a)
    _lib.check(lib.cs_nlc_bf16_to_ncl_f32(nlc.data_ptr(), back.data_ptr(), 3, 40, 77, _lib.stream_ptr()))
    torch.testing.assert_close(back, _bf(a))


def test_resnet1d_small_matches_cpu_reference():
    import copy
    from crossscale.models.resnet1d_wide import resnet1d_small
    torch.manual_seed(0)
    cpu = resnet1d_small(5)
    with torch.no_grad():
        for p in cpu.parameters():
            p.copy_(_bf(p))
    gpu = copy.deepcopy(cpu).to(DEV)
    x = _bf(torch.randn(16, 1, 360))
    y = torch.randint(0, 5, (16,))
    lc = F.cross_entropy(cpu(x), y); lc.backward()
    lg = F.cross_entropy(gpu(x.to(DEV)), y.to(DEV)); lg.backward()
    assert abs(float(lc) - float(lg)) < 3e-2
    num = den = 0.0
    for (n, a), b in zip(cpu.named_parameters(), gpu.parameters()):
        num += float((a.grad - b.grad.cpu()).double().pow(2).sum()); den += float(a.grad.double().pow(2).sum())
    assert (num / den) ** 0.5 < 0.1          # whole-gradient relative error through 5 bf16 conv layers + BN


def test_kernel_trainer_resnet_small_learns_and_grad_sync_single_rank():
    """ResNet-1D through KernelLocalTrainer (conv/BN/CE/SGD kernels), with the bucketed gradient
    all-reduce enabled (world 1: the collective degenerates to a scale by 1, exercising hooks/streams)."""
    from crossscale.data.synth import make_mitbih_like_beats
    from crossscale.fl.engine import KernelLocalTrainer
    from crossscale.models.resnet1d_wide import resnet1d_small
    torch.manual_seed(0)
    x, y = make_mitbih_like_beats(512, 360, 5, seed=5)
    x, y = torch.from_numpy(x).to(DEV).unsqueeze(1), torch.from_numpy(y).to(DEV)
    for sync in (False, True):
        tr = KernelLocalTrainer(resnet1d_small(5).to(DEV), DEV, lr=5e-2, sync_grads=sync, bucket_bytes=8 << 10)
        losses = []
        for step in range(12):
            i = (step * 64) % 448
            losses.append(float(tr.step(x[i:i + 64], y[i:i + 64])))
        assert losses[-1] < losses[0] and all(l == l for l in losses), losses
        if sync:
            assert tr.sync.launches >= 12 * len(tr.sync.buckets) and len(tr.sync.buckets) >= 2
            assert tr.sync.comm.status() == 0


def test_kernel_trainer_whole_step_graph_matches_eager():
    import copy
    from crossscale.fl.engine import KernelLocalTrainer
    from crossscale.models.resnet1d_wide import resnet1d_small
    torch.manual_seed(1)
    base = resnet1d_small(5).to(DEV)
    xs = [torch.randn(32, 1, 360, device=DEV) for _ in range(8)]
    ys = [torch.randint(0, 5, (32,), device=DEV) for _ in range(8)]
    outs = []
    for graph in (False, True):
        tr = KernelLocalTrainer(copy.deepcopy(base), DEV, lr=2e-2, use_graph=graph)
        losses = [float(tr.step(x, y)) for x, y in zip(xs, ys)]
        outs.append((tr.flat.flat.clone(), losses))
        assert (tr._graph is not None) == graph
    assert _rel(outs[1][0], outs[0][0]) < 2e-2                      # atomics order differs: not bit-exact
    assert all(abs(a - b) < 0.05 for a, b in zip(outs[0][1], outs[1][1]))


def test_fedavg_runner_with_resnet_small(synth_shards=None):
    import glob, os, tempfile
    from crossscale.data.shard_dataset import load_shards_to_gpu
    from crossscale.data.shard_prep import prepare_shards
    from crossscale.fl.engine import FedAvgRunner
    from crossscale.models.resnet1d_wide import resnet1d_small
    d = tempfile.mkdtemp()
    prepare_shards("mitbih_like", win_len=360, shard_size=512, out_dir=d, num_windows=1024, results_dir=None, verbose=False)
    xg, yg = load_shards_to_gpu(sorted(glob.glob(os.path.join(d, "ecg_*.bin"))), torch.device(DEV))
    r = FedAvgRunner(resnet1d_small(5).to(DEV), torch.device(DEV), xg, yg, batch_size=64, local_steps=4, config="G1",
                     model_name="resnet1d_small", wire="bf16")
    rows = r.run(3)
    assert not r.fused and r.kernel_generic and rows[-1]["avg_loss"] < rows[0]["avg_loss"] + 0.05
    assert rows[0]["comm_ms"] > 0 and r.agg.backend == "kernel"
    r.close()


@pytest.mark.parametrize("B,L,K", [(64, 500, 3), (512, 500, 7), (5, 1300, 32), (3, 40, 40)])
def test_conv1d_batch_valid_matches_reference_semantics(B, L, K):
    """GPU arm of the reference's native kernel (single-channel valid cross-correlation, conv1d_openmp_simd.c:21-61)
    against F.conv1d in fp32."""
    from crossscale.ops import _lib
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    x, w = torch.randn(B, L, device=dev), torch.randn(K, device=dev)
    y = torch.empty(B, L - K + 1, device=dev)
    _lib.check(_lib.cuda_lib().cs_conv1d_batch_valid(x.data_ptr(), w.data_ptr(), y.data_ptr(), B, L, K, _lib.stream_ptr()))
    ref = torch.nn.functional.conv1d(x.unsqueeze(1), w.view(1, 1, K)).squeeze(1)
    torch.testing.assert_close(y, ref, rtol=1e-5, atol=1e-5)
