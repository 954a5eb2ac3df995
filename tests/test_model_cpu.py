"""TinyECG: state_dict contract (Module_3/tiny_ecg_model.py:8-29), flat arena, checkpoints."""
import torch
import torch.nn as nn

from crossscale.models import TinyECG, Tiny1D, num_params, param_offsets, build_model
from crossscale.utils.misc import load_checkpoint, save_checkpoint


class StockTinyECG(nn.Module):
    """An independent stock-PyTorch statement of the reference topology (oracle)."""

    def __init__(self, num_classes=2):
        super().__init__()
        self.net = nn.Sequential(nn.Conv1d(1, 16, 7, padding=3), nn.ReLU(), nn.Conv1d(16, 16, 5, padding=2),
                                 nn.ReLU(), nn.AdaptiveAvgPool1d(1))
        self.head = nn.Linear(16, num_classes)

    def forward(self, x):
        return self.head(self.net(x).squeeze(-1))


def test_state_dict_contract():
    for c, total in ((2, 1458), (5, 1509)):
        m = TinyECG(num_classes=c)
        sd = m.state_dict()
        assert list(sd.keys()) == ["net.0.weight", "net.0.bias", "net.2.weight", "net.2.bias",
                                   "head.weight", "head.bias"]
        assert [tuple(v.shape) for v in sd.values()] == [(16, 1, 7), (16,), (16, 16, 5), (16,), (c, 16), (c,)]
        assert all(v.dtype == torch.float32 for v in sd.values())
        assert sum(v.numel() for v in sd.values()) == total == num_params(c)
        assert len(list(m.buffers())) == 0


def test_stock_state_dict_loads_and_matches():
    torch.manual_seed(0)
    stock, ours = StockTinyECG(5), TinyECG(5)
    ours.load_state_dict(stock.state_dict())
    x = torch.randn(8, 1, 360)
    torch.testing.assert_close(ours(x), stock(x))
    stock2 = StockTinyECG(5)
    stock2.load_state_dict(ours.state_dict())
    torch.testing.assert_close(stock2(x), stock(x))


def test_flat_arena_views_and_offsets():
    m = TinyECG(2)
    before = {k: v.clone() for k, v in m.state_dict().items()}
    flat = m.flatten_parameters()
    assert flat.numel == 1458 and flat.flat.numel() % 4 == 0
    for k, (off, n) in param_offsets(2).items():
        torch.testing.assert_close(flat.flat[off:off + n], before[k].reshape(-1))
    # views: writing the arena changes the module and vice versa
    flat.flat[:112] = 0.25
    assert float(m.net[0].weight.mean()) == 0.25
    with torch.no_grad():
        m.head.bias.fill_(3.0)
    assert float(flat.flat[1456]) == 3.0
    assert m.flatten_parameters() is flat                      # idempotent
    x = torch.randn(4, 1, 500)
    m(x).sum().backward()
    assert m.net[0].weight.grad is not None


def test_checkpoint_round_trip(tmp_path):
    a, b = TinyECG(5), TinyECG(5)
    a.flatten_parameters()
    p = str(tmp_path / "ck" / "tiny.pt")
    save_checkpoint(a, p)
    sd = torch.load(p, weights_only=True)
    assert list(sd.keys())[0] == "net.0.weight" and all(v.device.type == "cpu" for v in sd.values())
    b.flatten_parameters()
    load_checkpoint(b, p)
    x = torch.randn(3, 1, 360)
    torch.testing.assert_close(a(x), b(x))
    StockTinyECG(5).load_state_dict(sd)                        # a stock module accepts our checkpoint


def test_tiny1d_alias_and_registry():
    assert isinstance(Tiny1D(), TinyECG) and Tiny1D(in_ch=1, num_classes=5).num_classes == 5
    assert isinstance(build_model("tiny_ecg", 2), TinyECG)


def test_resnet1d_cpu_path_trains_and_matches_stock_layers():
    """ResNet-1D on CPU goes through plain PyTorch (F.conv1d / F.batch_norm behind Conv1dB200 / BNReLU1d / conv_bn): the
    oracle the CUDA kernels are tested against.  Parameter count of the wide model, tap-major weight views, one SGD step."""
    import torch.nn.functional as F
    from crossscale.models import build_model
    from crossscale.models.resnet1d_wide import resnet1d_wide_25m
    from crossscale.ops.conv1d import Conv1dB200
    from crossscale.ops.norm_act import BNReLU1d, conv_bn

    torch.manual_seed(0)
    m = build_model("resnet1d_small", 5)
    x, y = torch.randn(6, 1, 120), torch.randint(0, 5, (6,))
    opt = torch.optim.SGD(m.parameters(), lr=0.05, momentum=0.9)
    losses = []
    for _ in range(8):
        loss = F.cross_entropy(m(x), y)
        opt.zero_grad(); loss.backward(); opt.step()
        losses.append(float(loss))
    assert losses[-1] < losses[0] and all(p.grad is not None for p in m.parameters())

    conv, bn = Conv1dB200(8, 16, 5, padding=2, bias=False), BNReLU1d(16, relu=True)
    xc, res = torch.randn(3, 8, 40), torch.randn(3, 16, 40)
    ref = F.relu(F.batch_norm(F.conv1d(xc, conv.weight_oik, padding=2), None, None, bn.weight, bn.bias, True, 0.1, 1e-5) + res)
    assert torch.allclose(conv_bn(conv, bn, xc, residual=res), ref, atol=1e-6)
    assert tuple(conv.weight.shape) == (5, 16, 8) and tuple(conv.weight_oik.shape) == (16, 8, 5)

    wide = resnet1d_wide_25m(5)
    n = sum(p.numel() for p in wide.parameters())
    assert 24_000_000 < n < 25_000_000                      # "~25 M parameters" stress config of BASELINE.json
