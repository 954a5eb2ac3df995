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
