"""CPU behaviour of the op wrappers (fallbacks double as oracles)."""
import torch
import torch.nn.functional as F

from crossscale.models.tiny_ecg import FlatParams
from crossscale.ops.ce_scaler import cross_entropy


def test_cross_entropy_cpu_value_unscaled_gradient_scaled():
    torch.manual_seed(0)
    lg = torch.randn(9, 5, requires_grad=True)
    y = torch.randint(0, 5, (9,))
    loss = cross_entropy(lg, y, torch.tensor([8.0]))
    ref = F.cross_entropy(lg.detach().clone().requires_grad_(True), y)
    assert abs(float(loss) - float(ref)) < 1e-6
    loss.backward()
    lg2 = lg.detach().clone().requires_grad_(True)
    F.cross_entropy(lg2, y).backward()
    torch.testing.assert_close(lg.grad, 8.0 * lg2.grad)


def test_flat_params_alignment_and_rebind():
    m = torch.nn.Sequential(torch.nn.Linear(3, 5), torch.nn.Linear(5, 7))
    ref = [p.detach().clone() for p in m.parameters()]
    fp = FlatParams(m, align=4)
    assert all(o % 4 == 0 for o in fp.offsets) and fp.flat.numel() % 4 == 0
    for p, r in zip(m.parameters(), ref):
        torch.testing.assert_close(p.detach(), r)
    new = torch.zeros(fp.flat.numel() + 16)
    fp.rebind(new)
    for p, r, off in zip(m.parameters(), ref, fp.offsets):
        torch.testing.assert_close(p.detach(), r)
        assert p.data_ptr() == new.data_ptr() + 4 * off          # parameters are views of the new arena
    m(torch.randn(2, 3)).sum().backward()
    assert float(fp.grad.abs().sum()) > 0                          # grads land in the flat gradient arena
