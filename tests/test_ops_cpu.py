"""CPU behaviour of the op wrappers (fallbacks double as oracles)."""
import pytest
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


def test_round_sampler_permutation_mirror_and_bijection():
    """The persistent round kernel draws batches from a counter-based permutation; the Python mirror must be the
    same function as the C one (``cs_perm_index``) and a bijection on [0, n) for every epoch."""
    from crossscale.ops import _lib, tiny_ecg as tk
    lib = _lib.cuda_lib()
    for n in (1, 2, 3, 17, 256, 300, 1000, 20000):
        for epoch in (0, 1, 7):
            rows = [tk.perm_index(i, n, 1234, epoch) for i in range(min(n, 2000))]
            if n <= 2000:
                assert sorted(rows) == list(range(n))
            for i in range(0, min(n, 2000), 37):
                assert rows[i] == lib.cs_perm_index(i, n, 1234, epoch)
    a = [tk.perm_index(i, 1000, 1, 0) for i in range(1000)]
    b = [tk.perm_index(i, 1000, 1, 1) for i in range(1000)]
    c = [tk.perm_index(i, 1000, 2, 0) for i in range(1000)]
    assert a != b and a != c                                 # epochs and seeds give different orders
    # whole batches, ragged tail dropped, an epoch never repeats a row
    rows = tk.perm_batches(0, 256 * 2, 300, 64, 9)
    assert len(set(rows[:256])) == 256 and len(set(rows[256:])) == 256


def test_validate_labels_rejects_out_of_range():
    import torch
    from crossscale.ops import tiny_ecg as tk
    tk.validate_labels(torch.tensor([0, 1, 1, 0]), 2)
    with pytest.raises(ValueError):
        tk.validate_labels(torch.tensor([0, 4]), 2)
    with pytest.raises(ValueError):
        tk.validate_labels(torch.tensor([-1, 0]), 2)
