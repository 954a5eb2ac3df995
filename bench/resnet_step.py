"""Local training step of the widened ResNet-1D: this framework's kernels (tcgen05 conv, fused BN/ReLU, CE, flat
SGD) vs the same topology in stock PyTorch (nn.Conv1d / BatchNorm1d / AvgPool1d -> cuDNN, bf16 autocast, torch SGD).

    python bench/resnet_step.py [--batch 64 --L 360 --steps 10]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn as nn
import torch.nn.functional as F

from crossscale.fl.engine import KernelLocalTrainer
from crossscale.models.resnet1d_wide import resnet1d_wide_25m


class StockBlock(nn.Module):
    def __init__(self, cin, cout, k=7, down=False):
        super().__init__()
        self.pool = down
        self.c1 = nn.Conv1d(cin, cout, k, padding=k // 2, bias=False); self.b1 = nn.BatchNorm1d(cout)
        self.c2 = nn.Conv1d(cout, cout, k, padding=k // 2, bias=False); self.b2 = nn.BatchNorm1d(cout)
        self.d = None
        if cin != cout:
            self.d = nn.Conv1d(cin, cout, 1, bias=False); self.db = nn.BatchNorm1d(cout)

    def forward(self, x):
        if self.pool:
            x = F.avg_pool1d(x, 2)
        idn = x if self.d is None else self.db(self.d(x))
        y = F.relu(self.b1(self.c1(x)))
        return F.relu(self.b2(self.c2(y)) + idn)


class StockResNet(nn.Module):
    def __init__(self, nc=5, widths=(128, 256, 512, 768), stem=64, k=7):
        super().__init__()
        self.stem = nn.Conv1d(1, stem, 7, padding=3, bias=False); self.sb = nn.BatchNorm1d(stem)
        layers, cin = [], stem
        for i, w in enumerate(widths):
            for b in range(2):
                layers.append(StockBlock(cin, w, k, down=(b == 0 and i > 0))); cin = w
        self.layers = nn.Sequential(*layers)
        self.head = nn.Linear(cin, nc)

    def forward(self, x):
        y = F.relu(self.sb(self.stem(x)))
        return self.head(self.layers(y).mean(-1))


def timed(fn, steps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--L", type=int, default=360)
    ap.add_argument("--steps", type=int, default=10)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    x = torch.randn(a.batch, 1, a.L, device=dev)
    y = torch.randint(0, 5, (a.batch,), device=dev)
    ours = KernelLocalTrainer(resnet1d_wide_25m(5).to(dev), dev)
    t_ours = timed(lambda: ours.step(x, y), a.steps)
    graphed = KernelLocalTrainer(resnet1d_wide_25m(5).to(dev), dev, use_graph=True)
    for _ in range(5):
        graphed.step(x, y)
    t_graph = timed(lambda: graphed.step(x, y), a.steps)
    l0 = float(graphed.step(x, y)); [graphed.step(x, y) for _ in range(20)]; l1 = float(graphed.step(x, y))
    stock = StockResNet().to(dev)
    opt = torch.optim.SGD(stock.parameters(), lr=1e-2, momentum=0.9)

    def stock_step():
        opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            loss = F.cross_entropy(stock(x), y)
        loss.backward()
        opt.step()

    t_stock = timed(stock_step, a.steps)
    stock_cl = StockResNet().to(dev)
    print(f"   whole-step CUDA graph: {t_graph:.2f} ms/step ({a.batch / t_graph * 1e3:,.0f} samples/s), x{t_stock / t_graph:.2f} vs stock; "
          f"loss {l0:.4f} -> {l1:.4f} after 20 more graphed steps")
    print(f"ResNet-1D wide, B={a.batch} L={a.L}: ours {t_ours:.2f} ms/step ({a.batch / t_ours * 1e3:,.0f} samples/s) | "
          f"stock PyTorch/cuDNN bf16 autocast {t_stock:.2f} ms/step ({a.batch / t_stock * 1e3:,.0f} samples/s) | x{t_stock / t_ours:.2f}")


if __name__ == "__main__":
    main()
