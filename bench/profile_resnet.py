"""ncu launch-list driver: 2 warm + 1 measured ResNet-1D-wide training steps through KernelLocalTrainer."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from crossscale.fl.engine import KernelLocalTrainer
from crossscale.models.resnet1d_wide import resnet1d_wide_25m

dev = torch.device("cuda", 0)
torch.manual_seed(0)
B = int(os.environ.get("B", "64"))
x = torch.randn(B, 1, 360, device=dev)
y = torch.randint(0, 5, (B,), device=dev)
tr = KernelLocalTrainer(resnet1d_wide_25m(5).to(dev), dev)
for _ in range(3):
    tr.step(x, y)
torch.cuda.synchronize()
import time
t0 = time.perf_counter()
for _ in range(5):
    tr.step(x, y)
torch.cuda.synchronize()
print(f"{(time.perf_counter() - t0) / 5 * 1e3:.2f} ms/step wall")
