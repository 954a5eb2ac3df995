"""Small workload for compute-sanitizer --tool memcheck: every kernel family once at tiny sizes."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from crossscale.models import TinyECG
from crossscale.models.resnet1d_wide import resnet1d_small
from crossscale.ops import _lib, tiny_ecg as tk
from crossscale.parallel.fedavg import aggregate_local_group
from crossscale.parallel.symm import LocalGroup

dev = torch.device("cuda", 0)
torch.manual_seed(0)
m = TinyECG(5).to(dev)
x = torch.randn(12, 1, 360, device=dev)
y = torch.randint(0, 5, (12,), device=dev)
for amp in (tk.AMP_FP32, tk.AMP_BF16, tk.AMP_FP16):
    flat = m.flatten_parameters()
    st = tk.TinyTrainState(flat.flat.clone(), 5, amp=amp)
    st.step(x, y); st.step(x, y)
    out = tk.forward_logits(x, flat.flat, 5, amp)
    g = tk.backward_grads(x, flat.flat, torch.randn(12, 5, device=dev), 5, amp)
torch.cuda.synchronize()
print("tiny ok", float(out.abs().mean()))
r = resnet1d_small(5).to(dev)
loss = F.cross_entropy(r(x).float(), y)
loss.backward()
torch.cuda.synchronize()
print("resnet ok", float(loss))
for n in (1458, 70000):
    grp = LocalGroup(4, int(_lib.cuda_lib().cs_fedavg_arena_bytes(n, 0, 4)))
    prev = torch.randn(n, device=dev)
    th = [prev + 0.01 * torch.randn(n, device=dev) for _ in range(4)]
    aggregate_local_group(grp, th, [prev.clone() for _ in range(4)], [0.25] * 4, n)
    torch.cuda.synchronize()
    print("fedavg ok", n, grp.status())
    grp.close()
