"""Tiny driver for ncu: N fused TinyECG training steps at the headline shape (B=256, L=500, C=2)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from crossscale.models import TinyECG
from crossscale.ops import tiny_ecg as tk

ap = argparse.ArgumentParser()
ap.add_argument("--amp", default="bf16")
ap.add_argument("--steps", type=int, default=12)
ap.add_argument("--batch", type=int, default=256)
ap.add_argument("--L", type=int, default=500)
a = ap.parse_args()
dev = torch.device("cuda", 0)
torch.manual_seed(0)
m = TinyECG(2).to(dev)
flat = m.flatten_parameters()
st = tk.TinyTrainState(flat.flat, 2, amp=tk.AMP_BY_NAME[a.amp])
data = torch.randn(131072, a.L, device=dev)
lab = torch.randint(0, 2, (131072,), device=dev)
idx = torch.randint(0, 131072, (a.steps, a.batch), device=dev)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for s in range(3):
    st.step(data, lab, idx=idx[s])
torch.cuda.synchronize()
e0.record()
for s in range(a.steps):
    st.step(data, lab, idx=idx[s])
e1.record()
torch.cuda.synchronize()
import ctypes
from crossscale.ops import _lib
prof = torch.zeros(1024, dtype=torch.int64, device=dev)
lib = _lib.cuda_lib()
lib.cs_debug_set_tiny_profile.argtypes = [ctypes.c_void_p]
lib.cs_debug_set_tiny_profile(prof.data_ptr())
st.step(data, lab, idx=idx[0])
torch.cuda.synchronize()
lib.cs_debug_set_tiny_profile(None)
pt = prof.cpu().tolist()
names = ["start", "prologue", "x->xs", "conv1+planes", "fwd issued", "fwd done", "epi1+pool", "head/CE", "dz2 planes",
         "dg+wg issued", "dg done", "epi2 dz1", "wb issued", "wg done", "readout+atomics", "dealloc", "ticket"]
print("phase cycles (CTA 0):", [(names[i], pt[i] - pt[i - 1]) for i in range(1, 17) if pt[i] and pt[i - 1]], "total", pt[16] - pt[0])
import numpy as np
g = np.array(pt[64:64 + 2 * a.batch]).reshape(-1, 2)
t0 = g[:, 0].min()
print("CTA start ns: min 0 max", g[:, 0].max() - t0, "| CTA ticket ns: min", g[:, 1].min() - t0, "median", int(np.median(g[:, 1])) - t0,
      "max", g[:, 1].max() - t0, "| tail end ns", pt[40] - t0, "last CTA", pt[41])
print("tail ns: start", pt[46]-t0, "replicas summed", pt[47]-t0, "synced", pt[48]-t0, "gathered", pt[42]-t0, "checked", pt[43]-t0, "updated", pt[44]-t0, "zeroed", pt[45]-t0, "end", pt[40]-t0)
print(f"amp={a.amp} B={a.batch} L={a.L}: {e0.elapsed_time(e1) / a.steps * 1e3:.2f} us/step (eager launches), loss_sum={st.loss_sum():.4f}")
