"""Driver for the persistent FL-round kernel at the headline shape (B=256, L=500, C=2, 50 steps):
times rounds with CUDA events against the 50-launch CUDA-graph round, prints the per-phase clock64 timeline of
CTA 0 (middle step of the round) and the spread of the CTAs' arrival at that step's grid barrier."""
import argparse
import ctypes
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from crossscale.models import TinyECG
from crossscale.ops import _lib, tiny_ecg as tk
from crossscale.utils.clocks import ClockSampler

ap = argparse.ArgumentParser()
ap.add_argument("--amp", default="bf16")
ap.add_argument("--steps", type=int, default=50)
ap.add_argument("--batch", type=int, default=256)
ap.add_argument("--L", type=int, default=500)
ap.add_argument("--rounds", type=int, default=40)
ap.add_argument("--mode", type=int, default=tk.SAMPLE_PERM)
ap.add_argument("--ncu", action="store_true", help="few launches only (for ncu)")
a = ap.parse_args()
dev = torch.device("cuda", 0)
torch.manual_seed(0)
m = TinyECG(2).to(dev)
N = 131072
data = torch.randn(N, a.L, device=dev)
lab = torch.randint(0, 2, (N,), device=dev)
lib = _lib.cuda_lib()


def state():
    return tk.TinyTrainState(m.flatten_parameters().flat.clone(), 2, amp=tk.AMP_BY_NAME[a.amp],
                             scaler=tk.ScalerConfig(enabled=(a.amp == "fp16")))


st = state()
idx = torch.randint(0, N, (a.steps * a.batch,), device=dev)


def one_round():
    st.round(data, lab, a.steps, a.batch, idx=idx if a.mode == tk.SAMPLE_IDX else None, sample_mode=a.mode, seed=77)


if a.ncu:
    for _ in range(3):
        one_round()
    torch.cuda.synchronize()
    sys.exit(0)

sampler = ClockSampler(gpu_index=0).start()
for _ in range(200):
    one_round()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
times = []
for w in range(5):
    e0.record()
    for _ in range(a.rounds):
        one_round()
    e1.record()
    torch.cuda.synchronize()
    times.append(e0.elapsed_time(e1) / a.rounds)
st.check_errors()

# the first-generation round for comparison: `steps` launches replayed from one CUDA graph
st2 = state()
idx2 = idx.view(a.steps, a.batch)
for s in range(a.steps):
    st2.step(data, lab, idx=idx2[s])
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    for s in range(a.steps):
        st2.step(data, lab, idx=idx2[s])
for _ in range(50):
    g.replay()
torch.cuda.synchronize()
e0.record()
for _ in range(a.rounds):
    g.replay()
e1.record()
torch.cuda.synchronize()
graph_ms = e0.elapsed_time(e1) / a.rounds
clocks = sampler.stop()

# fixed cost vs per-step cost: rounds of 1, 2, 5, 10, 25 steps
step_scan = {}
for ns in (1, 2, 5, 10, 25, a.steps):
    idx_s = idx[: ns * a.batch]
    for _ in range(5):
        st.round(data, lab, ns, a.batch, idx=idx_s if a.mode == tk.SAMPLE_IDX else None, sample_mode=a.mode, seed=77)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(20):
        st.round(data, lab, ns, a.batch, idx=idx_s if a.mode == tk.SAMPLE_IDX else None, sample_mode=a.mode, seed=77)
    e1.record()
    torch.cuda.synchronize()
    step_scan[ns] = round(e0.elapsed_time(e1) / 20 * 1e3, 2)

prof = torch.zeros(4096, dtype=torch.int64, device=dev)
lib.cs_debug_set_tiny_profile.argtypes = [ctypes.c_void_p]
lib.cs_debug_set_tiny_profile(prof.data_ptr())
one_round()
torch.cuda.synchronize()
lib.cs_debug_set_tiny_profile(None)
pt = prof.cpu().tolist()
names = {0: "step start", 1: "xs + im2col plane staged", 2: "grid barrier wait + replica sum + SGD + W2 repack",
         3: "conv1 -> h1 planes, handoff", 4: "conv2 fwd MMAs done", 5: "epilogue 1 + pool", 6: "head + CE",
         7: "head grads + dz2 planes, handoff", 8: "dgrad MMAs done", 9: "epilogue 2 dz1 planes, handoff",
         10: "all wgrad MMAs done", 11: "readout + reductions + arrive"}
order = [0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11]
phases = [(names[k], pt[k] - pt[order[i - 1]]) for i, k in enumerate(order) if i > 0 and pt[k] and pt[order[i - 1]]]
lib.cs_debug_last_round_grid.restype = ctypes.c_int
nc = int(lib.cs_debug_last_round_grid())
per = np.array(pt[1024:1024 + 4 * nc]).reshape(nc, 4)
t0 = per[:, 0].min()
start, arrive, passed = per[:, 0] - t0, per[:, 1] - t0, per[:, 2] - t0
arr = per[:, 1]
out = {"amp": a.amp, "B": a.batch, "L": a.L, "steps": a.steps, "sample_mode": a.mode,
       "round_ms_median": float(np.median(times)), "round_ms_min": min(times), "round_ms_max": max(times),
       "us_per_step": float(np.median(times)) / a.steps * 1e3,
       "samples_per_s": a.steps * a.batch / (float(np.median(times)) / 1e3),
       "graph_round_ms": graph_ms, "speedup_vs_graph_round": graph_ms / float(np.median(times)),
       "clocks": clocks,
       "cta0_phase_cycles": phases, "cta0_step_cycles": (pt[11] - pt[0]) if pt[11] and pt[0] else None,
       "barrier_wait_to_update_cycles": (pt[21] - pt[20]) if pt[21] and pt[20] else None,
       "arrive_spread_ns": int(arr.max() - arr.min()) if arr.size else None,
       "arrive_p50_after_first_ns": int(np.median(arr) - arr.min()) if arr.size else None,
       "grid_ctas": nc, "us_per_round_by_steps": step_scan,
       "cta_step_start_ns": {"min": int(start.min()), "p50": int(np.median(start)), "max": int(start.max())},
       "cta_work_ns(start->arrive)": {"min": int((arrive - start).min()), "p50": int(np.median(arrive - start)), "max": int((arrive - start).max())},
       "cta_barrier_passed_ns(after step start)": {"min": int(passed.min()), "p50": int(np.median(passed)), "max": int(passed.max())},
       "loss_sum": st.loss_sum()}
print(json.dumps(out, indent=1))
