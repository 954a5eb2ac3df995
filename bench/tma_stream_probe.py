"""What does one SM's TMA unit deliver from cold DRAM?  (`cs_debug_tma_stream_probe`, csrc/cuda/tma_probe.cu.)
A persistent CTA per SM keeps `stages` loads in flight and recycles slots as they complete; 256 MB streamed per case after an
L2 flush.  Prints GB/s over the whole GPU and bytes per clock per SM for tiled boxes with 16 / 32 / 64 / 128-byte rows and for
1-D bulk copies, next to `torch` copy of the same bytes (read + write)."""
import ctypes, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from crossscale.ops import _lib
from crossscale.utils.clocks import ClockSampler

lib = _lib.cuda_lib()
lib.cs_debug_tma_stream_probe.argtypes = [ctypes.c_void_p, ctypes.c_longlong] + [ctypes.c_int] * 5 + [ctypes.c_void_p, ctypes.c_void_p]
dev = torch.device("cuda", 0)
N = 256 << 20
buf = torch.empty(N, dtype=torch.uint8, device=dev).random_(0, 255)
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
sampler = ClockSampler(gpu_index=0).start()
rows = []


def run(mode, row_bytes, nrows, chunk, stages):
    best = None
    for rep in range(3):
        flush.zero_(); torch.cuda.synchronize()
        ms, cyc = ctypes.c_float(0), ctypes.c_longlong(0)
        rc = lib.cs_debug_tma_stream_probe(buf.data_ptr(), N, mode, row_bytes, nrows, chunk, stages, ctypes.byref(ms), ctypes.byref(cyc))
        assert rc == 0, rc
        if best is None or ms.value < best[0]:
            best = (ms.value, cyc.value)
    per = (row_bytes * nrows) if mode == 0 else chunk
    r = {"mode": "tensor box" if mode == 0 else "1-D bulk", "row_bytes": row_bytes if mode == 0 else None, "rows_per_load": nrows if mode == 0 else None,
         "bytes_per_load": per, "stages": stages, "in_flight_kb_per_sm": round(per * stages / 1024, 1), "ms": round(best[0], 4),
         "gb_per_s": round(N / best[0] / 1e6, 1), "bytes_per_clk_per_sm": round(N / 148 / best[1], 2)}
    rows.append(r)
    print(r, flush=True)


for rb in (16, 32, 64, 128):
    for st in (4, 8):
        run(0, rb, 128, 0, st)
run(0, 128, 256, 0, 4)
for chunk in (4096, 16384):
    for st in (4, 8):
        run(1, 0, 0, chunk, st)
run(1, 0, 0, 32768, 4)
a, b = torch.empty(N, dtype=torch.uint8, device=dev), torch.empty(N, dtype=torch.uint8, device=dev)
ts = []
for _ in range(5):
    flush.zero_(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); b.copy_(a); e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
ms = sorted(ts)[2]
print(json.dumps({"rows": rows, "torch_copy_256MB": {"ms": ms, "read_gb_per_s": N / ms / 1e6, "read_plus_write_gb_per_s": 2 * N / ms / 1e6}, "clocks": sampler.stop()}))
