import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
torch.zeros(1, device="cuda")
from crossscale.ops import _lib
lib = _lib.cuda_lib()
lib.cs_debug_mma_probe.argtypes = [ctypes.c_int] * 4 + [ctypes.c_void_p]
out = (ctypes.c_longlong * 8)()
names = ["M128 N16 K", "M64 N16 MN", "M64 N32 MN", "M128 N64 K", "M128 N128 K"]
print("shape | rot issuers shift | per-warp (issue, total) cycles for 64 MMAs | cycles per MMA (all warps)")
for cfg in range(5):
    for rot in (1, 4):
        for issuers in (1, 2, 4):
            for shift in (0, 1):
                rc = lib.cs_debug_mma_probe(cfg, rot, issuers, shift, out)
                tot = max(out[2 * w + 1] for w in range(issuers))
                print(names[cfg], "|", rot, issuers, shift, "|", [(out[2 * w], out[2 * w + 1]) for w in range(issuers)], "|",
                      round(tot / (64 * issuers), 1), "rc", rc)
