"""Regenerate profiles/sass/: per-kernel mnemonic histograms + the FULL `cuobjdump -sass` listing of every hot kernel
(gzip-compressed, one file per kernel)."""
import collections
import gzip
import os
import re
import subprocess
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(REPO, "lib", "libcrossscale_cuda.so")
OUT = os.path.join(REPO, "profiles", "sass")
KERNELS = ["tiny_round_tf32_kernel", "conv1d_fwd_narrow_kernel", "conv1d_wgrad_narrow_kernel", "pack_weight_kernel", "unpack_wgrad_kernel",
           "mma_layout_probe_kernel", "tma_stream_probe_kernel", "tiny_round_kernelILi4ELb0E", "tiny_round_kernelILi4ELb1E", "fedavg_bucket_kernel", "fedavg_oneshot_kernelILb1E",
           "bn_stats_finalize", "bn_bwd_reduce_ws", "mma_probe_kernelILi64ELi16ELb1ELi1E", "tiny_umma_kernelILi4ELb0ELi2E", "tiny_umma_kernelILi4ELb1ELi2E", "tiny_ecg_f32_kernelILi2E", "conv1d_fwd_kernel",
           "conv1d_wgrad_kernel", "conv1d_fwd_sw128_kernel", "conv1d_fwd_pair_kernel", "conv1d_wgrad_sw128_kernel", "bn_apply_coef_kernel",
           "bn_bwd_apply_coef_kernel", "fedavg_oneshot_kernelILb0E", "fedavg_twoshot_kernelILb0E", "fedavg_twoshot_kernelILb1E",
           "fedavg_inplace_kernel", "fedavg_nvls_kernel", "barrier_kernel", "bn_stats_kernel", "bn_bwd_reduce_kernel",
           "sgd_momentum_kernel", "gather_rows_kernel", "ce_fwd_bwd_kernel", "cin1_fwd_kernel", "cin1_wgrad_kernel"]
MN = re.compile(r"^\s+/\*[0-9a-f]{4,5}\*/\s+(?:@!?U?P\d+\s+)?([A-Za-z0-9_.]+)")
KEEP = re.compile(r"UTC|LDTM|STTM|UTMA|SYNCS|REDG|ATOMG|MEMBAR|ERRBAR|ELECT|ACQBULK|UBLKCP|FFMA$|HMMA|LDGMC|STG.*MMEM|.*\.SYS|LDG\.E\.128|STG\.E\.128")

HEAD = """# SASS evidence (`cuobjdump -sass lib/libcrossscale_cuda.so`, sm_100a)

Histogram of the tensor-core / TMA / TMEM / NVLink-relevant mnemonics per kernel; the complete listing of each kernel is
next to this file as `<mangled name>.sass.gz` (`zcat` it).
`UTCHMMA` = tcgen05.mma kind::f16, `LDTM` = tcgen05.ld, `UTCBAR` = tcgen05.commit, `UTCATOMSWS` = tcgen05.alloc/dealloc,
`UTMALDG.4D` / `UTMALDG.3D` = cp.async.bulk.tensor (TMA load), `UTMASTG.3D` = TMA store, `SYNCS.*` = mbarrier, `REDG.E.ADD.F32x4` = red.global.add.v4.f32,
`STG/LDG.E.STRONG.SYS` + `MEMBAR.*.SYS` = the st.release.sys / ld.acquire.sys flag protocol of the cross-GPU barrier
(the data loads/stores on peer-mapped pointers are ordinary `LDG`/`STG`: the address aperture routes them over NVLink);
`LDGMC.E.ADD.F32x4...SYS` = multimem.ld_reduce (NVSwitch in-fabric reduction), the matching multimem.st shows as a `.SYS` store.
"""


def tf32_evidence(sass: str) -> str:
    """kind::tf32 is not a SASS mnemonic of its own: `UTCHMMA` takes the operand format from the instruction descriptor.  Show
    both ends: the PTX nvcc emits for tiny_ecg_round_tf32.cu and the descriptor constants in the kernel's SASS."""
    def idesc(fmt, M, N, am, bm):
        return (1 << 4) | (fmt << 7) | (fmt << 10) | (am << 15) | (bm << 16) | ((N >> 3) << 17) | ((M >> 4) << 24)
    src = os.path.join(REPO, "csrc", "cuda", "tiny_ecg_round_tf32.cu")
    ptx = subprocess.run(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-std=c++17", "-O3", "--expt-relaxed-constexpr",
                          "-I" + os.path.join(REPO, "include"), "-I" + os.path.join(REPO, "csrc", "cuda"), "-ptx", src, "-o", "/dev/stdout"],
                         capture_output=True, text=True).stdout
    mma = [l.strip() for l in ptx.split("\n") if "tcgen05.mma" in l]
    body = sass.split("Function : _ZN2cs22tiny_round_tf32_kernel", 1)
    body = body[1].split("Function : ", 1)[0] if len(body) > 1 else ""
    consts = {"forward / data gradient, M128 N16, K-major, a/b format 2 = tf32": idesc(2, 128, 16, 0, 0),
              "weight gradients, M128 N32, MN-major, tf32": idesc(2, 128, 32, 1, 1),
              "weight gradients tap 4, M64 N32, MN-major, tf32": idesc(2, 64, 32, 1, 1)}
    out = ["\n## kind::tf32 in `tiny_round_tf32_kernel` (the G0 / fp32 round)\n",
           f"PTX (`nvcc -ptx csrc/cuda/tiny_ecg_round_tf32.cu`): {len(mma)} `tcgen05.mma` instructions, all `kind::tf32`: "
           f"{sum('kind::tf32' in l for l in mma)}; first: `{mma[0] if mma else 'n/a'}`\n",
           "SASS: the same `UTCHMMA` as the 16-bit kernels; the operand format is bits 7-9 / 10-12 of the instruction descriptor "
           "(0 = f16, 1 = bf16, 2 = tf32), loaded as an immediate:\n", "| descriptor | value | `UMOV UR, value` in the kernel's SASS |", "|---|---|---|"]
    for k, v in consts.items():
        out.append(f"| {k} | 0x{v:x} | {len(re.findall(r'UMOV UR\d+, 0x%x' % v, body))} |")
    out.append(f"\nfor comparison the bf16 round kernel's forward descriptor is 0x{idesc(1, 128, 16, 0, 0):x} (format 1).")
    return "\n".join(out)


def main():
    os.makedirs(OUT, exist_ok=True)
    for old in os.listdir(OUT):                            # stale excerpts of kernels that changed signature
        if old.endswith(".sass.txt") or old.endswith(".sass.gz"):
            os.remove(os.path.join(OUT, old))
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    md = [HEAD]
    for f in re.split(r"\n\s*Function : ", sass)[1:]:
        name = f.split("\n", 1)[0].strip()
        if not any(k in name for k in KERNELS):
            continue
        lines = f.split("\n")
        cnt = collections.Counter(m.group(1) for m in map(MN.match, lines) if m)
        md.append(f"\n## `{name}`\n\ntotal SASS instructions: {sum(cnt.values())}\n\n| mnemonic | count |\n|---|---|")
        md += [f"| {k} | {cnt[k]} |" for k in sorted((k for k in cnt if KEEP.match(k)), key=lambda k: -cnt[k])]
        short = re.sub(r"[^A-Za-z0-9]+", "_", name)[:72]
        with gzip.open(os.path.join(OUT, short + ".sass.gz"), "wt") as fh:
            fh.write(f"// cuobjdump -sass of {name}\n" + "\n".join(l.rstrip() for l in lines) + "\n")
        md.append(f"\nfull listing: `{short}.sass.gz`")
    md.append(tf32_evidence(sass))
    with open(os.path.join(OUT, "README.md"), "w") as fh:
        fh.write("\n".join(md) + "\n")
    print("wrote", OUT)


if __name__ == "__main__":
    sys.exit(main())
