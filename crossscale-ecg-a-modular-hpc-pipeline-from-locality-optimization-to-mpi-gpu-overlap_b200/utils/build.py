"""In-tree native builds (nvcc / g++), no torch headers involved — seconds per file.

  lib/libcrossscale_cuda.so   csrc/cuda/*.cu     -gencode arch=compute_100a,code=sm_100a -lineinfo
  lib/libcrossscale_cpu.so    csrc/cpu/*.cpp     OpenMP + AVX2/AVX-512 (runtime dispatch)
  lib/libcrossscale_rt.so     csrc/runtime/*.cpp std::thread pinned-ring prefetcher

``.so`` files are git-ignored but travel with the gpurun snapshot.  Objects are cached under
``build/`` and rebuilt when a source or header is newer.
"""
from __future__ import annotations

import glob
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from typing import List

REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
LIB_DIR = os.path.join(REPO, "lib")
OBJ_DIR = os.path.join(REPO, "build")
NVCC_ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def _nvcc() -> str | None:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    return None


def _newer(target: str, deps: List[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def _run(cmd: List[str], verbose: bool) -> None:
    if verbose:
        print("[build]", " ".join(cmd), flush=True)
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"build failed: {' '.join(cmd)}\n{r.stdout}")
    if verbose and r.stdout.strip():
        print(r.stdout)


def cuda_lib_path() -> str:
    return os.path.join(LIB_DIR, "libcrossscale_cuda.so")


def cpu_lib_path() -> str:
    return os.path.join(LIB_DIR, "libcrossscale_cpu.so")


def rt_lib_path() -> str:
    return os.path.join(LIB_DIR, "libcrossscale_rt.so")


def build_cuda(verbose: bool = False, force: bool = False, ptxas_verbose: bool = False) -> str:
    nvcc = _nvcc()
    if nvcc is None:
        raise RuntimeError("nvcc not found")
    os.makedirs(LIB_DIR, exist_ok=True)
    os.makedirs(OBJ_DIR, exist_ok=True)
    srcs = sorted(glob.glob(os.path.join(REPO, "csrc", "cuda", "*.cu")))
    hdrs = (glob.glob(os.path.join(REPO, "csrc", "cuda", "*.cuh"))
            + glob.glob(os.path.join(REPO, "csrc", "cuda", "common", "*.cuh"))
            + glob.glob(os.path.join(REPO, "include", "*.h")))
    flags = ["-ccbin", "/usr/bin/g++"] if os.path.exists("/usr/bin/g++") else []
    flags += ["-O3", "-std=c++17", "-lineinfo", "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC",
             "-Xcompiler", "-fvisibility=default"] + NVCC_ARCH
    if ptxas_verbose:
        flags += ["-Xptxas", "-v"]
    jobs = []
    objs = []
    for s in srcs:
        o = os.path.join(OBJ_DIR, os.path.basename(s) + ".o")
        objs.append(o)
        if force or _newer(o, [s] + hdrs):
            jobs.append([nvcc] + flags + ["-c", s, "-o", o])
    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        list(ex.map(lambda c: _run(c, verbose), jobs))
    out = cuda_lib_path()
    if force or jobs or _newer(out, objs):
        _run([nvcc, "-shared", "-o", out] + objs + NVCC_ARCH + ["-cudart", "static"], verbose)
    return out


def _build_cpp(name: str, subdir: str, extra: List[str], verbose: bool, force: bool) -> str:
    os.makedirs(LIB_DIR, exist_ok=True)
    srcs = sorted(glob.glob(os.path.join(REPO, "csrc", subdir, "*.cpp")))
    hdrs = glob.glob(os.path.join(REPO, "include", "*.h")) + glob.glob(os.path.join(REPO, "csrc", subdir, "*.h"))
    out = os.path.join(LIB_DIR, name)
    if force or _newer(out, srcs + hdrs):
        # the image's $CXX (/opt/gcc) ships without libgomp.spec; prefer the distro compiler
        cxx = next((c for c in ("/usr/bin/g++", shutil.which("g++"), os.environ.get("CXX"))
                    if c and os.path.exists(c)), "g++")
        _run([cxx, "-O3", "-std=c++17", "-fPIC", "-shared", "-Wall", f"-I{os.path.join(REPO, 'include')}"]
             + extra + srcs + ["-o", out], verbose)
    return out


def build_cpu(verbose: bool = False, force: bool = False) -> str:
    return _build_cpp("libcrossscale_cpu.so", "cpu", ["-fopenmp", "-mavx2", "-mfma"], verbose, force)


def build_rt(verbose: bool = False, force: bool = False) -> str:
    return _build_cpp("libcrossscale_rt.so", "runtime", ["-pthread"], verbose, force)


def build_all(verbose: bool = False, force: bool = False) -> dict:
    out = {"cpu": build_cpu(verbose, force), "rt": build_rt(verbose, force)}
    out["cuda"] = build_cuda(verbose, force)
    return out


if __name__ == "__main__":
    res = build_all(verbose="-q" not in sys.argv, force="--force" in sys.argv)
    for k, v in res.items():
        print(f"{k}: {v}")
