"""bench.py / __graft_entry__ contracts that can be checked without a GPU."""
import json
import os
import subprocess
import sys

import pytest
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args):
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    return subprocess.run([sys.executable, os.path.join(REPO, "bench.py")] + args, cwd=REPO, env=env, capture_output=True,
                          text=True, timeout=300)


@pytest.mark.skipif(torch.cuda.is_available(), reason="CPU-only contract check")
def test_bench_without_gpu_prints_one_json_line_and_exits_zero():
    r = _run(["--steps", "2", "--warmup", "3"])
    assert r.returncode == 0, r.stderr[-2000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
    out = json.loads(line)
    assert "metric" in out and out.get("n_gpus") == 0 and "error" in out


@pytest.mark.skipif(torch.cuda.is_available(), reason="CPU-only contract check")
def test_reference_arm_reports_unavailable_without_gpu():
    r = _run(["--impl", "reference", "--steps", "2", "--warmup", "3"])
    assert r.returncode == 0, r.stderr[-2000:]
    out = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert out["impl"] == "reference" and "unavailable" in out


def test_graft_entry_has_build_and_smoke():
    sys.path.insert(0, REPO)
    import __graft_entry__ as g
    assert callable(g.build) and callable(g.smoke)
    from crossscale.utils import build as b
    for p in (b.cuda_lib_path(), b.cpu_lib_path(), b.rt_lib_path()):
        assert p.startswith(os.path.join(REPO, "lib"))          # built in-tree so the GPU box and the driver see it


def test_header_declares_every_exported_cuda_symbol():
    """include/crossscale.h is the C ABI: every extern "C" cs_* function of api.cu / noalloc must be declared there."""
    import re
    hdr = open(os.path.join(REPO, "include", "crossscale.h")).read()
    declared = set(re.findall(r"\b(cs_[a-z0-9_]+)\s*\(", hdr))
    defined = set()
    for f in ("api.cu", "tiny_ecg_noalloc.cu", "tiny_ecg_round.cu"):
        src = open(os.path.join(REPO, "csrc", "cuda", f)).read()
        defined |= set(re.findall(r"^(?:extern \"C\" )?(?:int|size_t|void\*?|const char\*)\s+(cs_[a-z0-9_]+)\s*\(", src, flags=re.M))
    missing = sorted(d for d in defined - declared if not d.startswith("cs_debug"))
    assert not missing, f"declare in include/crossscale.h: {missing}"
