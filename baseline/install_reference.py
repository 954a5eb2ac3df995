"""Put the UNMODIFIED reference under ``baseline/_ref`` (git-ignored, travels with gpurun).

``pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target baseline/_ref
/root/reference`` fails: the reference has neither ``setup.py`` nor ``pyproject.toml`` — it is a
tree of scripts, not a package (outcome recorded in DESIGN.md).  Its "installation" is therefore a
byte-for-byte copy of the source tree, verified by SHA-256 against ``/root/reference``.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import sys

SRC = "/root/reference"
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def _sha(p):
    h = hashlib.sha256()
    with open(p, "rb") as f:
        for blk in iter(lambda: f.read(1 << 20), b""):
            h.update(blk)
    return h.hexdigest()


def install() -> int:
    if not os.path.isdir(SRC):
        print(f"reference not mounted at {SRC}")
        return 1
    if os.path.isdir(DST):
        shutil.rmtree(DST)
    shutil.copytree(SRC, DST)
    bad = 0
    for root, _, files in os.walk(SRC):
        for fn in files:
            s = os.path.join(root, fn)
            d = os.path.join(DST, os.path.relpath(s, SRC))
            if _sha(s) != _sha(d):
                print("MISMATCH", d)
                bad += 1
    print(f"copied reference -> {DST} ({'OK, all files identical' if not bad else f'{bad} mismatches'})")
    return bad


if __name__ == "__main__":
    sys.exit(install())
