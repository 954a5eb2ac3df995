"""Importable alias for the framework package.

The framework lives in the directory
``crossscale-ecg-a-modular-hpc-pipeline-from-locality-optimization-to-mpi-gpu-overlap_b200/``
whose name (hyphens) is not a Python identifier.  This alias package points its
``__path__`` at that directory and executes its ``__init__.py`` so that
``import crossscale.models`` / ``python -m crossscale.cli.shard_prep`` resolve there.
There is no second copy of any module.
"""
import os as _os

_REAL_DIRNAME = "crossscale-ecg-a-modular-hpc-pipeline-from-locality-optimization-to-mpi-gpu-overlap_b200"
_root = _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))
_real = _os.path.join(_root, _REAL_DIRNAME)
if not _os.path.isdir(_real):  # pragma: no cover
    raise ImportError(f"framework directory missing: {_real}")
__path__.insert(0, _real)
REPO_ROOT = _root
PACKAGE_DIR = _real
_init = _os.path.join(_real, "__init__.py")
with open(_init, "r") as _f:
    exec(compile(_f.read(), _init, "exec"))
del _f, _init
