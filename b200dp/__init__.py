"""Importable alias for the framework package.

The source tree lives in the directory named after the upstream project
(``pg---diploma-project---distributed-ai-model-training-using-mpi-and-accelerated-gpu-_b200``),
which is not a valid Python identifier.  This module re-points ``b200dp.__path__`` at
that directory and executes its ``__init__``, so ``import b200dp.models`` etc. resolve to the
real sources and there is exactly one copy of every module.
"""
import os as _os

_ROOT = _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))
PACKAGE_DIRNAME = "pg---diploma-project---distributed-ai-model-training-using-mpi-and-accelerated-gpu-_b200"
_PKG_DIR = _os.path.join(_ROOT, PACKAGE_DIRNAME)

__path__ = [_PKG_DIR]
__file__ = _os.path.join(_PKG_DIR, "__init__.py")
with open(__file__, "r") as _f:
    exec(compile(_f.read(), __file__, "exec"), globals())
del _f
