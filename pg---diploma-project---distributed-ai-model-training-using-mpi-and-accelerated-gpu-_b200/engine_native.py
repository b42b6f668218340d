"""Native engine: the whole train / eval step runs in hand-written sm_100a kernels driven by the
C++ executor (csrc/engine.cpp).  Placeholder until the executor lands; ``available()`` gates it."""
from __future__ import annotations


def available() -> bool:
    return False
