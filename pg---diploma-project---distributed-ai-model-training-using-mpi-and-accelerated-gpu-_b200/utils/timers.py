"""Timing.

The reference brackets epochs with ``time.time()`` (cifar10_mpi_mobilenet_224.py:161-164, 227).
Device work is asynchronous, so here phases are bracketed with CUDA events on the launching
stream, read after a synchronize, and multi-rank numbers are reduced with MAX over ranks (the
slowest rank defines the step).  On CPU the same interface falls back to ``perf_counter``.
"""
from __future__ import annotations

import time
from collections import defaultdict
from typing import Dict, List, Optional

import torch


class PhaseTimer:
    """``with timer.phase("fwd"):`` accumulates per-phase device milliseconds."""

    def __init__(self, enabled: bool = True, cuda: Optional[bool] = None) -> None:
        self.enabled = enabled
        self.cuda = torch.cuda.is_available() if cuda is None else cuda
        self._pending: Dict[str, List] = defaultdict(list)
        self._ms: Dict[str, float] = defaultdict(float)
        self._count: Dict[str, int] = defaultdict(int)

    class _Phase:
        def __init__(self, t: "PhaseTimer", name: str) -> None:
            self.t, self.name = t, name

        def __enter__(self):
            if not self.t.enabled:
                return self
            if self.t.cuda:
                self.a = torch.cuda.Event(enable_timing=True)
                self.b = torch.cuda.Event(enable_timing=True)
                self.a.record()
            else:
                self.t0 = time.perf_counter()
            return self

        def __exit__(self, *exc):
            if not self.t.enabled:
                return False
            if self.t.cuda:
                self.b.record()
                self.t._pending[self.name].append((self.a, self.b))
            else:
                self.t._ms[self.name] += (time.perf_counter() - self.t0) * 1e3
            self.t._count[self.name] += 1
            return False

    def phase(self, name: str) -> "PhaseTimer._Phase":
        return PhaseTimer._Phase(self, name)

    def collect(self) -> Dict[str, float]:
        """Synchronize and fold pending event pairs into milliseconds per phase."""
        if self.cuda and self._pending:
            torch.cuda.synchronize()
            for name, pairs in self._pending.items():
                for a, b in pairs:
                    self._ms[name] += a.elapsed_time(b)
            self._pending.clear()
        return dict(self._ms)

    def reset(self) -> None:
        self._pending.clear()
        self._ms.clear()
        self._count.clear()


def max_over_ranks(value: float, device: torch.device) -> float:
    """MAX-reduce a scalar across the default process group (no-op when not distributed)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return float(value)
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(values, device: torch.device):
    """SUM-reduce float64 values across ranks -- the reference's ``[loss_sum, count]`` reduction
    (K-C6 / K-C7, cifar10_mpi_mobilenet_224.py:188-195, 216-223)."""
    import torch.distributed as dist
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized():
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return [float(x) for x in t.tolist()]
