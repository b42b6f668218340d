"""Host-only model of the fused all-reduce + Adam kernel (csrc/comm_adam.cu) -- SURVEY §4 "(a)".

Ranks are threads, "peer memory" is a list of numpy arrays every thread can index, the flag pads are
plain integer arrays.  The model follows the kernel launch for launch and block for block: the same
flag protocol (block b of rank r writes ``target`` into slot ``[b][r]`` of every peer's pad and
waits until its own pad shows ``target`` from everybody), the same range / slice arithmetic in units
of 4 elements, the same per-block zeroing pattern of the two-shot forms, the same ticket that lets
the LAST block of the step's last launch advance the step count, and the same failure rule (a
barrier that times out returns without touching anything and raises the sticky error flag).  It
exists so that this bookkeeping can be tested on a machine without GPUs; the arithmetic inside an
element is ``optim.adam_scalars`` + the textbook update, as in the kernel.

Not modelled: memory ordering (the GIL serialises everything), NVLS (its ownership arithmetic is the
two-shot one), bf16 gradients.
"""
from __future__ import annotations

import threading
import time
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

from ..optim import AdamHyper, adam_scalars

THREADS = 512            # kThreads of the kernel: elements are walked in units of one float4 per thread


@dataclass
class RankState:
    """What one rank owns: its arena (flags, gradients, weights) and its local optimiser state."""
    numel: int
    max_blocks: int
    world: int
    grads: np.ndarray = field(init=False)
    params: np.ndarray = field(init=False)
    m: np.ndarray = field(init=False)
    v: np.ndarray = field(init=False)
    flags: np.ndarray = field(init=False)          # [max_blocks][world]
    epoch: np.ndarray = field(init=False)          # [max_blocks]
    step: int = 0
    ticket: int = 0
    error: int = 0
    abort: int = 0                                 # set by ANY rank whose barrier timed out

    def __post_init__(self) -> None:
        assert self.numel % 4 == 0
        self.grads = np.zeros(self.numel, np.float32)
        self.params = np.zeros(self.numel, np.float32)
        self.m = np.zeros(self.numel, np.float32)
        self.v = np.zeros(self.numel, np.float32)
        self.flags = np.zeros((self.max_blocks, self.world), np.int64)
        self.epoch = np.zeros(self.max_blocks, np.int64)


def grid_for(n4_range: int, max_blocks: int = 148) -> int:
    """grid_for() of the kernel: sized by the whole range, one float4 per thread and eight trips."""
    blocks = (n4_range + THREADS * 8 - 1) // (THREADS * 8)
    return max(1, min(blocks, max_blocks))


class CommModel:
    def __init__(self, world: int, numel: int, hyper: Optional[AdamHyper] = None, algo: str = "one_shot",
                 max_blocks: int = 148, spin_timeout_s: float = 5.0) -> None:
        self.W, self.numel, self.h = world, numel, hyper or AdamHyper()
        self.algo = {"one_shot": 0, "two_shot": 1}[algo]
        self.ranks: List[RankState] = [RankState(numel, max_blocks, world) for _ in range(world)]
        self.max_blocks = max_blocks
        self.timeout = spin_timeout_s
        self._lock = threading.Lock()

    # ---- the device-side barrier of block b on rank r -----------------------------------------------------------
    def _barrier(self, r: int, b: int, target: int) -> bool:
        me = self.ranks[r]
        for peer in self.ranks:                        # thread t < W of the block: one remote store per peer
            peer.flags[b, r] = target
        deadline = time.monotonic() + self.timeout
        while True:
            if all(me.flags[b, q] - target >= 0 for q in range(self.W)):
                return True
            if time.monotonic() > deadline:            # spin limit: report, do not hang, do not touch anything
                me.error = 1 + next(q for q in range(self.W) if me.flags[b, q] - target < 0)
                for peer in self.ranks:                # and tell everybody: a late peer must not update anything either
                    peer.abort = me.error
                return False
            time.sleep(0)

    # ---- one block of one launch ----------------------------------------------------------------------------------
    def _block(self, r: int, b: int, grid: int, lo4: int, hi4: int, lr: float, bump: bool, skip_barriers: bool,
               zero_grads: bool) -> None:
        me, W = self.ranks[r], self.W
        def failed() -> bool:                          # own sticky flag, or a peer gave up on a barrier
            if not me.error and me.abort:
                me.error = me.abort
            return bool(me.error)

        if failed():
            return
        epoch = int(me.epoch[b])
        barriers = W > 1 and not skip_barriers
        if barriers and (not self._barrier(r, b, epoch + 1) or failed()):
            return
        step = me.step                                 # read before anybody can bump it (the ticket guarantees that)
        step_size, isb2 = adam_scalars(step + 1, lr, self.h.beta1, self.h.beta2)
        n4 = hi4 - lo4
        per = n4
        lo, hi = lo4, hi4
        if self.algo >= 1:
            per = (n4 + W - 1) // W
            lo = lo4 + per * r
            hi = min(lo + per, hi4)
        stride = grid * THREADS
        # elements of this block: i = lo + b*THREADS + t + k*stride  (t < THREADS)
        idx4 = self._block_indices(lo, hi, b, stride)
        if idx4.size:
            e = (idx4[:, None] * 4 + np.arange(4)[None, :]).ravel()
            acc = np.zeros(e.size, np.float32)
            for q in range(W):                         # fixed rank order 0..W-1: bit-identical sums everywhere
                acc = acc + self.ranks[q].grads[e]
            g = acc * np.float32(1.0 / W)
            if self.h.weight_decay:
                g = g + np.float32(self.h.weight_decay) * me.params[e]
            me.m[e] = np.float32(self.h.beta1) * me.m[e] + np.float32(1 - self.h.beta1) * g
            me.v[e] = np.float32(self.h.beta2) * me.v[e] + np.float32(1 - self.h.beta2) * g * g
            p = me.params[e] - np.float32(step_size) * me.m[e] / (np.sqrt(me.v[e]) * np.float32(isb2) + np.float32(self.h.eps))
            if self.algo == 0:
                me.params[e] = p
            else:
                for q in range(W):                     # all-gather: push the updated slice into every replica
                    self.ranks[q].params[e] = p
        second = barriers and (zero_grads or self.algo >= 1)
        if second and not self._barrier(r, b, epoch + 2):
            return
        if zero_grads:
            if self.algo == 0 or W == 1:
                z = self._block_indices(lo4, hi4, b, stride)
            else:                                      # the part of every slice q that block b of rank q has read
                z = np.concatenate([self._block_indices(lo4 + per * q, min(lo4 + per * (q + 1), hi4), b, stride)
                                    for q in range(W)])
            if z.size:
                me.grads[(z[:, None] * 4 + np.arange(4)[None, :]).ravel()] = 0
        if barriers:
            me.epoch[b] = epoch + 2
        if bump:
            with self._lock:
                me.ticket += 1
                if me.ticket == grid:                  # the last block of the step's last launch
                    me.ticket = 0
                    me.step = step + 1

    @staticmethod
    def _block_indices(lo: int, hi: int, b: int, stride: int) -> np.ndarray:
        out = []
        first = lo + b * THREADS
        while first < hi:
            out.append(np.arange(first, min(first + THREADS, hi)))
            first += stride
        return np.concatenate(out) if out else np.zeros(0, np.int64)

    # ---- a launch on every rank (one thread per rank; blocks of a rank run in a shuffled order) --------------------
    def launch(self, lr: float, range_lo: int = 0, range_hi: int = 0, bump: bool = True, skip_barriers: bool = False,
               zero_grads: bool = True, absent: tuple = (), seed: int = 0) -> None:
        lo4, hi4 = (range_lo // 4, range_hi // 4) if range_hi > 0 else (0, self.numel // 4)
        grid = grid_for(hi4 - lo4, self.max_blocks)

        def rank_main(r: int) -> None:
            order = np.random.RandomState(seed * 131 + r).permutation(grid)     # block scheduling order is not defined
            ths = [threading.Thread(target=self._block, args=(r, int(b), grid, lo4, hi4, lr, bump, skip_barriers, zero_grads))
                   for b in order]
            for t in ths:
                t.start()
            for t in ths:
                t.join()

        ths = [threading.Thread(target=rank_main, args=(r,)) for r in range(self.W) if r not in absent]
        for t in ths:
            t.start()
        for t in ths:
            t.join()
