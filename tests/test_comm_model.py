"""The fused all-reduce + Adam protocol on a host-only model (ranks = threads, peer memory = shared arrays): range /
slice / zeroing bookkeeping, the step-count ticket, bit-identical replicas, the failure rule -- SURVEY §4 "(a)"."""
import numpy as np
import pytest
import torch

from b200dp.optim import AdamHyper, FlatAdam
from b200dp.parallel.comm_model import CommModel, grid_for


def _setup(W, n, algo, seed=0):
    rng = np.random.RandomState(seed)
    m = CommModel(W, n, AdamHyper(lr=1e-3, weight_decay=0.01), algo=algo, max_blocks=6, spin_timeout_s=20.0)
    p0 = rng.randn(n).astype(np.float32)
    for r in m.ranks:
        r.params[:] = p0
    return m, p0, rng


def _reference(p0, grads_per_step, W, lrs):
    ref_p = torch.from_numpy(p0.copy())
    ref = FlatAdam(p0.size, "cpu", AdamHyper(lr=1e-3, weight_decay=0.01))
    for gs, lr in zip(grads_per_step, lrs):
        ref.step_torch(ref_p, torch.from_numpy(np.sum(gs, axis=0, dtype=np.float32)), lr, grad_scale=1.0 / W)
    return ref_p.numpy(), ref


@pytest.mark.parametrize("W,algo", [(1, "one_shot"), (2, "one_shot"), (4, "one_shot"), (2, "two_shot"), (4, "two_shot"), (3, "two_shot")])
@pytest.mark.parametrize("bucketed", [False, True])
def test_model_matches_flat_adam_and_replicas_are_identical(W, algo, bucketed):
    n = 4 * 5000 + 4 * 37                     # not a multiple of anything convenient
    m, p0, rng = _setup(W, n, algo)
    lrs = [1e-3, 1e-3, 1e-4]
    hist = []
    for s, lr in enumerate(lrs):
        gs = [rng.randn(n).astype(np.float32) for _ in range(W)]
        hist.append(gs)
        for r, g in zip(m.ranks, gs):
            r.grads[:] = g
        if bucketed:                          # three sub-range launches, the last one bumps the step count
            cuts = [0, 4 * 1111, 4 * 3333, n]
            for i in (2, 1, 0):
                m.launch(lr, cuts[i], cuts[i + 1], bump=(i == 0), seed=s * 7 + i)
        else:
            m.launch(lr, seed=s)
        for r in m.ranks:
            assert r.error == 0 and r.step == s + 1 and r.ticket == 0
            assert not r.grads.any()                                   # every element zeroed exactly by its owner-block
            assert np.array_equal(r.params, m.ranks[0].params)         # replicas bit-identical
    want, ref = _reference(p0, hist, W, lrs)
    assert np.allclose(m.ranks[0].params, want, rtol=3e-5, atol=3e-6)
    # moments: every rank has the full vectors (one-shot) or exactly its slices, and the slices tile the vector
    msum = np.sum([r.m for r in m.ranks], axis=0)
    if algo == "one_shot":
        assert np.allclose(m.ranks[0].m, ref.m.numpy(), rtol=1e-5, atol=1e-7)
    elif not bucketed:
        assert np.allclose(msum, ref.m.numpy(), rtol=1e-5, atol=1e-7)   # what state_dict() saves: the sum over ranks
        owned = np.sum([(r.m != 0) for r in m.ranks], axis=0)
        assert owned.max() == 1                                         # one owner per element


def test_a_missing_rank_is_an_error_and_nothing_is_updated():
    m, p0, rng = _setup(2, 4 * 600, "two_shot")
    m.timeout = 0.5
    for r in m.ranks:
        r.grads[:] = 1.0
    m.launch(1e-3, absent=(1,))                                        # rank 1 never launches (dead or stalled peer)
    r0 = m.ranks[0]
    assert r0.error == 2                                               # 1 + the rank that did not arrive
    assert np.array_equal(r0.params, p0) and not r0.m.any() and r0.step == 0 and r0.grads.all()
    # rank 1 arrives late: rank 0's flag from the failed launch is already there, so its first barrier would pass --
    # the abort word rank 0 left in rank 1's arena is what stops it from updating (and storing into rank 0's weights)
    m.launch(1e-3)
    assert np.array_equal(r0.params, p0) and m.ranks[1].error == 2 and np.array_equal(m.ranks[1].params, p0)
    assert not m.ranks[1].m.any() and m.ranks[1].step == 0


def test_skip_barriers_data_launch_then_final_launch():
    """The in-graph bucket form: an early bucket as a data launch without barriers or zeroing (a one-CTA barrier launch
    did the waiting), then the last bucket with both barriers."""
    W, n = 2, 4 * 3000
    m, p0, rng = _setup(W, n, "two_shot")
    gs = [rng.randn(n).astype(np.float32) for _ in range(W)]
    for r, g in zip(m.ranks, gs):
        r.grads[:] = g
    cut = 4 * 1000
    m.launch(1e-3, cut, n, bump=False, skip_barriers=True, zero_grads=False)
    assert all(r.step == 0 and r.grads[cut:].any() for r in m.ranks)
    m.launch(1e-3, 0, cut, bump=True)
    for r in m.ranks:
        r.grads[cut:] = 0                                              # the kernel's extra zero range
    want, _ = _reference(p0, [gs], W, [1e-3])
    assert all(r.step == 1 for r in m.ranks)
    assert np.allclose(m.ranks[0].params, want, rtol=3e-5, atol=3e-6) and np.array_equal(m.ranks[0].params, m.ranks[1].params)


def test_grid_sizing():
    assert grid_for(1) == 1 and grid_for(512 * 8) == 1 and grid_for(512 * 8 + 1) == 2 and grid_for(10 ** 9) == 148
