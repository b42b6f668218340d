"""Multi-rank check of the NATIVE ENGINE (run under b200dp.launch, one rank per GPU).

mode "train": 30 real training steps (preprocess, forward, backward, bucketed all-reduce + Adam inside the step
graph), each rank on its own data; after every step the parameter hash must be identical on all ranks, and the
global mean loss per step must match a run of the same steps with `--comm nccl` (NCCL all-reduce + flat Adam).
Gradients are accumulated with fp32 atomics, so no two runs are bit-equal and a randomly initialised network
amplifies the difference step by step: the first step (same weights) must agree to 2e-3, and the whole curve to within three
times the distance between two NCCL runs of the same steps (the measured noise floor).
mode "resume": the default multi-GPU optimiser shards Adam's moments (each rank updates the slices it owns); the saved
optimiser state must nevertheless be the FULL moments, and a run resumed from it must continue like the uninterrupted one.
mode "fault": rank 1 stops itself (SIGSTOP) after a few steps; every other rank must leave with CommError -- never
with silently updated weights -- within the barrier's spin limit.
"""
import os
import signal
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from b200dp.config import TrainConfig  # noqa: E402
from b200dp.data.datasets import SyntheticCIFAR10  # noqa: E402
from b200dp.data.loader import Batch  # noqa: E402
from b200dp.data.transforms import sample_aug_params  # noqa: E402
from b200dp.parallel.bootstrap import cleanup, setup_distributed  # noqa: E402
from b200dp.trainer import make_engine  # noqa: E402


def batches(rank: int, steps: int, B: int, dev):
    ds = SyntheticCIFAR10(steps * B, True, seed=100 + rank)
    g = torch.Generator().manual_seed(7 + rank)
    out = []
    for i in range(steps):
        sl = slice(i * B, (i + 1) * B)
        out.append(Batch(ds.images[sl].to(dev), ds.labels[sl].to(dev), sample_aug_params(B, generator=g).to(dev), B))
    return out


def run(cfg, ctx, data, check_hash: bool):
    eng = make_engine(cfg, ctx)
    losses = []
    for b in data:
        eng.train_step(b, 1e-3)
        ls, _, cnt = eng.read_metrics()
        g = eng.reduce([ls, cnt], "sum")
        losses.append(g[0] / g[1])
        if check_hash:
            h = eng.param_hash()
            hi, neg_lo = eng.reduce([h, -h], "max")
            assert hi == -neg_lo, f"replicas diverged at step {len(losses)}: [{-neg_lo!r}, {hi!r}]"
    info = (eng.comm_algo, eng.fused.arena.kind, eng.in_graph)
    eng.close()
    return losses, info


def main() -> int:
    mode = sys.argv[1] if len(sys.argv) > 1 else "train"
    ctx = setup_distributed()
    dev, r = ctx.device, ctx.rank
    B, steps = 16, 30
    data = batches(r, steps, B, dev)
    base = dict(mode="ddp", engine="native", batch_size=B, dataset="synthetic", dropout=0.0, seed=3)
    if mode == "fault":
        os.environ["B200DP_COMM_SPIN_LIMIT"] = str(20_000_000)
        eng = make_engine(TrainConfig(**base), ctx)
        for i, b in enumerate(data * 50):
            if i == 5 and r == 1:
                torch.cuda.synchronize()
                os.kill(os.getpid(), signal.SIGSTOP)      # a stalled rank: its host stops launching steps
            eng.train_step(b, 1e-3)
            if i % 8 == 7:
                eng.read_metrics()
        print("FAULT NOT DETECTED", flush=True)
        return 0
    if mode == "resume":
        cfg = TrainConfig(comm="fused", comm_algo="two_shot", **base)
        a = make_engine(cfg, ctx)
        for b in data[:6]:
            a.train_step(b, 1e-3)
        sd_model, sd_opt = a.state_dict(), a.optimizer_state()         # collective: moments are summed over ranks
        local_m = a.fused.m.clone()
        for b in data[6:10]:
            a.train_step(b, 1e-3)
        torch.cuda.synchronize()
        want = a.core.params.clone()
        # this rank only ever touched its own slices of m; the saved state covers everything that has a gradient
        own = (local_m != 0).float().mean().item()
        full = (sd_opt["m"] != 0).float().mean().item()
        assert full > 1.5 * own and sd_opt["step"] == 6 and sd_opt["algo"] == "two_shot" and sd_opt["world"] == ctx.world_size, (own, full)
        a.close()

        def resumed(opt_state):
            e = make_engine(cfg, ctx)
            e.load_state_dict(sd_model)
            e.broadcast_parameters()
            e.load_optimizer_state(opt_state)
            for b in data[6:10]:
                e.train_step(b, 1e-3)
            torch.cuda.synchronize()
            return e, ((e.core.params - want).norm() / want.norm()).item()

        # gradients are accumulated with fp32 atomics and Adam normalises them, so two identical continuations differ:
        # the second continuation from the same state is the noise floor.  A state that has lost the peers' moments
        # (what a rank holds locally -- the round-1 bug) must land clearly further away.
        c, noise = resumed(sd_opt)
        c.close()
        lossy = dict(sd_opt, m=local_m.cpu(), v=torch.zeros_like(sd_opt["v"]))
        d, bad = resumed(lossy)
        d.close()
        b2, rel = resumed(sd_opt)
        assert rel < max(3 * noise, 1e-3) and bad > 3 * max(rel, noise), (rel, noise, bad)
        h = b2.param_hash()
        hi, neg_lo = b2.reduce([h, -h], "max")
        assert hi == -neg_lo
        if r == 0:
            print(f"resume: moments nonzero {own:.2f} locally, {full:.2f} in the saved state; distance to the uninterrupted run "
                  f"{rel:.1e} (second continuation {noise:.1e}, with the peers' moments lost {bad:.1e})")
            print("RESUME CHECK OK", flush=True)
        b2.close()
        cleanup(ctx)
        return 0
    algos = ["auto", "one_shot", "two_shot"]
    rel = lambda x, y: [abs(a - b) / max(abs(b), 1e-6) for a, b in zip(x, y)]
    ref, _ = run(TrainConfig(comm="nccl", **base), ctx, data, False)
    ref2, _ = run(TrainConfig(comm="nccl", **base), ctx, data, False)
    noise = max(rel(ref2, ref))
    if r == 0:
        print(f"noise floor (nccl vs nccl): {noise:.2e}", flush=True)
    for algo in algos:
        got, info = run(TrainConfig(comm="fused", comm_algo=algo, **base), ctx, data, True)
        d = rel(got, ref)
        worst = max(d)
        if r == 0:
            print(f"algo={algo} -> {info}: loss {got[0]:.4f} -> {got[-1]:.4f}, first steps {d[0]:.1e} {d[1]:.1e} {d[2]:.1e}, "
                  f"max rel diff vs nccl {worst:.2e}", flush=True)
        assert d[0] < 2e-3 and max(d[:3]) < max(3 * noise, 1e-2), (algo, d[:3])
        assert worst < max(3 * noise, 5e-3), (algo, worst, noise, got, ref)
        assert info[2], "the fused optimiser must run inside the step graph"
    if r == 0:
        print("ENGINE CHECK OK", flush=True)
    cleanup(ctx)
    return 0


if __name__ == "__main__":
    sys.exit(main())
