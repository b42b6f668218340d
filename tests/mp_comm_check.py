"""Multi-rank check of the peer-memory kernels (run under b200dp.launch / torchrun, one rank per GPU).

Asserts: fused all-reduce + Adam == NCCL all-reduce + flat Adam to tolerance, parameters are
BIT-IDENTICAL across ranks after several steps (one-shot and two-shot), plain all-reduce ==
dist.all_reduce, broadcast and barrier work, and prints achieved bus bandwidth for a size sweep.
"""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from b200dp.optim import AdamHyper, FlatAdam  # noqa: E402
from b200dp.parallel.bootstrap import cleanup, setup_distributed  # noqa: E402
from b200dp.parallel.comm import FusedAllReduceAdam  # noqa: E402


def same_on_all_ranks(t: torch.Tensor) -> bool:
    ref = t.clone()
    dist.broadcast(ref, src=0)
    ok = torch.tensor([1 if torch.equal(ref, t) else 0], device=t.device)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    return bool(ok.item())


def main() -> int:
    ctx = setup_distributed()
    dev, W, r = ctx.device, ctx.world_size, ctx.rank
    torch.manual_seed(1234)
    n = 2_242_624
    hyper = AdamHyper(lr=1e-3)
    p0 = torch.randn(n, device=dev)
    probe = FusedAllReduceAdam(ctx, 1024, hyper)
    kind = probe.arena.kind
    probe.close()
    if r == 0:
        print("arena:", kind)
    algos = ("one_shot", "two_shot", "nvls") if kind == "vmm+nvls" else ("one_shot", "two_shot")
    # bucketed = every step as three sub-range launches (the in-graph form).  Ownership of the two-shot slices -- and
    # with it which rank holds the moments of an element -- depends on the ranges, so a run uses one form throughout.
    for algo, gdtype, bucketed in [(a, g, b) for a in algos for g in (torch.float32, torch.bfloat16) for b in (False, True)]:
        if True:
            if (algo == "nvls" or bucketed) and gdtype != torch.float32:
                continue
            fused = FusedAllReduceAdam(ctx, n, hyper, params=p0.clone(), work_bf16=algo != "nvls", grad_dtype=gdtype, algo=algo)
            ref_p, ref = p0.clone(), FlatAdam(n, dev, hyper)
            gen = torch.Generator(device=dev).manual_seed(100 + r)
            for step in range(4):
                g = torch.randn(n, device=dev, generator=gen)
                fused.grads.copy_(g)
                gsum = fused.grads.float().clone()
                dist.all_reduce(gsum)
                if bucketed:       # three sub-range launches with a small fixed grid, the step counter bumps after the last
                    cuts = [0, 1_000_000, 2_000_000, n]
                    for i in (2, 1, 0):
                        a = fused.args
                        a.range_lo, a.range_hi, a.bump_step, a.exact_blocks = cuts[i], cuts[i + 1], int(i == 0), 16
                        fused.step(1e-3)
                    fused.args.range_lo = fused.args.range_hi = fused.args.exact_blocks = 0
                    fused.args.bump_step = 1
                else:
                    fused.step(1e-3)
                ref.step_torch(ref_p, gsum, 1e-3, grad_scale=1.0 / W)
                torch.cuda.synchronize()
                fused.check_error()
                assert torch.allclose(fused.params, ref_p, rtol=3e-5, atol=3e-6), (algo, gdtype, bucketed, step)
                assert torch.count_nonzero(fused.grads) == 0
                assert same_on_all_ranks(fused.params), f"replicas diverged ({algo}, {gdtype}, {bucketed}, step {step})"
            if fused.work is not None:
                assert torch.equal(fused.work, fused.params.to(torch.bfloat16))
            # plain all-reduce
            fused.grads.copy_(torch.full((n,), float(r + 1), device=dev))
            avg = fused.all_reduce_mean()
            torch.cuda.synchronize()
            assert torch.allclose(avg, torch.full((n,), (W + 1) / 2.0, device=dev))
            fused.barrier()
            fused.close()
    # broadcast of a parameter-sized vector from rank 0 into every rank's arena
    fused = FusedAllReduceAdam(ctx, n, hyper, params=p0.clone(), algo="two_shot", extra_numel=4096)
    fused.params.fill_(7.0 if r == 0 else float(r))
    fused.extra.fill_(3.0 if r == 0 else -1.0)
    torch.cuda.synchronize(); dist.barrier()
    fused.broadcast_params(0)
    fused.broadcast_extra(0)
    torch.cuda.synchronize()
    assert torch.all(fused.params == 7.0) and torch.all(fused.extra == 3.0)
    # small double reductions on the peer slot tables (K-C6 / K-C7, timers)
    got = fused.reduce_f64(torch.tensor([float(r + 1), 10.0 * r, 0.5]), "sum").cpu()
    assert torch.equal(got, torch.tensor([W * (W + 1) / 2.0, 10.0 * W * (W - 1) / 2, 0.5 * W], dtype=torch.float64))
    got = fused.reduce_f64(torch.tensor([float(r), -float(r)]), "max").cpu()
    assert torch.equal(got, torch.tensor([W - 1.0, 0.0], dtype=torch.float64))
    fused.check_error()
    fused.close()
    # bus-bandwidth sweep of the fused kernel vs NCCL all-reduce (device-timed, max over ranks)
    if r == 0:
        print("bytes,algo,fused_us,nccl_us")
    sweep = [(nb, al) for nb in (64 << 10, 1 << 20, 8970496, 64 << 20, 256 << 20) for al in algos]
    if os.environ.get("B200DP_COMM_SWEEP", "1") == "0":
        sweep = []
    for nbytes, algo in sweep:
        m = nbytes // 4
        f = FusedAllReduceAdam(ctx, m, hyper, algo=algo)
        t = torch.randn(m, device=dev)

        def timed(fn, iters=20):
            for _ in range(5):
                fn()
            torch.cuda.synchronize(); dist.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(iters):
                fn()
            b.record(); torch.cuda.synchronize()
            v = torch.tensor([a.elapsed_time(b) / iters * 1e3], device=dev)
            dist.all_reduce(v, op=dist.ReduceOp.MAX)
            return float(v.item())
        fu = timed(lambda: f.step(1e-3))
        nc = timed(lambda: dist.all_reduce(t))
        if r == 0:
            print(f"{nbytes},{algo},{fu:.1f},{nc:.1f}")
        f.close()
    dist.barrier()
    if r == 0:
        print("COMM CHECK OK")
    cleanup(ctx)
    return 0


if __name__ == "__main__":
    sys.exit(main())
