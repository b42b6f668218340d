"""GPU numerics: each native kernel against a plain PyTorch fp32 reference of the same op."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def ops():
    from b200dp import ops as o
    from b200dp.ops import _lib
    if _lib.load() is None:
        pytest.fail(f"native library must load on a GPU box: {_lib.load_error()}")
    return o


def _levels(a, b):
    """difference in 8-bit levels between two normalised images (NCHW fp32)."""
    std = torch.tensor([0.229, 0.224, 0.225], device=a.device).view(1, 3, 1, 1) * 255
    return (a - b).abs() * std


def test_preprocess_plain_matches_torch(ops):
    from b200dp.data.datasets import SyntheticCIFAR10
    from b200dp.data.transforms import apply_pipeline_torch
    img = SyntheticCIFAR10(16, True).images.cuda()
    for ch in (3, 4):
        out = ops.preprocess(img, None, channels=ch)
        assert out.shape == (16, 224, 224, ch) and out.dtype == torch.bfloat16
        ref = apply_pipeline_torch(img, None)
        got = out[..., :3].permute(0, 3, 1, 2).float()
        d = _levels(got, ref)
        assert (d > 1.6).float().mean() < 1e-3, float(d.max())      # bf16 output rounding + rare .5 ties
        if ch == 4:
            assert torch.all(out[..., 3] == 0)


def test_preprocess_augment_matches_torch(ops):
    from b200dp.data.datasets import SyntheticCIFAR10
    from b200dp.data.transforms import apply_pipeline_torch, identity_params, sample_aug_params
    img = SyntheticCIFAR10(32, True).images.cuda()
    g = torch.Generator().manual_seed(3)
    p = sample_aug_params(32, g)
    p[0] = identity_params(1)[0]            # a disabled row inside an augmenting batch
    out = ops.preprocess(img, p.cuda(), channels=4)[..., :3].permute(0, 3, 1, 2).float()
    ref = apply_pipeline_torch(img, p.cuda())
    d = _levels(out, ref)
    # nearest-neighbour rotation and rounding ties may pick a neighbouring pixel on a few positions
    assert (d > 3.0).float().mean() < 0.01, (float(d.mean()), float((d > 3.0).float().mean()))
    assert d.mean() < 0.5


def _im2col_ref(x):
    """NHWC bf16 image [B,S,S,4] -> rows of the 3x3 stride-2 pad-1 taps in (ky, kx, channel) order, 27 + 5 zeros."""
    B, S = x.shape[0], x.shape[1]
    xp = torch.nn.functional.pad(x[..., :3].float(), (0, 0, 1, 1, 1, 1))            # pad W and H by one
    taps = [xp[:, ky:ky + S:2, kx:kx + S:2, :] for ky in range(3) for kx in range(3)]  # 9 x [B, S/2, S/2, 3]
    col = torch.cat(taps, dim=-1).reshape(B * (S // 2) ** 2, 27)
    return torch.nn.functional.pad(col, (0, 5))


@pytest.mark.parametrize("size", [224, 64])
def test_preprocess_col_equals_im2col_of_preprocess(ops, size):
    """The fused input kernel (image never materialised) against the unfused pair: preprocess, then im2col in torch."""
    from b200dp.data.datasets import SyntheticCIFAR10
    from b200dp.data.transforms import identity_params, sample_aug_params
    img = SyntheticCIFAR10(24, True).images.cuda()
    p = sample_aug_params(24, torch.Generator().manual_seed(5), size=size)
    p[3] = identity_params(1)[0]
    for params in (None, p.cuda()):
        ref = _im2col_ref(ops.preprocess(img, params, size=size, channels=4))
        col, sums = ops.preprocess_col(img, params, size=size, want_sums=True)
        d = (col.float() - ref).abs()
        assert (d > 0).float().mean() < 2e-4, (float(d.max()), float((d > 0).float().mean()))   # fused-multiply-add ties
        assert d.max() < 0.2
        assert torch.allclose(sums, col.float().sum(0), rtol=1e-3, atol=0.5)
        assert torch.all(col[:, 27:] == 0)


def test_softmax_xent_matches_torch(ops):
    torch.manual_seed(0)
    for B, C in ((128, 10), (40, 10), (7, 1000)):
        logits = (torch.randn(B, C, device="cuda") * 3).contiguous()
        labels = torch.randint(0, C, (B,), device="cuda")
        metrics = torch.zeros(3, dtype=torch.float64, device="cuda")
        loss, dlog = ops.softmax_xent(logits, labels, metrics)
        lr = logits.clone().requires_grad_(True)
        ref = torch.nn.functional.cross_entropy(lr, labels, reduction="none")
        ref.mean().backward()
        assert torch.allclose(loss, ref.detach(), atol=2e-5, rtol=1e-5)
        assert torch.allclose(dlog, lr.grad, atol=1e-6, rtol=1e-4)
        m = metrics.tolist()
        assert abs(m[0] - float(ref.sum())) < 1e-3 * B and m[2] == B
        assert m[1] == int((logits.argmax(1) == labels).sum())


def test_softmax_topk_matches_torch(ops):
    torch.manual_seed(1)
    logits = torch.randn(33, 10, device="cuda")
    probs, tv, ti = ops.softmax_topk(logits, 3)
    rp = torch.softmax(logits, 1)
    rv, ri = torch.topk(rp, 3, dim=1)
    assert torch.allclose(probs, rp, atol=1e-6) and torch.allclose(tv, rv, atol=1e-6)
    assert torch.equal(ti.long(), ri)


def _single_ctx():
    from b200dp.parallel.bootstrap import DistContext, RankInfo
    return DistContext(RankInfo(0, 1, 0, "single"), torch.device("cuda", 0), "nccl", False)


@pytest.mark.parametrize("gdtype", [torch.float32, torch.bfloat16])
def test_fused_adam_world1_matches_flat_adam(ops, gdtype):
    from b200dp.optim import AdamHyper, FlatAdam
    from b200dp.parallel.comm import FusedAllReduceAdam
    torch.manual_seed(0)
    n = 2_242_624
    hyper = AdamHyper(lr=1e-3, weight_decay=0.01)
    p0 = torch.randn(n, device="cuda")
    fused = FusedAllReduceAdam(_single_ctx(), n, hyper, params=p0.clone(), work_bf16=True, grad_dtype=gdtype)
    ref_p = p0.clone()
    ref = FlatAdam(n, "cuda", hyper)
    for step in range(5):
        g = torch.randn(n, device="cuda")
        fused.grads.copy_(g)
        gq = fused.grads.float().clone()
        lr = 1e-3 if step < 3 else 1e-4
        fused.step(lr)
        ref.step_torch(ref_p, gq, lr)
        assert torch.allclose(fused.params, ref_p, rtol=2e-5, atol=2e-6), step
        assert torch.count_nonzero(fused.grads) == 0              # zero_grad fused in
    assert torch.allclose(fused.m, ref.m, rtol=1e-5, atol=1e-7) and torch.allclose(fused.v, ref.v, rtol=2e-5, atol=1e-9)
    assert torch.equal(fused.work, fused.params.to(torch.bfloat16))
    assert fused.steps == 5
    # bucketed form: three launches over sub-ranges, the step counter advances on the last one only
    g = torch.randn(n, device="cuda")
    fused.grads.copy_(g)
    gq = fused.grads.float().clone()
    cuts = [0, 1_000_000, 2_000_000, n]
    for i in (2, 1, 0):
        fused.args.range_lo, fused.args.range_hi, fused.args.bump_step, fused.args.exact_blocks = cuts[i], cuts[i + 1], int(i == 0), 8
        fused.step(1e-4)
    fused.args.range_lo = fused.args.range_hi = fused.args.exact_blocks = 0
    fused.args.bump_step = 1
    ref.step_torch(ref_p, gq, 1e-4)
    assert torch.allclose(fused.params, ref_p, rtol=2e-5, atol=2e-6) and fused.steps == 6
    assert torch.count_nonzero(fused.grads) == 0
    assert torch.equal(fused.reduce_f64(torch.tensor([1.5, 2.0]), "sum").cpu(), torch.tensor([1.5, 2.0], dtype=torch.float64))
    fused.check_error()
    avg = None
    fused.grads.copy_(torch.arange(n, device="cuda") % 7)
    avg = fused.all_reduce_mean()
    assert torch.equal(avg, (torch.arange(n, device="cuda") % 7).float())
    fused.close()


def test_torch_engine_step_on_gpu_uses_native_kernels(ops):
    """The GPU path must go through the native preprocess + fused optimiser kernels."""
    from b200dp.config import TrainConfig
    from b200dp.data.datasets import SyntheticCIFAR10
    from b200dp.data.loader import Batch
    from b200dp.data.transforms import sample_aug_params
    from b200dp.engine_torch import TorchEngine
    cfg = TrainConfig(mode="gpu", engine="torch", batch_size=16)
    eng = TorchEngine(cfg, _single_ctx())
    ds = SyntheticCIFAR10(16, True)
    b = Batch(ds.images.cuda(), ds.labels.cuda(), sample_aug_params(16).cuda(), 16)
    ops.reset_launch_count()
    before = eng.flat_p.clone()
    losses = []
    for _ in range(8):
        eng.train_step(b, 1e-3)
        l, c, n = eng.read_metrics()
        losses.append(l / n)
    assert ops.launch_count() == 16                     # preprocess + fused Adam per step
    assert not torch.equal(before, eng.flat_p) and losses[-1] < losses[0]
    sd = eng.state_dict()
    assert len(sd) == 314
    eng.close()


@pytest.mark.multigpu
@pytest.mark.parametrize("world", [2])
def test_fused_allreduce_adam_multi_rank(world):
    if torch.cuda.device_count() < world:
        pytest.skip("not enough GPUs")
    env = dict(os.environ, PYTHONPATH=ROOT)
    r = subprocess.run([sys.executable, "-m", "b200dp.launch", "-np", str(world), "--",
                        os.path.join(ROOT, "tests", "mp_comm_check.py")], env=env, capture_output=True, text=True,
                       timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-4000:] + r.stderr[-4000:]
    assert "COMM CHECK OK" in r.stdout


@pytest.mark.parametrize("rows,C", [(1000, 16), (12544 * 3, 32), (777, 96), (49 * 5, 320), (5, 64), (3136 * 2, 160)])
def test_narrow_passes(rows, C):
    """block output / gradient join / BatchNorm-backward apply over the bottleneck tensors (csrc/bn_ops.cu)."""
    from b200dp import ops
    torch.manual_seed(rows + C)
    dev = "cuda"
    bf = lambda *s: torch.randn(*s, device=dev).to(torch.bfloat16)
    raw, res, g1, g2 = bf(rows, C), bf(rows, C), bf(rows, C), bf(rows, C)
    coef2 = torch.randn(2, C, device=dev)
    coef3 = torch.randn(3, C, device=dev)
    # block_out, with and without residual
    for r in (None, res):
        colsum = torch.zeros(C, device=dev)
        out = ops.block_out(raw, r, coef2, colsum)
        ref = raw.float() * coef2[0] + coef2[1] + (r.float() if r is not None else 0)
        assert torch.allclose(out.float(), ref, rtol=1e-2, atol=2e-2)
        assert torch.allclose(colsum, out.float().sum(0), rtol=1e-3, atol=0.05)
    # grad_join_stats, with and without the second branch
    for b in (None, g2):
        sums = torch.zeros(2, C, device=dev)
        dout = ops.grad_join_stats(g1, b, raw, sums)
        ref = g1.float() + (b.float() if b is not None else 0)
        assert torch.allclose(dout.float(), ref, rtol=1e-2, atol=2e-2)
        assert torch.allclose(sums[0], dout.float().sum(0), rtol=1e-3, atol=0.05)
        assert torch.allclose(sums[1], (dout.float() * raw.float()).sum(0), rtol=1e-3, atol=0.1)
    # affine: three-term form
    out = ops.affine(g1, raw, coef3, True)
    ref = coef3[0] * g1.float() + coef3[1] * raw.float() + coef3[2]
    assert torch.allclose(out.float(), ref, rtol=1e-2, atol=3e-2)
