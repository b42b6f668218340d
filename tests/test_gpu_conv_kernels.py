"""Depthwise / stem / weight-gradient kernels against fp32 PyTorch (SURVEY W5 wgrad, W7, W8)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    from b200dp import ops as o
    from b200dp.ops import _lib
    if _lib.load() is None:
        pytest.fail(f"native library must load on a GPU box: {_lib.load_error()}")
    return o


def _bf(t):
    return t.to(torch.bfloat16)


def _relerr(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()


@pytest.mark.parametrize("impl", ["simt", "tile"])
@pytest.mark.parametrize("B,H,C,stride", [(4, 28, 192, 1), (3, 56, 96, 2), (2, 14, 576, 1), (5, 7, 960, 1),
                                          (2, 112, 32, 1), (3, 15, 144, 2), (130, 14, 384, 1), (2, 28, 144, 2),
                                          (7, 7, 960, 1), (3, 14, 576, 2), (2, 112, 96, 2), (1, 56, 144, 1)])
def test_dw_fwd(ops, B, H, C, stride, impl):
    torch.manual_seed(C + H)
    x = _bf(torch.randn(B, H, H, C, device="cuda") * 2)
    coef = torch.stack([torch.rand(C, device="cuda") + 0.5, torch.randn(C, device="cuda")]).contiguous()
    w = (torch.randn(9, C, device="cuda") * 0.3).contiguous()
    stats = torch.zeros(2, C, device="cuda")
    y = ops.dw_fwd(x, coef, w, stride, stats, impl=impl)
    a = (x.float() * coef[0] + coef[1]).clamp(0, 6).permute(0, 3, 1, 2)
    ref = F.conv2d(a, w.t().reshape(C, 1, 3, 3), stride=stride, padding=1, groups=C).permute(0, 2, 3, 1)
    assert y.shape == ref.shape
    assert (y.float() - ref).abs().max().item() <= 0.02 * ref.abs().max().item() + 1e-2
    assert torch.allclose(stats[0], y.float().sum((0, 1, 2)), rtol=2e-3, atol=0.5)
    assert torch.allclose(stats[1], (y.float() ** 2).sum((0, 1, 2)), rtol=2e-3, atol=0.5)


@pytest.mark.parametrize("impl", ["simt", "tile"])
@pytest.mark.parametrize("B,H,C,stride", [(4, 28, 192, 1), (3, 56, 96, 2), (2, 14, 576, 1), (3, 7, 960, 1),
                                          (2, 112, 32, 1), (3, 15, 144, 2), (70, 14, 384, 1), (2, 56, 144, 1),
                                          (5, 7, 960, 1), (2, 112, 96, 2), (3, 28, 192, 2), (4, 14, 576, 2), (2, 56, 144, 2)])
def test_dw_bwd(ops, B, H, C, stride, impl):
    torch.manual_seed(C + H + 1)
    Ho = (H - 1) // stride + 1
    e_raw = _bf(torch.randn(B, H, H, C, device="cuda") * 2)
    ecoef = torch.stack([torch.rand(C, device="cuda") + 0.5, torch.randn(C, device="cuda")]).contiguous()
    dzd = _bf(torch.randn(B, Ho, Ho, C, device="cuda"))
    d_raw = _bf(torch.randn(B, Ho, Ho, C, device="cuda"))
    dcoef = (torch.randn(3, C, device="cuda") * 0.5).contiguous()
    w = (torch.randn(9, C, device="cuda") * 0.3).contiguous()
    stats = torch.zeros(2, C, device="cuda")
    dw = torch.zeros(9, C, device="cuda")
    dze = ops.dw_bwd(e_raw, ecoef, dzd, d_raw, dcoef, w, stride, stats, dw, impl=impl)
    # reference with autograd
    z = e_raw.float() * ecoef[0] + ecoef[1]
    a = z.clamp(0, 6).permute(0, 3, 1, 2).detach().requires_grad_(True)
    wt = w.t().reshape(C, 1, 3, 3).detach().requires_grad_(True)
    y = F.conv2d(a, wt, stride=stride, padding=1, groups=C)
    dd = (dcoef[0] * dzd.float() + dcoef[1] * d_raw.float() + dcoef[2]).permute(0, 3, 1, 2)
    y.backward(dd)
    mask = ((z > 0) & (z < 6)).float()
    ref = a.grad.permute(0, 2, 3, 1) * mask
    assert (dze.float() - ref).abs().max().item() <= 0.02 * ref.abs().max().item() + 1e-2
    assert _relerr(dw, wt.grad.reshape(C, 9).t()) < 5e-3
    assert torch.allclose(stats[0], dze.float().sum((0, 1, 2)), rtol=2e-3, atol=0.5)
    assert torch.allclose(stats[1], (dze.float() * e_raw.float()).sum((0, 1, 2)), rtol=2e-3, atol=1.0)


@pytest.mark.parametrize("B,S", [(4, 224), (3, 64), (9, 32)])
def test_stem_fwd_and_wgrad(ops, B, S):
    torch.manual_seed(S)
    x = torch.randn(B, S, S, 3, device="cuda")
    x4 = torch.zeros(B, S, S, 4, device="cuda", dtype=torch.bfloat16)
    x4[..., :3] = _bf(x)
    w = (torch.randn(27, 32, device="cuda") * 0.2).contiguous()
    stats = torch.zeros(2, 32, device="cuda")
    y = ops.stem_fwd(x4, w, stats)
    xin = x4[..., :3].float().permute(0, 3, 1, 2)
    wt = w.reshape(3, 3, 3, 32).permute(3, 2, 0, 1).contiguous().requires_grad_(True)   # [co,ci,ky,kx]
    ref = F.conv2d(xin, wt, stride=2, padding=1)
    assert (y.float() - ref.permute(0, 2, 3, 1)).abs().max().item() <= 0.02 * ref.abs().max().item() + 1e-2
    assert torch.allclose(stats[0], y.float().sum((0, 1, 2)), rtol=2e-3, atol=0.5)
    dz = _bf(torch.randn_like(y.float()))
    coef = (torch.randn(3, 32, device="cuda") * 0.5).contiguous()
    dw = torch.zeros(27, 32, device="cuda")
    ops.stem_wgrad(x4, dz, y, coef, dw)
    dy = (coef[0] * dz.float() + coef[1] * y.float() + coef[2]).permute(0, 3, 1, 2)
    ref.backward(dy)
    refdw = wt.grad.permute(2, 3, 1, 0).reshape(27, 32)
    assert _relerr(dw, refdw) < 5e-3


@pytest.mark.parametrize("M,N,K", [(6272, 96, 16), (1000, 192, 32), (3000, 576, 96), (777, 960, 160), (500, 1280, 320),
                                   (128 * 64, 144, 32), (200, 384, 64)])
def test_pw_wgrad_mode0(ops, M, N, K):
    """expand / last conv: A pure [M,K], G = BN-backward(dz, y_raw) [M,N] -> dW [N,K]"""
    torch.manual_seed(M)
    a = _bf(torch.randn(M, K, device="cuda"))
    dz, yraw = _bf(torch.randn(M, N, device="cuda")), _bf(torch.randn(M, N, device="cuda") * 2)
    coef = (torch.randn(3, N, device="cuda") * 0.5).contiguous()
    dw = torch.zeros(N, K, device="cuda")
    ops.pw_wgrad(dw, a=a, dz=dz, y_raw=yraw, coef=coef)
    g = _bf(coef[0] * dz.float() + coef[1] * yraw.float() + coef[2]).float()
    ref = g.t() @ a.float()
    assert _relerr(dw, ref) < 5e-3, _relerr(dw, ref)


@pytest.mark.parametrize("M,N,K", [(6272, 16, 96), (1000, 32, 144), (3000, 96, 576), (777, 160, 960), (500, 320, 960),
                                   (128 * 64, 16, 32), (200, 64, 384)])
def test_pw_wgrad_mode1(ops, M, N, K):
    """project conv: G pure [M,N], A = relu6(BN(d_raw)) [M,K] -> dW [N,K]"""
    torch.manual_seed(M + 1)
    g = _bf(torch.randn(M, N, device="cuda"))
    draw = _bf(torch.randn(M, K, device="cuda") * 2)
    coef = torch.stack([torch.rand(K, device="cuda") + 0.5, torch.randn(K, device="cuda")]).contiguous()
    dw = torch.zeros(N, K, device="cuda")
    ops.pw_wgrad(dw, g=g, a_raw=draw, coef=coef)
    a = _bf((draw.float() * coef[0] + coef[1]).clamp(0, 6)).float()
    ref = g.float().t() @ a
    assert _relerr(dw, ref) < 5e-3, _relerr(dw, ref)


@pytest.mark.parametrize("M,N,K,pack", [(128 * 64, 16, 32, 4), (4000, 16, 32, 2), (1024, 32, 64, 2), (128 * 98, 16, 32, 0)])
def test_pw_wgrad_mode1_packed(ops, M, N, K, pack):
    """project conv with row packing: `pack` consecutive pixels form one row, only diagonal blocks are accumulated"""
    torch.manual_seed(M + 2)
    g = _bf(torch.randn(M, N, device="cuda"))
    draw = _bf(torch.randn(M, K, device="cuda") * 2)
    coef = torch.stack([torch.rand(K, device="cuda") + 0.5, torch.randn(K, device="cuda")]).contiguous()
    dw = torch.zeros(N, K, device="cuda")
    ops.pw_wgrad(dw, g=g, a_raw=draw, coef=coef, pack=pack)
    a = _bf((draw.float() * coef[0] + coef[1]).clamp(0, 6)).float()
    ref = g.float().t() @ a
    assert _relerr(dw, ref) < 5e-3, _relerr(dw, ref)


@pytest.mark.parametrize("M,N,K,pack", [(128 * 64, 32, 32, 4), (4000, 32, 16, 2), (128 * 50, 64, 32, 2)])
def test_pw_wgrad_mode0_packed(ops, M, N, K, pack):
    """factored BatchNorm-backward form with row packing (the cc * sum_m A term is shared out over the diagonal blocks)"""
    torch.manual_seed(M + 3)
    a = _bf(torch.randn(M, K, device="cuda"))
    dz, yraw = _bf(torch.randn(M, N, device="cuda")), _bf(torch.randn(M, N, device="cuda") * 2)
    coef = (torch.randn(3, N, device="cuda") * 0.5).contiguous()
    dw = torch.zeros(N, K, device="cuda")
    ops.pw_wgrad(dw, a=a, dz=dz, y_raw=yraw, coef=coef, pack=pack)
    g = _bf(coef[0] * dz.float() + coef[1] * yraw.float() + coef[2]).float()
    ref = g.t() @ a.float()
    assert _relerr(dw, ref) < 5e-3, _relerr(dw, ref)
