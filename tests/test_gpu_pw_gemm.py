"""tcgen05 pointwise-conv GEMM vs fp32 PyTorch: every A mode and epilogue mode, the
MobileNetV2 channel shapes, ragged M, multi-N-tile and deep-K cases."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    from b200dp import ops as o
    from b200dp.ops import _lib
    if _lib.load() is None:
        pytest.fail(f"native library must load on a GPU box: {_lib.load_error()}")
    return o


def _rand(shape, scale=1.0):
    return (torch.randn(shape, device="cuda") * scale).to(torch.bfloat16)


SHAPES = [  # (M, K, N)
    (128 * 5, 16, 96), (128 * 3 + 40, 32, 192), (1000, 64, 384), (777, 96, 576), (512, 160, 960),
    (300, 320, 1280), (128 * 9, 96, 32), (650, 144, 32), (900, 576, 160), (260, 960, 320), (4096, 32, 16),
    (128 * 200, 16, 96),
]


@pytest.mark.parametrize("M,K,N", SHAPES)
def test_fwd_tma_stats(ops, M, K, N):
    torch.manual_seed(M + K + N)
    a, w = _rand((M, K)), _rand((N, K), 0.2)
    stats = torch.zeros(2, N, device="cuda")
    out = ops.pw_gemm(w, a=a, emode=ops.E_STATS, stats=stats)
    ref = a.float() @ w.float().t()
    err = (out.float() - ref).abs().max().item()
    tol = 0.02 * ref.abs().max().item() + 1e-2
    assert err <= tol, (err, tol)
    o = out.float()
    assert torch.allclose(stats[0], o.sum(0), rtol=2e-3, atol=2e-2 * (M ** 0.5))
    assert torch.allclose(stats[1], (o * o).sum(0), rtol=2e-3, atol=1e-2 * M ** 0.5)


@pytest.mark.parametrize("M,K,N", [(640, 96, 32), (1000, 144, 32), (390, 32, 16), (515, 960, 320), (2048, 384, 64)])
@pytest.mark.parametrize("relu6", [True, False])
def test_fwd_bnrelu_stats(ops, M, K, N, relu6):
    torch.manual_seed(1)
    raw, w = _rand((M, K), 2.0), _rand((N, K), 0.2)
    coef = torch.stack([torch.rand(K, device="cuda") + 0.5, torch.randn(K, device="cuda")]).contiguous()
    stats = torch.zeros(2, N, device="cuda")
    out = ops.pw_gemm(w, a_raw=raw, acoef=coef, relu6=relu6, emode=ops.E_STATS, stats=stats)
    act = raw.float() * coef[0] + coef[1]
    if relu6:
        act = act.clamp(0, 6)
    ref = act.to(torch.bfloat16).float() @ w.float().t()
    err = (out.float() - ref).abs().max().item()
    assert err <= 0.02 * ref.abs().max().item() + 1e-2, err
    assert torch.allclose(stats[0], out.float().sum(0), rtol=2e-3, atol=0.5)


@pytest.mark.parametrize("M,K,N", [(640, 96, 16), (1000, 144, 32), (333, 1280, 320), (2048, 576, 96)])
def test_dgrad_bngrad_plain(ops, M, K, N):
    torch.manual_seed(2)
    dz, x, w = _rand((M, K)), _rand((M, K), 2.0), _rand((N, K), 0.2)
    coef = torch.randn(3, K, device="cuda").contiguous()
    out = ops.pw_gemm(w, a_dz=dz, a_raw=x, acoef=coef)
    a = (coef[0] * dz.float() + coef[1] * x.float() + coef[2]).to(torch.bfloat16).float()
    ref = a @ w.float().t()
    err = (out.float() - ref).abs().max().item()
    assert err <= 0.02 * ref.abs().max().item() + 1e-2, err


@pytest.mark.parametrize("M,K,N", [(640, 16, 96), (1000, 32, 144), (300, 160, 960), (2048, 64, 384), (128 * 40 + 7, 32, 192)])
def test_dgrad_maskstats(ops, M, K, N):
    torch.manual_seed(3)
    g, w = _rand((M, K)), _rand((N, K), 0.2)
    aux = _rand((M, N), 2.0)
    ecoef = torch.stack([torch.rand(N, device="cuda") + 0.5, torch.randn(N, device="cuda")]).contiguous()
    stats = torch.zeros(2, N, device="cuda")
    out = ops.pw_gemm(w, a=g, emode=ops.E_MASKSTATS, aux=aux, ecoef=ecoef, stats=stats)
    z = aux.float() * ecoef[0] + ecoef[1]
    mask = ((z > 0) & (z < 6)).float()
    ref = (g.float() @ w.float().t()) * mask
    err = (out.float() - ref).abs().max().item()
    assert err <= 0.02 * ref.abs().max().item() + 1e-2, err
    o = out.float()
    assert torch.allclose(stats[0], o.sum(0), rtol=2e-3, atol=0.5)
    assert torch.allclose(stats[1], (o * aux.float()).sum(0), rtol=2e-3, atol=1.0)


def test_small_grid_many_tiles_per_cta(ops):
    """Few CTAs, many tiles each: exercises the stage / accumulator phase wrap-around."""
    torch.manual_seed(4)
    M, K, N = 128 * 37 + 5, 144, 192
    a, w = _rand((M, K)), _rand((N, K), 0.2)
    stats = torch.zeros(2, N, device="cuda")
    out = ops.pw_gemm(w, a=a, emode=ops.E_STATS, stats=stats, max_ctas=3)
    ref = a.float() @ w.float().t()
    assert (out.float() - ref).abs().max().item() <= 0.02 * ref.abs().max().item() + 1e-2
    assert torch.allclose(stats[0], out.float().sum(0), rtol=2e-3, atol=0.5)


# ---- row packing: r consecutive pixels as one row against blockdiag_r(W) (narrow layers) ----------------------------
@pytest.mark.parametrize("M,K,N,pack", [(128 * 6, 16, 96, 4), (1000, 32, 16, 4), (1284, 32, 144, 2), (2048, 16, 32, 4),
                                        (128 * 100, 16, 96, 4), (770, 32, 192, 2)])
def test_packed_fwd_tma_stats(ops, M, K, N, pack):
    torch.manual_seed(M + K + N)
    a, w = _rand((M, K)), _rand((N, K), 0.2)
    stats = torch.zeros(2, N, device="cuda")
    out = ops.pw_gemm(w, a=a, emode=ops.E_STATS, stats=stats, pack=pack)
    ref = a.float() @ w.float().t()
    assert (out.float() - ref).abs().max().item() <= 0.02 * ref.abs().max().item() + 1e-2
    o = out.float()
    assert torch.allclose(stats[0], o.sum(0), rtol=2e-3, atol=2e-2 * (M ** 0.5))
    assert torch.allclose(stats[1], (o * o).sum(0), rtol=2e-3, atol=1e-2 * M ** 0.5)


@pytest.mark.parametrize("M,K,N,pack", [(1000, 32, 16, 4), (640, 96, 32, 2), (1284, 144, 32, 2), (2048, 192, 32, 2)])
def test_packed_fwd_bnrelu_stats(ops, M, K, N, pack):
    torch.manual_seed(5)
    raw, w = _rand((M, K), 2.0), _rand((N, K), 0.2)
    coef = torch.stack([torch.rand(K, device="cuda") + 0.5, torch.randn(K, device="cuda")]).contiguous()
    stats = torch.zeros(2, N, device="cuda")
    out = ops.pw_gemm(w, a_raw=raw, acoef=coef, relu6=True, emode=ops.E_STATS, stats=stats, pack=pack)
    act = (raw.float() * coef[0] + coef[1]).clamp(0, 6)
    ref = act.to(torch.bfloat16).float() @ w.float().t()
    assert (out.float() - ref).abs().max().item() <= 0.02 * ref.abs().max().item() + 1e-2
    assert torch.allclose(stats[0], out.float().sum(0), rtol=2e-3, atol=0.5)


@pytest.mark.parametrize("M,K,N,pack", [(640, 96, 16, 4), (1000, 144, 32, 2), (2048, 192, 32, 2)])
def test_packed_dgrad_bngrad(ops, M, K, N, pack):
    torch.manual_seed(6)
    dz, x, w = _rand((M, K)), _rand((M, K), 2.0), _rand((N, K), 0.2)
    coef = torch.randn(3, K, device="cuda").contiguous()
    out = ops.pw_gemm(w, a_dz=dz, a_raw=x, acoef=coef, pack=pack)
    a = (coef[0] * dz.float() + coef[1] * x.float() + coef[2]).to(torch.bfloat16).float()
    ref = a @ w.float().t()
    assert (out.float() - ref).abs().max().item() <= 0.02 * ref.abs().max().item() + 1e-2


@pytest.mark.parametrize("M,K,N,pack", [(640, 16, 32, 4), (1000, 32, 144, 2), (128 * 40 + 6, 32, 192, 2), (2048, 16, 96, 4)])
def test_packed_dgrad_maskstats(ops, M, K, N, pack):
    torch.manual_seed(7)
    g, w = _rand((M, K)), _rand((N, K), 0.2)
    aux = _rand((M, N), 2.0)
    ecoef = torch.stack([torch.rand(N, device="cuda") + 0.5, torch.randn(N, device="cuda")]).contiguous()
    stats = torch.zeros(2, N, device="cuda")
    out = ops.pw_gemm(w, a=g, emode=ops.E_MASKSTATS, aux=aux, ecoef=ecoef, stats=stats, pack=pack)
    z = aux.float() * ecoef[0] + ecoef[1]
    mask = ((z > 0) & (z < 6)).float()
    ref = (g.float() @ w.float().t()) * mask
    assert (out.float() - ref).abs().max().item() <= 0.02 * ref.abs().max().item() + 1e-2
    o = out.float()
    assert torch.allclose(stats[0], o.sum(0), rtol=2e-3, atol=0.5)
    assert torch.allclose(stats[1], (o * aux.float()).sum(0), rtol=2e-3, atol=1.0)


@pytest.mark.parametrize("M,K,N", [(640, 96, 16), (1000, 144, 32), (2048, 576, 96), (128 * 9 + 5, 960, 160)])
def test_dgrad_bngrad_with_stats_epilogue(ops, M, K, N):
    """A_BNGRAD prologue (dx = ca*dz + cb*x + cc on the A tile) together with the E_MASKSTATS epilogue, as the executor uses
    it with an always-open mask (scale 0, shift 3) to take the sums of the next BatchNorm backward inside the GEMM."""
    torch.manual_seed(5)
    dz, x, w = _rand((M, K)), _rand((M, K), 2.0), _rand((N, K), 0.2)
    coef = torch.randn(3, K, device="cuda").contiguous()
    aux = _rand((M, N), 2.0)
    ecoef = torch.stack([torch.zeros(N, device="cuda"), torch.full((N,), 3.0, device="cuda")]).contiguous()
    stats = torch.zeros(2, N, device="cuda")
    out = ops.pw_gemm(w, a_dz=dz, a_raw=x, acoef=coef, emode=ops.E_MASKSTATS, aux=aux, ecoef=ecoef, stats=stats)
    a = (coef[0] * dz.float() + coef[1] * x.float() + coef[2]).to(torch.bfloat16).float()
    ref = a @ w.float().t()
    err = (out.float() - ref).abs().max().item()
    assert err <= 0.02 * ref.abs().max().item() + 1e-2, err
    o = out.float()
    assert torch.allclose(stats[0], o.sum(0), rtol=2e-3, atol=0.5)
    assert torch.allclose(stats[1], (o * aux.float()).sum(0), rtol=2e-3, atol=1.0)
