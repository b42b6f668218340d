"""Flat Adam == torch.optim.Adam."""
import torch

from b200dp.optim import AdamHyper, FlatAdam


def test_flat_adam_matches_torch_adam():
    torch.manual_seed(0)
    n = 1000
    p0 = torch.randn(n)
    ref_p = torch.nn.Parameter(p0.clone())
    opt = torch.optim.Adam([ref_p], lr=1e-3)
    ours_p = p0.clone()
    adam = FlatAdam(n, "cpu", AdamHyper(lr=1e-3))
    for step in range(20):
        g = torch.randn(n)
        ref_p.grad = g.clone()
        opt.step()
        adam.step_torch(ours_p, g, lr=1e-3)
        assert torch.allclose(ours_p, ref_p.detach(), rtol=1e-6, atol=1e-7), step
    assert adam.step == 20


def test_grad_scale_is_world_average():
    n = 64
    p1, p2 = torch.ones(n), torch.ones(n)
    a1, a2 = FlatAdam(n, "cpu"), FlatAdam(n, "cpu")
    g = torch.randn(n)
    a1.step_torch(p1, g * 4, grad_scale=0.25)
    a2.step_torch(p2, g)
    assert torch.allclose(p1, p2)
