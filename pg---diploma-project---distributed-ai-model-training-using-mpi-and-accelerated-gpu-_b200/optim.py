"""Adam and StepLR on flat arenas.

The reference uses ``optim.Adam(params, lr=1e-4)`` (defaults betas (0.9, 0.999), eps 1e-8, no
weight decay, no amsgrad; cifar10_mpi_mobilenet_224.py:148) and ``StepLR(10, 0.1)`` stepped once
per epoch (:149, :226).  Here parameters, gradients and both moments are single flat fp32
vectors, so one optimiser step is one pass over ``[0, total)``:

* ``FlatAdam.step_torch`` -- five ATen vector ops; the CPU mode and the oracle for the kernel;
* the GPU path never calls it: ``parallel.comm.FusedAllReduceAdam`` runs the same arithmetic inside
  the all-reduce kernel (csrc/comm_adam.cu), reading gradients from peer memory.

The arithmetic (bias correction folded into a step size and an epsilon-side factor) is written
once in ``adam_scalars`` and shared by both, and matches ``torch.optim.Adam`` bit-for-bit in the
scalar factors.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, Tuple

import torch


def step_lr(base_lr: float, epoch: int, step_size: int = 10, gamma: float = 0.1) -> float:
    """LR used during ``epoch`` (0-based) when ``scheduler.step()`` runs at each epoch end (ref :226)."""
    return base_lr * (gamma ** (epoch // step_size))


def adam_scalars(step: int, lr: float, beta1: float, beta2: float) -> Tuple[float, float]:
    """(step_size, inv_sqrt_bias2) for 1-based ``step``:
    ``p -= step_size * m / (sqrt(v) * inv_sqrt_bias2 + eps)`` as in torch's single-tensor Adam."""
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    return lr / bc1, 1.0 / math.sqrt(bc2)


@dataclass
class AdamHyper:
    lr: float = 1e-4
    beta1: float = 0.9
    beta2: float = 0.999
    eps: float = 1e-8
    weight_decay: float = 0.0


class FlatAdam:
    """Adam state for one flat parameter vector."""

    def __init__(self, numel: int, device, hyper: AdamHyper | None = None) -> None:
        self.h = hyper or AdamHyper()
        self.m = torch.zeros(numel, dtype=torch.float32, device=device)
        self.v = torch.zeros(numel, dtype=torch.float32, device=device)
        self.step = 0

    @torch.no_grad()
    def step_torch(self, params: torch.Tensor, grads: torch.Tensor, lr: float | None = None,
                   grad_scale: float = 1.0) -> None:
        """Reference implementation with ATen ops.  ``grad_scale`` is the 1/world factor."""
        h = self.h
        lr = h.lr if lr is None else lr
        self.step += 1
        g = grads.to(torch.float32)
        if grad_scale != 1.0:
            g = g * grad_scale
        if h.weight_decay:
            g = g + h.weight_decay * params
        self.m.mul_(h.beta1).add_(g, alpha=1 - h.beta1)
        self.v.mul_(h.beta2).addcmul_(g, g, value=1 - h.beta2)
        step_size, isb2 = adam_scalars(self.step, lr, h.beta1, h.beta2)
        denom = self.v.sqrt().mul_(isb2).add_(h.eps)
        params.addcdiv_(self.m, denom, value=-step_size)

    def state_dict(self) -> Dict[str, object]:
        return {"m": self.m.detach().cpu(), "v": self.v.detach().cpu(), "step": self.step,
                "hyper": vars(self.h).copy()}

    def load_state_dict(self, sd: Dict[str, object]) -> None:
        self.m.copy_(sd["m"])
        self.v.copy_(sd["v"])
        self.step = int(sd["step"])
