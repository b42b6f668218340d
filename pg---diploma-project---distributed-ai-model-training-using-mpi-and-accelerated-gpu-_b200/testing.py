"""Oracles shared by the GPU tests and the diagnostic scripts."""
from __future__ import annotations

from typing import List, Tuple

import torch
import torch.nn.functional as F


def cos(a: torch.Tensor, b: torch.Tensor) -> float:
    return F.cosine_similarity(a.flatten().double(), b.flatten().double(), dim=0).item()


def _ste(t: torch.Tensor, native: torch.Tensor) -> torch.Tensor:
    """value of ``native``, gradient of ``t``"""
    return native + (t - t.detach())


def teacher_forced_backward(ref, core, x: torch.Tensor, labels: torch.Tensor, n: int
                            ) -> Tuple[float, List[Tuple[str, float]]]:
    """Run ``ref`` (our torch MobileNetV2, train mode, dropout 0) forward with every conv output and block
    output replaced by the native executor's value (straight-through), then backward.  Leaves gradients in
    ``ref.parameters()``; returns (loss, per-conv forward relative error of torch-on-native-input vs native)."""
    spec = core.layout.spec
    errs: List[Tuple[str, float]] = []

    def raw(i: int) -> torch.Tensor:
        c = spec.convs[i]
        cp = (c.cout + 15) // 16 * 16
        t = core.tap(0, i, core.batch * c.hout * c.hout * cp)[: n * c.hout * c.hout * cp]
        return t.view(n, c.hout, c.hout, cp)[..., : c.cout].float().permute(0, 3, 1, 2).contiguous()

    def xout(bi: int) -> torch.Tensor:
        b = spec.blocks[bi]
        cp = (b.cout + 15) // 16 * 16
        t = core.tap(3, bi, core.batch * b.hout * b.hout * cp)[: n * b.hout * b.hout * cp]
        return t.view(n, b.hout, b.hout, cp)[..., : b.cout].float().permute(0, 3, 1, 2).contiguous()

    ci = [0]

    def conv_bn(h, conv, bn, relu, bf16_input):
        i = ci[0]
        ci[0] += 1
        hin = h.to(torch.bfloat16).float() if bf16_input else h     # GEMM operands are bf16
        hin = h + (hin - h).detach()
        w = conv.weight
        if conv.kernel_size == (1, 1):
            w = w + (w.to(torch.bfloat16).float() - w).detach()
        y = F.conv2d(hin, w, None, conv.stride, conv.padding, 1, conv.groups)
        nat = raw(i)
        errs.append((spec.convs[i].conv_key, ((y.detach() - nat).norm() / nat.norm()).item()))
        y = _ste(y, nat)
        y = F.batch_norm(y, None, None, bn.weight, bn.bias, True, 0.0, bn.eps)
        return F.relu6(y) if relu else y

    ref.train()
    ref.zero_grad()
    f = ref.features
    h = conv_bn(x, f[0][0], f[0][1], True, False)
    for bi, blk in enumerate(f[1:-1]):
        seq = blk.conv
        inp = h
        k = 0
        if len(seq) == 4:
            h = conv_bn(h, seq[0][0], seq[0][1], True, True)
            k = 1
        h = conv_bn(h, seq[k][0], seq[k][1], True, False)
        h = conv_bn(h, seq[k + 1], seq[k + 2], False, True)
        if blk.use_res:
            h = h + inp
        h = _ste(h, xout(bi))
    h = conv_bn(h, f[-1][0], f[-1][1], True, True)
    pooled = h.mean((2, 3))
    logits = ref.classifier[1](pooled)
    loss = F.cross_entropy(logits, labels)
    loss.backward()
    return loss.item(), errs
