"""Image pipeline P1-P7 (SURVEY §2.4 "CPU image ops", W11, W12).

The reference runs seven PIL transforms per image inside DataLoader workers
(cifar10_mpi_mobilenet_224.py:72-82 train, :84-89 test) -- that CPU pipeline is what bounds its
throughput.  Here the same pipeline is *parameterised*: the random draws of one
sample are a 12-float row (``sample_aug_params``), and two implementations consume it:

* ``apply_pipeline_torch`` -- batched ATen ops; used by the serial-CPU mode and as the oracle;
* ``b200dp_preprocess`` (csrc/preprocess.cu) -- one sm_100a kernel, uint8 32x32x3 in,
  normalised bf16 NHWC 224x224 out, everything composed per output pixel.

Semantics follow torchvision/PIL: Resize = bilinear with half-pixel centres; RandomResizedCrop
(scale 0.7-1, ratio 3/4-4/3) = integer crop box then bilinear resize back to 224; horizontal
flip p=0.5; ColorJitter(0.3,0.3,0.3,0.1) = the four ops in a random order per image; rotation
+-15 deg with nearest sampling and zero fill; ToTensor; Normalize with ImageNet statistics.
(The reference's Gradio app normalises with CIFAR statistics by mistake, GROUP03.pdf p.26; this
framework uses ImageNet statistics everywhere -- SURVEY §2.1 quirks.)
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn.functional as F

from .. import IMAGENET_MEAN, IMAGENET_STD

NPARAM = 12
# columns of the parameter row
P_TOP, P_LEFT, P_H, P_W, P_FLIP, P_BRIGHT, P_CONTRAST, P_SAT, P_HUE, P_ORDER, P_ANGLE, P_ENABLE = range(NPARAM)

_PERMS = [(a, b, c, d) for a in range(4) for b in range(4) for c in range(4) for d in range(4)
          if len({a, b, c, d}) == 4]   # 24 orders of (brightness, contrast, saturation, hue)


def identity_params(batch: int) -> torch.Tensor:
    """Parameter rows of the test transform: Resize + ToTensor + Normalize only (ref :84-89)."""
    p = torch.zeros(batch, NPARAM, dtype=torch.float32)
    p[:, P_H] = 224
    p[:, P_W] = 224
    p[:, P_BRIGHT] = 1
    p[:, P_CONTRAST] = 1
    p[:, P_SAT] = 1
    return p


def sample_aug_params(batch: int, generator: Optional[torch.Generator] = None, size: int = 224,
                      scale=(0.7, 1.0), ratio=(3.0 / 4.0, 4.0 / 3.0), brightness: float = 0.3,
                      contrast: float = 0.3, saturation: float = 0.3, hue: float = 0.1,
                      degrees: float = 15.0) -> torch.Tensor:
    """Draw the train-time random parameters (ref :74-78) for ``batch`` samples."""
    g = generator

    def U(lo, hi, n=batch):
        return lo + (hi - lo) * torch.rand(n, generator=g)

    p = identity_params(batch)
    area = float(size * size)
    log_r = (math.log(ratio[0]), math.log(ratio[1]))
    done = torch.zeros(batch, dtype=torch.bool)
    top = torch.zeros(batch)
    left = torch.zeros(batch)
    hh = torch.full((batch,), float(size))
    ww = torch.full((batch,), float(size))
    for _ in range(10):  # torchvision retries up to 10 times, then falls back to a centre crop
        ta = area * U(scale[0], scale[1])
        ar = torch.exp(U(log_r[0], log_r[1]))
        w = torch.round(torch.sqrt(ta * ar))
        h = torch.round(torch.sqrt(ta / ar))
        ok = (w > 0) & (w <= size) & (h > 0) & (h <= size) & ~done
        i = torch.floor(torch.rand(batch, generator=g) * (size - h + 1))
        j = torch.floor(torch.rand(batch, generator=g) * (size - w + 1))
        top = torch.where(ok, i, top)
        left = torch.where(ok, j, left)
        hh = torch.where(ok, h, hh)
        ww = torch.where(ok, w, ww)
        done |= ok
    p[:, P_TOP], p[:, P_LEFT], p[:, P_H], p[:, P_W] = top, left, hh, ww
    p[:, P_FLIP] = (torch.rand(batch, generator=g) < 0.5).float()
    p[:, P_BRIGHT] = U(1 - brightness, 1 + brightness)
    p[:, P_CONTRAST] = U(1 - contrast, 1 + contrast)
    p[:, P_SAT] = U(1 - saturation, 1 + saturation)
    p[:, P_HUE] = U(-hue, hue)
    p[:, P_ORDER] = torch.floor(torch.rand(batch, generator=g) * 24).clamp_(0, 23)
    p[:, P_ANGLE] = U(-degrees, degrees)
    p[:, P_ENABLE] = 1
    return p


# ------------------------------------------------------------------------------------------
# torch implementation (CPU mode + oracle)
# ------------------------------------------------------------------------------------------

def _round_u8(x: torch.Tensor) -> torch.Tensor:
    return torch.floor(x + 0.5).clamp_(0, 255)


def resize_bilinear_u8(images_u8: torch.Tensor, size: int = 224) -> torch.Tensor:
    """P1: uint8 NHWC [B,32,32,3] -> float NCHW [B,3,size,size] holding rounded 0..255 values."""
    x = images_u8.permute(0, 3, 1, 2).to(torch.float32)
    x = F.interpolate(x, size=(size, size), mode="bilinear", align_corners=False)
    return _round_u8(x)


def _crop_resize(x: torch.Tensor, p: torch.Tensor) -> torch.Tensor:
    """P2: per-sample integer crop box, bilinear resize back to SxS (clamped inside the box)."""
    B, C, S, _ = x.shape
    dev = x.device
    o = torch.arange(S, device=dev, dtype=torch.float32) + 0.5

    def axis(start, length):
        pos = start[:, None] + o[None, :] * (length[:, None] / S) - 0.5
        lo = start[:, None]
        hi = start[:, None] + length[:, None] - 1
        pos = torch.minimum(torch.maximum(pos, lo), hi)
        i0 = torch.floor(pos)
        f = pos - i0
        i1 = torch.minimum(i0 + 1, hi)
        return i0.long(), i1.long(), f

    y0, y1, fy = axis(p[:, P_TOP], p[:, P_H])
    x0, x1, fx = axis(p[:, P_LEFT], p[:, P_W])
    bidx = torch.arange(B, device=dev)[:, None, None]

    def g(yi, xi):
        return x[bidx, :, yi[:, :, None], xi[:, None, :]].permute(0, 3, 1, 2)  # [B,C,S,S]

    fy_ = fy[:, None, :, None]
    fx_ = fx[:, None, None, :]
    top = g(y0, x0) * (1 - fx_) + g(y0, x1) * fx_
    bot = g(y1, x0) * (1 - fx_) + g(y1, x1) * fx_
    return _round_u8(top * (1 - fy_) + bot * fy_)


def _gray(x: torch.Tensor) -> torch.Tensor:
    return 0.299 * x[:, 0:1] + 0.587 * x[:, 1:2] + 0.114 * x[:, 2:3]


def _blend(a: torch.Tensor, b: torch.Tensor, f: torch.Tensor) -> torch.Tensor:
    return (a + f * (b - a)).clamp_(0, 255)


def _rgb_to_hsv(x: torch.Tensor):
    r, g, b = x[:, 0], x[:, 1], x[:, 2]
    mx = torch.maximum(torch.maximum(r, g), b)
    mn = torch.minimum(torch.minimum(r, g), b)
    d = mx - mn
    s = torch.where(mx > 0, d / mx.clamp_min(1e-12), torch.zeros_like(mx))
    dz = d.clamp_min(1e-12)
    hr = ((g - b) / dz) % 6.0
    hg = (b - r) / dz + 2.0
    hb = (r - g) / dz + 4.0
    h = torch.where(mx == r, hr, torch.where(mx == g, hg, hb))
    h = torch.where(d > 0, h / 6.0, torch.zeros_like(h))
    return h, s, mx


def _hsv_to_rgb(h, s, v):
    h6 = (h % 1.0) * 6.0
    i = torch.floor(h6)
    f = h6 - i
    p = v * (1 - s)
    q = v * (1 - s * f)
    t = v * (1 - s * (1 - f))
    i = i.long() % 6
    r = torch.where(i == 0, v, torch.where(i == 1, q, torch.where(i == 2, p, torch.where(i == 3, p, torch.where(i == 4, t, v)))))
    g = torch.where(i == 0, t, torch.where(i == 1, v, torch.where(i == 2, v, torch.where(i == 3, q, torch.where(i == 4, p, p)))))
    b = torch.where(i == 0, p, torch.where(i == 1, p, torch.where(i == 2, t, torch.where(i == 3, v, torch.where(i == 4, v, q)))))
    return torch.stack([r, g, b], dim=1)


def color_jitter(x: torch.Tensor, p: torch.Tensor) -> torch.Tensor:
    """P4 on float 0..255 NCHW images; the four ops run in each sample's own order."""
    order = torch.tensor(_PERMS, device=x.device)[p[:, P_ORDER].long()]   # [B,4]
    fb = p[:, P_BRIGHT].view(-1, 1, 1, 1)
    fc = p[:, P_CONTRAST].view(-1, 1, 1, 1)
    fs = p[:, P_SAT].view(-1, 1, 1, 1)
    fh = p[:, P_HUE].view(-1, 1, 1)
    x = x.clone()
    for pos in range(4):
        # each op only on the samples that have it at this position (a quarter of the batch on average): same values as
        # evaluating all four on everything and selecting, a quarter of the work
        for op in range(4):
            idx = (order[:, pos] == op).nonzero(as_tuple=True)[0]
            if idx.numel() == 0:
                continue
            xi = x[idx]
            if op == 0:
                xo = (xi * fb[idx]).clamp(0, 255)
            elif op == 1:
                mean = torch.floor(_gray(xi).mean(dim=(1, 2, 3), keepdim=True) + 0.5)
                xo = _blend(mean.expand_as(xi), xi, fc[idx])
            elif op == 2:
                xo = _blend(_gray(xi).expand_as(xi), xi, fs[idx])
            else:
                h, sat, v = _rgb_to_hsv(xi)
                xo = _hsv_to_rgb(h + fh[idx], sat, v)
            x[idx] = xo
    return x


def _rotate_nearest(x: torch.Tensor, p: torch.Tensor) -> torch.Tensor:
    """P5: PIL ``rotate(angle, NEAREST, expand=False, fill=0)`` about the image centre."""
    B, C, S, _ = x.shape
    dev = x.device
    ang = -p[:, P_ANGLE] * (math.pi / 180.0)
    ca, sa = torch.cos(ang), torch.sin(ang)
    # PIL: matrix = [cos, sin, ., -sin, cos, .] of (-angle), centred at (S/2, S/2)
    a, b, d, e = ca, sa, -sa, ca
    cx = cy = S / 2.0
    c = cx - (a * cx + b * cy)
    f = cy - (d * cx + e * cy)
    xs = torch.arange(S, device=dev, dtype=torch.float32) + 0.5
    X = xs[None, None, :]
    Y = xs[None, :, None]
    xin = a[:, None, None] * X + b[:, None, None] * Y + c[:, None, None]
    yin = d[:, None, None] * X + e[:, None, None] * Y + f[:, None, None]
    xi = torch.floor(xin).long()
    yi = torch.floor(yin).long()
    ok = (xi >= 0) & (xi < S) & (yi >= 0) & (yi < S)
    xi = xi.clamp(0, S - 1)
    yi = yi.clamp(0, S - 1)
    bidx = torch.arange(B, device=dev)[:, None, None]
    out = x[bidx, :, yi, xi].permute(0, 3, 1, 2)
    return out * ok[:, None].to(x.dtype)


def apply_pipeline_torch(images_u8: torch.Tensor, params: Optional[torch.Tensor] = None,
                         size: int = 224) -> torch.Tensor:
    """uint8 NHWC [B,32,32,3] -> normalised fp32 NCHW [B,3,size,size].

    ``params`` None (or all rows with enable=0) gives the test transform (C7); rows from
    ``sample_aug_params`` give the train transform (C6)."""
    x = resize_bilinear_u8(images_u8, size)
    if params is not None and bool((params[:, P_ENABLE] > 0).any()):
        p = params.to(x.device)
        x = _crop_resize(x, p)
        flip = p[:, P_FLIP].view(-1, 1, 1, 1) > 0
        x = torch.where(flip, x.flip(-1), x)
        x = color_jitter(x, p)
        x = _rotate_nearest(x, p)
    mean = torch.tensor(IMAGENET_MEAN, device=x.device).view(1, 3, 1, 1) * 255.0
    std = torch.tensor(IMAGENET_STD, device=x.device).view(1, 3, 1, 1) * 255.0
    return (x - mean) / std
