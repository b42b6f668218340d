"""Python face of the hand-written kernels (one function per C entry point in include/b200dp.h).

Each wrapper checks shapes / dtypes / contiguity, passes raw device pointers and the current
CUDA stream, and raises on a non-zero return code.  Nothing here falls back to ATen: if the
library is missing on a GPU machine the call fails loudly.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import _lib
from ._lib import check, current_stream, ptr


def native_available() -> bool:
    """True when the native library is built and a CUDA device is present."""
    return torch.cuda.is_available() and _lib.load() is not None


def launch_count() -> int:
    lib = _lib.load()
    return int(lib.b200dp_launch_count()) if lib is not None else 0


def reset_launch_count() -> None:
    lib = _lib.load()
    if lib is not None:
        lib.b200dp_reset_launch_count()


def _chk(t: torch.Tensor, dtype, name: str) -> torch.Tensor:
    if t.dtype != dtype:
        raise TypeError(f"{name}: expected {dtype}, got {t.dtype}")
    if not t.is_cuda:
        raise ValueError(f"{name}: expected a CUDA tensor")
    if not t.is_contiguous():
        raise ValueError(f"{name}: expected a contiguous tensor")
    return t


def preprocess(images_u8: torch.Tensor, params: Optional[torch.Tensor] = None, size: int = 224,
               out_dtype=torch.bfloat16, channels: int = 4, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """uint8 [B,32,32,3] -> normalised bf16 NHWC [B,size,size,channels].
    ``params`` None = test transform; else fp32 [B,12] rows from ``data.transforms.sample_aug_params``."""
    lib = _lib.require()
    _chk(images_u8, torch.uint8, "images")
    B = images_u8.shape[0]
    if tuple(images_u8.shape[1:]) != (32, 32, 3):
        raise ValueError(f"images: expected [B,32,32,3], got {tuple(images_u8.shape)}")
    if out_dtype != torch.bfloat16:
        raise TypeError("preprocess writes bf16")
    if params is not None:
        _chk(params, torch.float32, "params")
        if tuple(params.shape) != (B, _lib.AUG_NPARAM):
            raise ValueError(f"params: expected [{B},{_lib.AUG_NPARAM}]")
    if out is None:
        out = torch.empty((B, size, size, channels), dtype=torch.bfloat16, device=images_u8.device)
    check(lib.b200dp_preprocess(ptr(images_u8), ptr(params), ptr(out), B, size, channels, current_stream()))
    return out


def softmax_xent(logits: torch.Tensor, labels: torch.Tensor, metrics: Optional[torch.Tensor] = None,
                 want_grad: bool = True, mean_over: Optional[float] = None
                 ) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """Per-sample cross-entropy, d(mean loss)/d(logits), and on-device metric accumulation
    (``metrics`` float64 [3] += loss sum, correct count, batch)."""
    lib = _lib.require()
    _chk(logits, torch.float32, "logits")
    _chk(labels, torch.int64, "labels")
    B, Cn = logits.shape
    loss = torch.empty(B, dtype=torch.float32, device=logits.device)
    dlog = torch.empty_like(logits) if want_grad else None
    if metrics is not None:
        _chk(metrics, torch.float64, "metrics")
    check(lib.b200dp_softmax_xent(ptr(logits), ptr(labels), ptr(loss), ptr(dlog), ptr(metrics), B, Cn,
                                  float(B if mean_over is None else mean_over), current_stream()))
    return loss, dlog


def softmax_topk(logits: torch.Tensor, k: int = 3) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """(probs [B,C], top-k probabilities [B,k], top-k indices int32 [B,k])."""
    lib = _lib.require()
    _chk(logits, torch.float32, "logits")
    B, Cn = logits.shape
    probs = torch.empty_like(logits)
    tv = torch.empty((B, k), dtype=torch.float32, device=logits.device)
    ti = torch.empty((B, k), dtype=torch.int32, device=logits.device)
    check(lib.b200dp_softmax_topk(ptr(logits), ptr(probs), ptr(tv), ptr(ti), B, Cn, k, current_stream()))
    return probs, tv, ti
