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


def preprocess_col(images_u8: torch.Tensor, params: Optional[torch.Tensor] = None, size: int = 224,
                   want_sums: bool = False) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """uint8 [B,32,32,3] -> the stem's im2col matrix, bf16 [B*(size/2)^2, 32] (27 taps in (ky, kx, channel) order + 5
    zeros), without materialising the image; optionally the fp32 [32] column sums."""
    lib = _lib.require()
    _chk(images_u8, torch.uint8, "images")
    B = images_u8.shape[0]
    if params is not None:
        _chk(params, torch.float32, "params")
    col = torch.empty((B * (size // 2) ** 2, 32), dtype=torch.bfloat16, device=images_u8.device)
    sums = torch.zeros(32, dtype=torch.float32, device=images_u8.device) if want_sums else None
    check(lib.b200dp_preprocess_col(ptr(images_u8), ptr(params), ptr(col), ptr(sums), B, size, current_stream()))
    return col, sums


def resize_normalize(images_u8: torch.Tensor, size: int = 224, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """uint8 [B,H,W,3] of any size -> normalised bf16 NHWC [B,size,size,4]: PIL-style anti-aliased bilinear resize,
    ToTensor, Normalize (the reference's inference transform on the original image)."""
    lib = _lib.require()
    _chk(images_u8, torch.uint8, "images")
    if images_u8.dim() != 4 or images_u8.shape[-1] != 3:
        raise ValueError(f"images: expected [B,H,W,3], got {tuple(images_u8.shape)}")
    B, H, W = images_u8.shape[:3]
    if out is None:
        out = torch.empty((B, size, size, 4), dtype=torch.bfloat16, device=images_u8.device)
    check(lib.b200dp_resize_normalize(ptr(images_u8), B, H, W, ptr(out), size, current_stream()))
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


A_TMA, A_BNRELU, A_BNGRAD = 0, 1, 2
E_PLAIN, E_STATS, E_MASKSTATS = 0, 1, 2


def pw_gemm(w: torch.Tensor, *, a: Optional[torch.Tensor] = None, a_raw: Optional[torch.Tensor] = None,
            a_dz: Optional[torch.Tensor] = None, acoef: Optional[torch.Tensor] = None, relu6: bool = True,
            emode: int = E_PLAIN, aux: Optional[torch.Tensor] = None, ecoef: Optional[torch.Tensor] = None,
            stats: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None, max_ctas: int = 0,
            pack: int = 1) -> torch.Tensor:
    """Pointwise-conv GEMM ``out[M,N] = Aeff[M,K] @ w[N,K]^T`` on tcgen05 (csrc/pw_gemm.cu).

    A operand: ``a`` (bf16 [M,K], TMA path), or ``a_raw`` + ``acoef`` [2,K] (BatchNorm scale/shift
    [+ReLU6] applied on load), or ``a_dz`` + ``a_raw`` + ``acoef`` [3,K] (BatchNorm input gradient
    ``ca*dz + cb*x + cc`` formed on load).  ``emode``: E_STATS accumulates ``stats[0] += sum(out)``,
    ``stats[1] += sum(out^2)`` per channel; E_MASKSTATS masks with ReLU6'(ecoef[0]*aux + ecoef[1]) and
    accumulates ``sum(out)``, ``sum(out*aux)``.
    """
    lib = _lib.require()
    _chk(w, torch.bfloat16, "w")
    N, K = w.shape
    if a is not None:
        amode, src = A_TMA, _chk(a, torch.bfloat16, "a")
    elif a_dz is not None:
        amode, src = A_BNGRAD, _chk(a_dz, torch.bfloat16, "a_dz")
        _chk(a_raw, torch.bfloat16, "a_raw")
    else:
        amode, src = A_BNRELU, _chk(a_raw, torch.bfloat16, "a_raw")
    M = src.shape[0]
    if src.shape[1] != K:
        raise ValueError(f"A has {src.shape[1]} channels, weights expect {K}")
    if amode != A_TMA:
        _chk(acoef, torch.float32, "acoef")
        if tuple(acoef.shape) != ((3 if amode == A_BNGRAD else 2), K):
            raise ValueError("acoef shape")
    if out is None:
        out = torch.empty((M, N), dtype=torch.bfloat16, device=w.device)
    if emode != E_PLAIN:
        _chk(stats, torch.float32, "stats")
        if tuple(stats.shape) != (2, N):
            raise ValueError("stats must be [2, N]")
    if emode == E_MASKSTATS:
        _chk(aux, torch.bfloat16, "aux")
        _chk(ecoef, torch.float32, "ecoef")
    a0 = a_dz if amode == A_BNGRAD else a_raw
    a1 = a_raw if amode == A_BNGRAD else None
    if pack > 1:     # r consecutive rows as one row of r*K channels against blockdiag_r(w) (narrow layers)
        wbd = torch.block_diag(*([w.float()] * pack)).to(torch.bfloat16).contiguous()
        check(lib.b200dp_pw_gemm_packed(amode, emode, M, N, K, ptr(a), ptr(a0), ptr(a1), ptr(acoef), 1 if relu6 else 0,
                                        ptr(wbd), ptr(out), ptr(aux), ptr(ecoef), ptr(stats), N, max_ctas, pack,
                                        current_stream()))
        return out
    check(lib.b200dp_pw_gemm(amode, emode, M, N, K, ptr(a), K, ptr(a0), ptr(a1), ptr(acoef), 1 if relu6 else 0,
                             ptr(w), ptr(out), N, ptr(aux), ptr(ecoef), ptr(stats), N, max_ctas,
                             current_stream()))
    return out


def pw_wgrad(dw: torch.Tensor, *, a: Optional[torch.Tensor] = None, dz: Optional[torch.Tensor] = None,
             y_raw: Optional[torch.Tensor] = None, g: Optional[torch.Tensor] = None,
             a_raw: Optional[torch.Tensor] = None, coef: torch.Tensor = None, max_ctas: int = 0,
             pack: int = 1) -> torch.Tensor:
    """``dw[N,K] += G^T @ A`` on tcgen05 with MN-major operands (csrc/pw_wgrad.cu).
    mode 0: ``a`` pure [M,K], G = coef[0]*dz + coef[1]*y_raw + coef[2] ([M,N]);
    mode 1: ``g`` pure [M,N], A = relu6(coef[0]*a_raw + coef[1]) ([M,K])."""
    lib = _lib.require()
    _chk(dw, torch.float32, "dw")
    N, K = dw.shape
    if a is not None:
        mode, pure, t0, t1, ldp, ldt = 0, _chk(a, torch.bfloat16, "a"), _chk(dz, torch.bfloat16, "dz"), \
            _chk(y_raw, torch.bfloat16, "y_raw"), K, N
    else:
        mode, pure, t0, t1, ldp, ldt = 1, _chk(g, torch.bfloat16, "g"), _chk(a_raw, torch.bfloat16, "a_raw"), None, N, K
    M = pure.shape[0]
    _chk(coef, torch.float32, "coef")
    asum = a.float().sum(0).contiguous() if mode == 0 else None     # the BN-backward affine is factored out of the reduction
    if pack != 1:    # row packing (0 = let the library choose)
        check(lib.b200dp_pw_wgrad_packed(mode, M, N, K, ptr(pure), ptr(t0), ptr(t1), ptr(coef), ptr(asum), ptr(dw),
                                         max_ctas, pack, current_stream()))
        return dw
    check(lib.b200dp_pw_wgrad(mode, M, N, K, ptr(pure), ldp, ptr(t0), ptr(t1), ldt, ptr(coef), ptr(asum), ptr(dw), K,
                              max_ctas, current_stream()))
    return dw


def dw_fwd(x: torch.Tensor, coef: torch.Tensor, w: torch.Tensor, stride: int,
           stats: Optional[torch.Tensor] = None, impl: str = "simt") -> torch.Tensor:
    """Depthwise 3x3 on NHWC bf16 ``x`` [B,H,W,C] with relu6(coef[0]*x+coef[1]) applied on load; ``w`` fp32 [9,C].

    ``impl``: "simt" = register sliding-window kernel (dwconv.cu), "tile" = TMA halo tiles + shared-memory stencil
    (dw_tile.cu)."""
    lib = _lib.require()
    fn = {"tile": lib.b200dp_dw_fwd_tile, "simt": lib.b200dp_dw_fwd}[impl]
    _chk(x, torch.bfloat16, "x"); _chk(coef, torch.float32, "coef"); _chk(w, torch.float32, "w")
    B, H, W, Cn = x.shape
    Ho, Wo = (H - 1) // stride + 1, (W - 1) // stride + 1
    y = torch.empty((B, Ho, Wo, Cn), dtype=torch.bfloat16, device=x.device)
    check(fn(ptr(x), ptr(coef), coef.data_ptr() + 4 * Cn, ptr(w), ptr(y), ptr(stats), Cn, B, H, W, Cn, stride,
             current_stream()))
    return y


def dw_bwd(e_raw: torch.Tensor, ecoef: torch.Tensor, dzd: torch.Tensor, d_raw: torch.Tensor, dcoef: torch.Tensor,
           w: torch.Tensor, stride: int, stats: torch.Tensor, dw: torch.Tensor, impl: str = "simt") -> torch.Tensor:
    lib = _lib.require()
    fn = lib.b200dp_dw_bwd_tile if impl == "tile" else lib.b200dp_dw_bwd
    B, H, W, Cn = e_raw.shape
    dze = torch.empty_like(e_raw)
    check(fn(ptr(e_raw), ptr(ecoef), ecoef.data_ptr() + 4 * Cn, ptr(dzd), ptr(d_raw), ptr(dcoef), ptr(w),
                            ptr(dze), ptr(stats), Cn, ptr(dw), B, H, W, Cn, stride, current_stream()))
    return dze


def stem_fwd(x4: torch.Tensor, w: torch.Tensor, stats: Optional[torch.Tensor] = None) -> torch.Tensor:
    lib = _lib.require()
    B, H, W, _ = x4.shape
    y = torch.empty((B, (H - 1) // 2 + 1, (W - 1) // 2 + 1, 32), dtype=torch.bfloat16, device=x4.device)
    check(lib.b200dp_stem_fwd(ptr(x4), ptr(w), ptr(y), ptr(stats), 32, B, H, W, current_stream()))
    return y


def stem_wgrad(x4: torch.Tensor, dz: torch.Tensor, yraw: torch.Tensor, coef: torch.Tensor, dw: torch.Tensor):
    lib = _lib.require()
    B, H, W, _ = x4.shape
    check(lib.b200dp_stem_wgrad(ptr(x4), ptr(dz), ptr(yraw), ptr(coef), ptr(dw), B, H, W, current_stream()))
    return dw


def affine(a: torch.Tensor, b: Optional[torch.Tensor], coef: torch.Tensor, three: bool) -> torch.Tensor:
    """``three``: out = coef[0]*a + coef[1]*b + coef[2] (BatchNorm backward apply); else out = coef[0]*a + coef[1] (+ b)."""
    lib = _lib.require()
    C = a.shape[-1]
    out = torch.empty_like(a)
    check(lib.b200dp_affine(ptr(a), ptr(b), ptr(coef), int(three), ptr(out), a.numel() // C, C, current_stream()))
    return out


def grad_join_stats(g1: torch.Tensor, g2: Optional[torch.Tensor], x: torch.Tensor, sums: torch.Tensor):
    """dout = g1 (+ g2); sums[0] += sum dout, sums[1] += sum dout*x per channel.  Returns dout (g1 itself if g2 is None)."""
    lib = _lib.require()
    C = g1.shape[-1]
    dout = torch.empty_like(g1) if g2 is not None else None
    check(lib.b200dp_grad_join_stats(ptr(g1), ptr(g2), ptr(x), ptr(dout), ptr(sums), C, g1.numel() // C, C,
                                     current_stream()))
    return dout if dout is not None else g1


def block_out(raw: torch.Tensor, res: Optional[torch.Tensor], coef: torch.Tensor,
              colsum: Optional[torch.Tensor] = None) -> torch.Tensor:
    """out = coef[0]*raw + coef[1] (+ res); colsum[C] += column sums of the rounded result."""
    lib = _lib.require()
    C = raw.shape[-1]
    out = torch.empty_like(raw)
    check(lib.b200dp_block_out(ptr(raw), ptr(res), ptr(coef), ptr(out), ptr(colsum), raw.numel() // C, C,
                               current_stream()))
    return out
