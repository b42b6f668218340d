"""Micro-benchmarks of single kernels at MobileNetV2 layer shapes (batch 128), CUDA-event timed with an L2
flush between iterations; also the target for `ncu` captures (select one case with --only).

    python scripts/bench_kernels.py [--only NAME] [--iters 10]
"""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200dp import ops

HBM = 6553.3e9
ap = argparse.ArgumentParser()
ap.add_argument("--only", default="")
ap.add_argument("--iters", type=int, default=10)
ns = ap.parse_args()
dev = "cuda"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
bf = lambda *s: (torch.randn(*s, device=dev)).to(torch.bfloat16)


def timeit(fn):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(ns.iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


cases = {}
B = int(os.environ.get("BENCH_B", "128"))


def pw_fwd_expand(h, cin, cout):
    M = B * h * h
    a, w = bf(M, cin), bf(cout, cin)
    stats = torch.zeros(2, cout, device=dev)
    out = torch.empty(M, cout, dtype=torch.bfloat16, device=dev)
    return (lambda: ops.pw_gemm(w, a=a, emode=ops.E_STATS, stats=stats, out=out)), M * (cin + cout) * 2


def pw_fwd_project(h, cin, cout):
    M = B * h * h
    a, w = bf(M, cin), bf(cout, cin)
    coef = torch.rand(2, cin, device=dev)
    stats = torch.zeros(2, cout, device=dev)
    out = torch.empty(M, cout, dtype=torch.bfloat16, device=dev)
    return (lambda: ops.pw_gemm(w, a_raw=a, acoef=coef, emode=ops.E_STATS, stats=stats, out=out)), M * (cin + cout) * 2


def pw_dgrad_project(h, hidden, cout):
    M = B * h * h
    g, w, aux = bf(M, cout), bf(hidden, cout), bf(M, hidden)
    ecoef = torch.rand(2, hidden, device=dev)
    stats = torch.zeros(2, hidden, device=dev)
    out = torch.empty(M, hidden, dtype=torch.bfloat16, device=dev)
    return (lambda: ops.pw_gemm(w, a=g, emode=ops.E_MASKSTATS, aux=aux, ecoef=ecoef, stats=stats, out=out)), M * (cout + 2 * hidden) * 2


def pw_dgrad_expand(h, cin, hidden):
    M = B * h * h
    dz, x, w = bf(M, hidden), bf(M, hidden), bf(cin, hidden)
    coef = torch.rand(3, hidden, device=dev)
    out = torch.empty(M, cin, dtype=torch.bfloat16, device=dev)
    return (lambda: ops.pw_gemm(w, a_dz=dz, a_raw=x, acoef=coef, out=out)), M * (2 * hidden + cin) * 2


def pw_wgrad_expand(h, cin, hidden):
    M = B * h * h
    a, dz, y = bf(M, cin), bf(M, hidden), bf(M, hidden)
    coef = torch.rand(3, hidden, device=dev)
    dw = torch.zeros(hidden, cin, device=dev)
    return (lambda: ops.pw_wgrad(dw, a=a, dz=dz, y_raw=y, coef=coef)), M * (2 * hidden + cin) * 2


def pw_wgrad_project(h, hidden, cout):
    M = B * h * h
    g, d = bf(M, cout), bf(M, hidden)
    coef = torch.rand(2, hidden, device=dev)
    dw = torch.zeros(cout, hidden, device=dev)
    return (lambda: ops.pw_wgrad(dw, g=g, a_raw=d, coef=coef)), M * (hidden + cout) * 2


def dw_fwd(h, c, stride, impl="simt"):
    x = bf(B, h, h, c)
    coef = torch.rand(2, c, device=dev)
    w = torch.randn(9, c, device=dev)
    stats = torch.zeros(2, c, device=dev)
    ho = (h - 1) // stride + 1
    return (lambda: ops.dw_fwd(x, coef, w, stride, stats, impl=impl)), B * (h * h + ho * ho) * c * 2


def dw_bwd(h, c, stride, impl="simt"):
    ho = (h - 1) // stride + 1
    e, dzd, d = bf(B, h, h, c), bf(B, ho, ho, c), bf(B, ho, ho, c)
    ecoef, dcoef = torch.rand(2, c, device=dev), torch.rand(3, c, device=dev)
    w = torch.randn(9, c, device=dev)
    stats, dw = torch.zeros(2, c, device=dev), torch.zeros(9, c, device=dev)
    return (lambda: ops.dw_bwd(e, ecoef, dzd, d, dcoef, w, stride, stats, dw, impl=impl)), B * (2 * h * h + 2 * ho * ho) * c * 2


cases = {
    "pw_fwd_expand_112_16_96": lambda: pw_fwd_expand(112, 16, 96),
    "pw_fwd_expand_56_32_144": lambda: pw_fwd_expand(56, 32, 144),
    "pw_fwd_expand_14_64_384": lambda: pw_fwd_expand(14, 64, 384),
    "pw_fwd_project_112_32_16": lambda: pw_fwd_project(112, 32, 16),
    "pw_fwd_project_56_144_32": lambda: pw_fwd_project(56, 144, 32),
    "pw_fwd_project_14_576_96": lambda: pw_fwd_project(14, 576, 96),
    "pw_dgrad_project_56_144_32": lambda: pw_dgrad_project(56, 144, 32),
    "pw_dgrad_project_14_576_96": lambda: pw_dgrad_project(14, 576, 96),
    "pw_dgrad_expand_112_16_96": lambda: pw_dgrad_expand(112, 16, 96),
    "pw_dgrad_expand_14_96_576": lambda: pw_dgrad_expand(14, 96, 576),
    "pw_wgrad_expand_112_16_96": lambda: pw_wgrad_expand(112, 16, 96),
    "pw_wgrad_expand_14_96_576": lambda: pw_wgrad_expand(14, 96, 576),
    "pw_wgrad_project_112_32_16": lambda: pw_wgrad_project(112, 32, 16),
    "pw_wgrad_project_56_144_32": lambda: pw_wgrad_project(56, 144, 32),
    "pw_wgrad_project_14_576_96": lambda: pw_wgrad_project(14, 576, 96),
    "pw_fwd_expand_7_160_960": lambda: pw_fwd_expand(7, 160, 960),
    "pw_fwd_project_7_960_160": lambda: pw_fwd_project(7, 960, 160),
    "pw_fwd_expand_14_96_576": lambda: pw_fwd_expand(14, 96, 576),
    "dw_fwd_56_144_s1": lambda: dw_fwd(56, 144, 1),
    "dw_fwd_112_96_s2": lambda: dw_fwd(112, 96, 2),
    "dw_fwd_14_384_s1": lambda: dw_fwd(14, 384, 1),
    "dw_fwd_simt_112_32_s1": lambda: dw_fwd(112, 32, 1),
    "dw_fwd_simt_56_144_s2": lambda: dw_fwd(56, 144, 2),
    "dw_fwd_simt_28_192_s1": lambda: dw_fwd(28, 192, 1),
    "dw_fwd_simt_28_192_s2": lambda: dw_fwd(28, 192, 2),
    "dw_fwd_simt_14_576_s1": lambda: dw_fwd(14, 576, 1),
    "dw_fwd_simt_14_576_s2": lambda: dw_fwd(14, 576, 2),
    "dw_fwd_simt_7_960_s1": lambda: dw_fwd(7, 960, 1),
    "dw_fwd_tile_112_32_s1": lambda: dw_fwd(112, 32, 1, "tile"),
    "dw_fwd_tile_56_144_s1": lambda: dw_fwd(56, 144, 1, "tile"),
    "dw_fwd_tile_28_192_s1": lambda: dw_fwd(28, 192, 1, "tile"),
    "dw_fwd_tile_14_384_s1": lambda: dw_fwd(14, 384, 1, "tile"),
    "dw_fwd_tile_14_576_s1": lambda: dw_fwd(14, 576, 1, "tile"),
    "dw_fwd_tile_7_960_s1": lambda: dw_fwd(7, 960, 1, "tile"),
    "dw_fwd_s2_simt_112_96": lambda: dw_fwd(112, 96, 2, "simt"),
    "dw_fwd_s2_simt_56_144": lambda: dw_fwd(56, 144, 2, "simt"),
    "dw_fwd_s2_simt_28_192": lambda: dw_fwd(28, 192, 2, "simt"),
    "dw_fwd_s2_simt_14_576": lambda: dw_fwd(14, 576, 2, "simt"),
    "dw_fwd_s2_tile_112_96": lambda: dw_fwd(112, 96, 2, "tile"),
    "dw_fwd_s2_tile_56_144": lambda: dw_fwd(56, 144, 2, "tile"),
    "dw_fwd_s2_tile_28_192": lambda: dw_fwd(28, 192, 2, "tile"),
    "dw_fwd_s2_tile_14_576": lambda: dw_fwd(14, 576, 2, "tile"),
    "dw_bwd_s2_simt_112_96": lambda: dw_bwd(112, 96, 2, "simt"),
    "dw_bwd_s2_simt_56_144": lambda: dw_bwd(56, 144, 2, "simt"),
    "dw_bwd_s2_simt_28_192": lambda: dw_bwd(28, 192, 2, "simt"),
    "dw_bwd_s2_simt_14_576": lambda: dw_bwd(14, 576, 2, "simt"),
    "dw_bwd_s2_tile_112_96": lambda: dw_bwd(112, 96, 2, "tile"),
    "dw_bwd_s2_tile_56_144": lambda: dw_bwd(56, 144, 2, "tile"),
    "dw_bwd_s2_tile_28_192": lambda: dw_bwd(28, 192, 2, "tile"),
    "dw_bwd_s2_tile_14_576": lambda: dw_bwd(14, 576, 2, "tile"),
    "dw_bwd_56_144_s1": lambda: dw_bwd(56, 144, 1),
    "dw_bwd_112_96_s2": lambda: dw_bwd(112, 96, 2),
    "dw_bwd_14_384_s1": lambda: dw_bwd(14, 384, 1),
    "dw_bwd_simt_112_32_s1": lambda: dw_bwd(112, 32, 1, "simt"),
    "dw_bwd_simt_56_144_s1": lambda: dw_bwd(56, 144, 1, "simt"),
    "dw_bwd_simt_28_192_s1": lambda: dw_bwd(28, 192, 1, "simt"),
    "dw_bwd_simt_14_384_s1": lambda: dw_bwd(14, 384, 1, "simt"),
    "dw_bwd_simt_14_576_s1": lambda: dw_bwd(14, 576, 1, "simt"),
    "dw_bwd_simt_7_960_s1": lambda: dw_bwd(7, 960, 1, "simt"),
    "dw_bwd_tile_112_32_s1": lambda: dw_bwd(112, 32, 1, "tile"),
    "dw_bwd_tile_56_144_s1": lambda: dw_bwd(56, 144, 1, "tile"),
    "dw_bwd_tile_28_192_s1": lambda: dw_bwd(28, 192, 1, "tile"),
    "dw_bwd_tile_14_384_s1": lambda: dw_bwd(14, 384, 1, "tile"),
    "dw_bwd_tile_14_576_s1": lambda: dw_bwd(14, 576, 1, "tile"),
    "dw_bwd_tile_7_960_s1": lambda: dw_bwd(7, 960, 1, "tile"),
}
for name, mk in cases.items():
    if ns.only and ns.only not in name:
        continue
    fn, nbytes = mk()
    ms = timeit(fn)
    print(f"{name:32s} {ms * 1e3:8.1f} us  {nbytes / 1e6:8.1f} MB  {nbytes / ms / 1e6:7.1f} GB/s  {nbytes / (ms * 1e-3) / HBM * 100:5.1f}% of measured HBM copy bw")
    del fn
    torch.cuda.empty_cache()
