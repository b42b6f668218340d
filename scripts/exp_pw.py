import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200dp import ops
dev = "cuda"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
bf = lambda *s: torch.randn(*s, device=dev).to(torch.bfloat16)
def timeit(fn, iters=8):
    for _ in range(3): fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    ts.sort(); return ts[len(ts)//2] * 1e3
M = 128 * 112 * 112
for K, N in ((16, 96), (64, 96), (16, 32), (16, 128), (64, 128), (128, 128)):
    a, w = bf(M, K), bf(N, K)
    out = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
    stats = torch.zeros(2, N, device=dev)
    t_plain = timeit(lambda: ops.pw_gemm(w, a=a, out=out))
    t_stats = timeit(lambda: ops.pw_gemm(w, a=a, emode=ops.E_STATS, stats=stats, out=out))
    by = M * (K + N) * 2
    print(f"K={K:4d} N={N:4d}  plain {t_plain:7.1f} us ({by/t_plain/1e3:6.0f} GB/s)   stats {t_stats:7.1f} us ({by/t_stats/1e3:6.0f} GB/s)")
    del a, w, out
print("--- A via register path (identity BN) vs TMA, narrow K")
for K, N in ((16, 96), (32, 144), (32, 192)):
    hh = 112 if K == 16 else 56
    M2 = 128 * hh * hh
    a, w = bf(M2, K), bf(N, K)
    out = torch.empty(M2, N, dtype=torch.bfloat16, device=dev)
    stats = torch.zeros(2, N, device=dev)
    ident = torch.stack([torch.ones(K, device=dev), torch.zeros(K, device=dev)]).contiguous()
    t_tma = timeit(lambda: ops.pw_gemm(w, a=a, emode=ops.E_STATS, stats=stats, out=out))
    t_reg = timeit(lambda: ops.pw_gemm(w, a_raw=a, acoef=ident, relu6=False, emode=ops.E_STATS, stats=stats, out=out))
    by = M2 * (K + N) * 2
    print(f"K={K:4d} N={N:4d} M={M2}  tma {t_tma:7.1f} us ({by/t_tma/1e3:6.0f} GB/s)   reg {t_reg:7.1f} us ({by/t_reg/1e3:6.0f} GB/s)")
