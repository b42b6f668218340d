"""Where does the fixed cost of a small pointwise-conv GEMM go?  CTA 0 stamps %globaltimer at its milestones."""
import os, sys, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200dp import ops
from b200dp.ops import _lib

lib = _lib.require()
lib.b200dp_debug_pw_trace.argtypes = [ctypes.c_void_p]
lib.b200dp_debug_pw_trace.restype = None
dev = torch.device("cuda", 0)
trace = torch.zeros(32, dtype=torch.int64, device=dev)
names = ["entry", "setup done", "first TMA issued", "first tile landed (MMA sees full)", "MMAs of tile 0 committed",
         "epilogue sees tmem_full", "epilogue tile 0 stored (TMA store issued)", "roles done (thread 0)",
         "teardown barrier passed", "TMEM released", "epilogue warp: tile loop done", "epilogue warp: sums flushed to smem",
         "epilogue warp: TMA stores complete", "warp 2: after stats atomics, before dealloc",
         "MMA thread: chunk 4 ready", "MMA thread: chunk 12 ready", "transform: raw chunk 4 landed", "transform: raw chunk 12 landed"]
for (B, h, cin, cout) in [(2, 7, 160, 960), (128, 7, 160, 960), (128, 14, 96, 576), (128, 28, 32, 192)]:
    M = B * h * h
    a = torch.randn(M, cin, device=dev).to(torch.bfloat16)
    w = torch.randn(cout, cin, device=dev).to(torch.bfloat16)
    stats = torch.zeros(2, cout, device=dev)
    out = torch.empty(M, cout, dtype=torch.bfloat16, device=dev)
    for _ in range(3):
        ops.pw_gemm(w, a=a, emode=ops.E_STATS, stats=stats, out=out)
    torch.cuda.synchronize()
    lib.b200dp_debug_pw_trace(trace.data_ptr())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); ops.pw_gemm(w, a=a, emode=ops.E_STATS, stats=stats, out=out); e1.record()
    torch.cuda.synchronize()
    lib.b200dp_debug_pw_trace(None)
    t = trace.cpu().tolist()
    print(f"--- M={M} K={cin} N={cout}: events {e0.elapsed_time(e1) * 1e3:.1f} us")
    for i in sorted(range(len(names)), key=lambda i: t[i]):
        print(f"  {names[i]:48s} +{(t[i] - t[0]) / 1e3:7.2f} us")


print("=== project conv forward (A = relu6(BN(raw)) rewritten in shared memory) ===")
for (B, h, cin, cout) in [(128, 7, 960, 160), (128, 14, 576, 96), (128, 14, 384, 64)]:
    M = B * h * h
    raw = torch.randn(M, cin, device=dev).to(torch.bfloat16)
    w = torch.randn(cout, cin, device=dev).to(torch.bfloat16)
    coef = torch.rand(2, cin, device=dev)
    stats = torch.zeros(2, cout, device=dev)
    out = torch.empty(M, cout, dtype=torch.bfloat16, device=dev)
    fn = lambda: ops.pw_gemm(w, a_raw=raw, acoef=coef, relu6=True, emode=ops.E_STATS, stats=stats, out=out)
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    trace.zero_()
    lib.b200dp_debug_pw_trace(trace.data_ptr())
    fn()
    torch.cuda.synchronize()
    lib.b200dp_debug_pw_trace(None)
    t = trace.cpu().tolist()
    print(f"--- M={M} K={cin} N={cout}")
    for i in sorted([i for i in range(len(names)) if t[i]], key=lambda i: t[i]):
        print(f"  {names[i]:48s} +{(t[i] - t[0]) / 1e3:7.2f} us")

print("=== weight gradient (pw_wgrad) ===")
wnames = ["entry", "setup done", "-", "first stage ready (MMA sees full)", "MMAs of item 0 committed",
          "epilogue sees tmem_full", "epilogue item 0 done (red.adds issued)", "epilogue loop done", "MMA loop done",
          "TMEM released"]
for (B, h, cin, cout, mode) in [(128, 7, 160, 960, 0), (128, 14, 96, 576, 0), (128, 14, 576, 96, 1), (128, 28, 192, 32, 1),
                                (128, 112, 32, 16, 1)]:
    M = B * h * h
    hidden, narrow = (cout, cin) if mode == 0 else (cin, cout)
    bf = lambda *s: torch.randn(*s, device=dev).to(torch.bfloat16)
    dw = torch.zeros(cout, cin, device=dev)
    if mode == 0:
        a, dz, y = bf(M, cin), bf(M, cout), bf(M, cout)
        coef = torch.rand(3, cout, device=dev)
        fn = lambda: ops.pw_wgrad(dw, a=a, dz=dz, y_raw=y, coef=coef)
    else:
        g, d = bf(M, cout), bf(M, cin)
        coef = torch.rand(2, cin, device=dev)
        fn = lambda: ops.pw_wgrad(dw, g=g, a_raw=d, coef=coef)
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    trace.zero_()
    lib.b200dp_debug_pw_trace(trace.data_ptr())
    fn()
    torch.cuda.synchronize()
    lib.b200dp_debug_pw_trace(None)
    t = trace.cpu().tolist()
    print(f"--- mode {mode} M={M} {cin}->{cout}")
    for i in sorted([i for i in range(len(wnames)) if t[i]], key=lambda i: t[i]):
        print(f"  {wnames[i]:48s} +{(t[i] - t[0]) / 1e3:7.2f} us")
