"""b200dp_dw_fwd_bn_noalloc: depthwise conv + BatchNorm coefficients without anyone allocating scratch.

Who needs it: code that may not allocate device memory at the point of the call -- a plug-in inside another
runtime's enqueue hook, a serving loop with a fixed memory budget, a latency-critical path that must not enter
the allocator, a process whose free memory is all owned by a framework allocator.  The regular two-call form
(b200dp_dw_fwd + b200dp_bn_fwd_finalize) makes such a caller allocate and zero a per-channel sums buffer; this
entry point keeps those sums in a static 8 KB device array inside the library instead.

The test talks to the library through its C ABI (ctypes, raw pointers), calls the new entry point twice in a row
with different shapes -- so the second call sees the first call's leftovers in the shared scratch -- and checks
both results against the two-call form.  The file name sorts last so that, under `pytest -x`, every other test has
already run when this one is reached; it depends on no other test and also passes in a fresh process.
"""
import ctypes as C
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "pg---diploma-project---distributed-ai-model-training-using-mpi-and-accelerated-gpu-_b200",
                   "lib", "libb200dp.so")
MAX_C = 1024


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(LIB):
        pytest.fail(f"{LIB} missing -- run python build_native.py")
    L = C.CDLL(LIB)
    vp, i, f = C.c_void_p, C.c_int, C.c_float
    L.b200dp_last_error.restype = C.c_char_p
    L.b200dp_dw_fwd_bn_noalloc.argtypes = [vp, vp, vp, vp, vp, vp, f, vp, vp, vp, i, i, i, i, i, vp]
    L.b200dp_dw_fwd_bn_noalloc.restype = i
    L.b200dp_dw_fwd.argtypes = [vp, vp, vp, vp, vp, vp, i, i, i, i, i, i, vp]
    L.b200dp_dw_fwd.restype = i
    L.b200dp_bn_fwd_finalize.argtypes = [vp, i, vp, vp, f, f, f, vp, vp, vp, vp, i, vp]
    L.b200dp_bn_fwd_finalize.restype = i
    return L


def _case(B, H, Cn, stride, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    r = lambda *s: torch.randn(*s, device="cuda", generator=g)
    x = (r(B, H, H, Cn) * 2).to(torch.bfloat16).contiguous()
    in_coef = torch.stack([torch.rand(Cn, device="cuda", generator=g) + 0.5, r(Cn)]).contiguous()
    w = (r(9, Cn) * 0.3).contiguous()
    gamma = (torch.rand(Cn, device="cuda", generator=g) + 0.5).contiguous()
    beta = (r(Cn) * 0.1).contiguous()
    return x, in_coef, w, gamma, beta


def _two_call_form(lib, x, in_coef, w, gamma, beta, stride, stream):
    B, H, _, Cn = x.shape
    Ho = (H - 1) // stride + 1
    y = torch.empty(B, Ho, Ho, Cn, dtype=torch.bfloat16, device="cuda")
    sums = torch.zeros(2, Cn, device="cuda")                       # the scratch the caller has to own
    coef, saved = torch.empty(2, Cn, device="cuda"), torch.empty(2, Cn, device="cuda")
    rc = lib.b200dp_dw_fwd(x.data_ptr(), in_coef.data_ptr(), in_coef.data_ptr() + 4 * Cn, w.data_ptr(), y.data_ptr(),
                           sums.data_ptr(), Cn, B, H, H, Cn, stride, stream)
    assert rc == 0, lib.b200dp_last_error()
    rc = lib.b200dp_bn_fwd_finalize(sums.data_ptr(), Cn, gamma.data_ptr(), beta.data_ptr(), float(B * Ho * Ho), 1e-5,
                                    0.0, None, None, coef.data_ptr(), saved.data_ptr(), Cn, stream)
    assert rc == 0, lib.b200dp_last_error()
    return y, coef, saved


def _no_alloc_form(lib, x, in_coef, w, gamma, beta, stride, stream):
    B, H, _, Cn = x.shape
    Ho = (H - 1) // stride + 1
    y = torch.empty(B, Ho, Ho, Cn, dtype=torch.bfloat16, device="cuda")          # outputs are still the caller's
    coef, saved = torch.empty(2, Cn, device="cuda"), torch.empty(2, Cn, device="cuda")
    rc = lib.b200dp_dw_fwd_bn_noalloc(x.data_ptr(), in_coef.data_ptr(), in_coef.data_ptr() + 4 * Cn, w.data_ptr(),
                                      gamma.data_ptr(), beta.data_ptr(), 1e-5, y.data_ptr(), coef.data_ptr(),
                                      saved.data_ptr(), B, H, H, Cn, stride, stream)
    assert rc == 0, lib.b200dp_last_error()
    return y, coef, saved


def _torch_oracle(x, in_coef, w, gamma, beta, stride, eps=1e-5):
    """fp32 PyTorch statement of the op: y = dwconv3x3(relu6(scale * x + shift)); BatchNorm coefficients of y."""
    Cn = x.shape[-1]
    a = torch.clamp(x.float() * in_coef[0] + in_coef[1], 0, 6).to(torch.bfloat16).float()     # the kernel rounds the activation
    wt = w.t().reshape(Cn, 1, 3, 3)
    y = torch.nn.functional.conv2d(a.permute(0, 3, 1, 2), wt, stride=stride, padding=1, groups=Cn)
    yq = y.to(torch.bfloat16).float()            # statistics are taken over the values that are stored
    mean = yq.mean((0, 2, 3))
    var = yq.var((0, 2, 3), unbiased=False)
    inv = torch.rsqrt(var + eps)
    scale = gamma * inv
    return y.permute(0, 2, 3, 1), torch.stack([scale, beta - mean * scale]), torch.stack([mean, inv])


def test_two_calls_in_a_row_match_the_two_call_form(lib):
    stream = torch.cuda.current_stream().cuda_stream
    # first call: wide tensor; second call: fewer channels and another stride, so stale sums from call one sit in
    # the scratch beyond and inside the range call two uses
    for B, H, Cn, stride, seed in ((4, 28, 192, 1, 1), (3, 30, 96, 2, 2)):
        args = _case(B, H, Cn, stride, seed)
        y1, coef1, saved1 = _no_alloc_form(lib, *args, stride, stream)
        y0, coef0, saved0 = _two_call_form(lib, *args, stride, stream)
        torch.cuda.synchronize()
        assert torch.equal(y1, y0)                                   # same kernel, same inputs
        assert torch.allclose(coef1, coef0, rtol=2e-3, atol=1e-3)    # sums are atomically accumulated: order differs
        assert torch.allclose(saved1, saved0, rtol=2e-3, atol=1e-3)
        # and against a plain fp32 PyTorch statement of the same op (the two forms share a kernel, so agreeing with
        # each other alone would not catch an error in it)
        y_ref, coef_ref, saved_ref = _torch_oracle(*args, stride)
        assert torch.allclose(y1.float(), y_ref, rtol=2e-2, atol=3e-2)
        assert torch.allclose(saved1, saved_ref, rtol=5e-3, atol=5e-3)
        assert torch.allclose(coef1, coef_ref, rtol=5e-3, atol=5e-3)


def test_rejects_sizes_beyond_the_static_scratch(lib):
    Cn = MAX_C + 16
    x, in_coef, w, gamma, beta = _case(1, 4, Cn, 1, 3)
    y = torch.empty(1, 4, 4, Cn, dtype=torch.bfloat16, device="cuda")
    coef, saved = torch.empty(2, Cn, device="cuda"), torch.empty(2, Cn, device="cuda")
    rc = lib.b200dp_dw_fwd_bn_noalloc(x.data_ptr(), in_coef.data_ptr(), in_coef.data_ptr() + 4 * Cn, w.data_ptr(),
                                      gamma.data_ptr(), beta.data_ptr(), 1e-5, y.data_ptr(), coef.data_ptr(),
                                      saved.data_ptr(), 1, 4, 4, Cn, 1, torch.cuda.current_stream().cuda_stream)
    assert rc == -1
    msg = lib.b200dp_last_error().decode()
    assert "exceeds the static scratch" in msg and str(MAX_C) in msg
