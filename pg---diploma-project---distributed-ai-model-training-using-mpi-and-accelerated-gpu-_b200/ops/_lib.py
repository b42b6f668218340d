"""ctypes binding of ``lib/libb200dp.so`` (C ABI in ``include/b200dp.h``).

The library is built in-tree by ``build_native.py`` and loaded from the package directory, never
from site-packages.  On a machine with a GPU a missing library is an error (``require()``), not a
silent fallback to ATen.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

MAX_WORLD = 16
COMM_MAX_BLOCKS = 256
REDUCE_MAX = 32
AUG_NPARAM = 12

_PKG_DIR = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.environ.get("B200DP_LIB") or os.path.join(_PKG_DIR, "lib", "libb200dp.so")   # override: A/B runs of two builds

_lib: Optional[C.CDLL] = None
_load_error: Optional[str] = None


class CommAdamArgs(C.Structure):
    _fields_ = [
        ("world", C.c_int), ("rank", C.c_int), ("numel", C.c_longlong),
        ("grad_dtype", C.c_int), ("algo", C.c_int),
        ("grads", C.c_void_p * MAX_WORLD),
        ("flags", C.c_void_p * MAX_WORLD),
        ("params", C.c_void_p * MAX_WORLD),
        ("work_bf16", C.c_void_p * MAX_WORLD),
        ("m", C.c_void_p), ("v", C.c_void_p), ("lr", C.c_void_p),
        ("step_state", C.c_void_p), ("epoch_state", C.c_void_p), ("error_flag", C.c_void_p),
        ("beta1", C.c_double), ("beta2", C.c_double), ("eps", C.c_float), ("weight_decay", C.c_float),
        ("zero_grads", C.c_int), ("apply_adam", C.c_int),
        ("out_avg", C.c_void_p),
        ("spin_limit", C.c_ulonglong),
        ("max_blocks", C.c_int),
        ("range_lo", C.c_longlong), ("range_hi", C.c_longlong),
        ("bump_step", C.c_int), ("skip_barriers", C.c_int), ("exact_blocks", C.c_int),
        ("zero_lo", C.c_longlong), ("zero_hi", C.c_longlong),
        ("mc_grads", C.c_void_p), ("mc_params", C.c_void_p),
        ("error_host", C.c_void_p),
        ("slots", C.c_void_p * MAX_WORLD),
        ("abort_words", C.c_void_p * MAX_WORLD),
    ]


def _declare(lib: C.CDLL) -> None:
    vp, i, ll, f = C.c_void_p, C.c_int, C.c_longlong, C.c_float
    lib.b200dp_last_error.restype = C.c_char_p
    lib.b200dp_version.restype = i
    lib.b200dp_launch_count.restype = C.c_ulonglong
    lib.b200dp_reset_launch_count.restype = None
    sigs = {
        "b200dp_arena_alloc": [C.c_size_t, C.POINTER(vp)],
        "b200dp_arena_free": [vp],
        "b200dp_ipc_export": [vp, C.c_char_p],
        "b200dp_ipc_open": [C.c_char_p, C.POINTER(vp)],
        "b200dp_ipc_close": [vp],
        "b200dp_enable_peer_access": [i],
        "b200dp_comm_adam_step": [C.POINTER(CommAdamArgs), vp],
        "b200dp_comm_adam_grid": [C.POINTER(CommAdamArgs)],
        "b200dp_comm_barrier": [C.POINTER(CommAdamArgs), vp],
        "b200dp_comm_broadcast": [C.POINTER(CommAdamArgs), i, C.POINTER(vp), vp, vp, ll, vp],
        "b200dp_comm_reduce_f64": [C.POINTER(CommAdamArgs), vp, vp, i, i, vp],
        "b200dp_vmm_query": [C.POINTER(i), C.POINTER(C.c_size_t)],
        "b200dp_vmm_alloc": [C.c_size_t, C.POINTER(vp), C.POINTER(i), C.POINTER(C.c_ulonglong)],
        "b200dp_vmm_import": [i, C.c_size_t, C.POINTER(vp)],
        "b200dp_vmm_free": [vp],
        "b200dp_mc_create": [i, C.c_size_t, C.POINTER(C.c_ulonglong), C.POINTER(i)],
        "b200dp_mc_import": [i, C.c_size_t, C.POINTER(C.c_ulonglong)],
        "b200dp_mc_add_device": [C.c_ulonglong],
        "b200dp_mc_bind": [C.c_ulonglong, C.c_ulonglong, C.c_size_t],
        "b200dp_mc_map": [C.c_ulonglong, C.c_size_t, C.POINTER(vp)],
        "b200dp_mc_release": [C.c_ulonglong],
        "b200dp_close_fd": [i],
        "b200dp_preprocess": [vp, vp, vp, i, i, i, vp],
        "b200dp_preprocess_col": [vp, vp, vp, vp, i, i, vp],
        "b200dp_preprocess_col_supported": [i],
        "b200dp_resize_normalize": [vp, i, i, i, vp, i, vp],
        "b200dp_softmax_xent": [vp, vp, vp, vp, vp, i, i, f, vp],
        "b200dp_softmax_topk": [vp, vp, vp, vp, i, i, i, vp],
        "b200dp_pw_gemm": [i, i, i, i, i, vp, i, vp, vp, vp, i, vp, vp, i, vp, vp, vp, i, i, vp],
        "b200dp_pw_gemm_packed": [i, i, i, i, i, vp, vp, vp, vp, i, vp, vp, vp, vp, vp, i, i, i, vp],
        "b200dp_pw_gemm_pick_pack": [i, i, i],
        "b200dp_pw_wgrad_packed": [i, i, i, i, vp, vp, vp, vp, vp, vp, i, i, vp],
        "b200dp_pw_wgrad": [i, i, i, i, vp, i, vp, vp, i, vp, vp, vp, i, i, vp],
        "b200dp_dw_fwd": [vp, vp, vp, vp, vp, vp, i, i, i, i, i, i, vp],
        "b200dp_dw_fwd_tile": [vp, vp, vp, vp, vp, vp, i, i, i, i, i, i, vp],
        "b200dp_dw_bwd": [vp, vp, vp, vp, vp, vp, vp, vp, vp, i, vp, i, i, i, i, i, vp],
        "b200dp_dw_bwd_tile": [vp, vp, vp, vp, vp, vp, vp, vp, vp, i, vp, i, i, i, i, i, vp],
        "b200dp_stem_fwd": [vp, vp, vp, vp, i, i, i, i, vp],
        "b200dp_stem_wgrad": [vp, vp, vp, vp, vp, i, i, i, vp],
        "b200dp_bn_fwd_finalize": [vp, i, vp, vp, f, f, f, vp, vp, vp, vp, i, vp],
        "b200dp_bn_bwd_finalize": [vp, i, vp, vp, f, vp, vp, vp, i, vp],
        "b200dp_affine": [vp, vp, vp, i, vp, ll, i, vp],
        "b200dp_block_out": [vp, vp, vp, vp, vp, ll, i, vp],
        "b200dp_grad_join_stats": [vp, vp, vp, vp, vp, i, ll, i, vp],
    }
    for name, args in sigs.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = i


def load() -> Optional[C.CDLL]:
    """Load the library once; returns None (and remembers why) if it is not built."""
    global _lib, _load_error
    if _lib is not None or _load_error is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        _load_error = f"{LIB_PATH} not found -- run `python build_native.py`"
        return None
    try:
        lib = C.CDLL(LIB_PATH)
        _declare(lib)
        for extra in _EXTRA_DECLS:
            extra(lib)
        _lib = lib
    except (OSError, AttributeError) as e:
        _load_error = f"cannot load {LIB_PATH}: {e}"
    return _lib


_EXTRA_DECLS = []   # other binding modules register their own declarations here


def register_decls(fn) -> None:
    """Register extra ctypes declarations; applied now if the library is already loaded."""
    _EXTRA_DECLS.append(fn)
    if _lib is not None:
        fn(_lib)


def load_error() -> Optional[str]:
    return _load_error


def require() -> C.CDLL:
    lib = load()
    if lib is None:
        raise RuntimeError(f"b200dp native library unavailable: {_load_error}")
    return lib


def check(rc: int) -> int:
    if rc < 0:
        msg = require().b200dp_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"b200dp native call failed ({rc}): {msg}")
    return rc


def ptr(t) -> int:
    """Device (or host) address of a torch tensor, 0 for None."""
    return 0 if t is None else t.data_ptr()


def current_stream() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream
