"""Native engine: the whole train / eval step runs in hand-written sm_100a kernels driven by the
C++ executor (csrc/engine.cu), and gradient sync + Adam is the fused peer-memory kernel.

What the reference does per step (cifar10_mpi_mobilenet_224.py:173-184) and what happens here:

* ``imgs.to(device)`` 77 MB fp32                   -> 393 KB of uint8 (+labels, +12 floats/sample)
* ``zero_grad``                                    -> fused into the optimiser kernel
* ``ddp_model(imgs)`` (~160 cuDNN/ATen kernels)    -> ``b200dp_engine_train_step`` (one CUDA-graph launch:
  ``criterion`` / ``backward`` (~500 kernels)         preprocess, stem, tcgen05 pointwise GEMMs, depthwise
                                                      stencils, head, and the mirror-image backward)
* DDP bucket all-reduce + ``optimizer.step``       -> ``FusedAllReduceAdam.step`` (one kernel)
* ``loss.item()`` / ``.sum().item()``              -> device-side double accumulators, read once per epoch

Parameters live in one flat fp32 arena (``models/arena.py``); ``state_dict`` / ``load_state_dict``
convert to / from the torchvision layout so checkpoints interchange with the reference.
"""
from __future__ import annotations

import ctypes as C
from collections import OrderedDict
from typing import Dict, List, Optional, Tuple

import torch

from .models.arena import ArenaLayout
from .models.mobilenetv2 import build_model
from .models.spec import BN_EPS, BN_MOMENTUM, build_spec
from .ops import _lib
from .ops._lib import check
from .optim import AdamHyper


def _declare(lib: C.CDLL) -> None:
    vp, i, ll = C.c_void_p, C.c_int, C.c_longlong
    lib.b200dp_engine_create.argtypes = [C.POINTER(vp), i, i, i, C.POINTER(ll), i, ll, ll, vp, vp, vp,
                                         C.c_float, C.c_float, C.c_float, C.c_uint, i]
    lib.b200dp_engine_create.restype = i
    lib.b200dp_engine_destroy.argtypes = [vp]
    lib.b200dp_engine_destroy.restype = i
    lib.b200dp_engine_workspace_bytes.argtypes = [vp]
    lib.b200dp_engine_workspace_bytes.restype = C.c_size_t
    lib.b200dp_engine_buffer.argtypes = [vp, i]
    lib.b200dp_engine_buffer.restype = vp
    lib.b200dp_engine_tap.argtypes = [vp, i, i]
    lib.b200dp_engine_tap.restype = vp
    lib.b200dp_engine_train_step.argtypes = [vp, vp, vp, vp, i, vp]
    lib.b200dp_engine_train_step.restype = i
    lib.b200dp_engine_eval_step.argtypes = [vp, vp, vp, i, vp]
    lib.b200dp_engine_eval_step.restype = i
    lib.b200dp_engine_invalidate.argtypes = [vp]
    lib.b200dp_engine_invalidate.restype = i
    lib.b200dp_engine_profile_step.argtypes = [vp, vp, vp, vp, i, C.c_char_p, i, C.POINTER(C.c_float), i,
                                               C.POINTER(i), vp]
    lib.b200dp_engine_profile_step.restype = i
    lib.b200dp_engine_forward_preprocessed.argtypes = [vp, vp, i, vp]
    lib.b200dp_engine_forward_preprocessed.restype = i
    lib.b200dp_engine_set_comm.argtypes = [vp, vp, C.POINTER(i), i, i]
    lib.b200dp_engine_set_comm.restype = i


_lib.register_decls(_declare)


def available() -> bool:
    if not torch.cuda.is_available():
        return False
    lib = _lib.load()
    return lib is not None and hasattr(lib, "b200dp_engine_create")


class _DevView:
    """torch view of raw device memory owned by the native engine."""

    def __init__(self, owner, ptr: int, numel: int, typestr: str) -> None:
        self._owner = owner
        self.__cuda_array_interface__ = {"shape": (numel,), "typestr": typestr, "data": (ptr, False),
                                         "version": 2, "strides": None}


class NativeCore:
    """Owns the parameter / buffer arenas and the C++ executor; shared by training and inference."""

    def __init__(self, device: torch.device, batch: int, img_size: int, num_classes: int, dropout: float,
                 seed: int, grads: Optional[torch.Tensor] = None, use_graphs: bool = True,
                 params: Optional[torch.Tensor] = None, buffers: Optional[torch.Tensor] = None) -> None:
        self.lib = _lib.require()
        self.device = device
        self.batch = batch
        self.layout = ArenaLayout(build_spec(img_size, num_classes))
        L = self.layout
        # `params` / `grads` may live in a peer-mapped comm arena (two-shot optimiser: peers store the updated slices)
        self.params = params if params is not None else torch.zeros(L.total, dtype=torch.float32, device=device)
        self.buffers = buffers if buffers is not None else torch.zeros(L.buffer_total, dtype=torch.float32, device=device)
        self.grads = grads if grads is not None else torch.zeros(L.total, dtype=torch.float32, device=device)
        self.num_batches_tracked = 0
        table = L.offsets_table()
        arr = (C.c_longlong * len(table))(*table)
        h = C.c_void_p()
        check(self.lib.b200dp_engine_create(C.byref(h), batch, img_size, num_classes, arr, len(table), L.total,
                                            L.buffer_total, self.params.data_ptr(), self.grads.data_ptr(),
                                            self.buffers.data_ptr(), dropout, BN_EPS, BN_MOMENTUM, seed & 0xFFFFFFFF,
                                            1 if use_graphs else 0))
        self.h = h
        self.num_classes = num_classes
        self.metrics_train = self._view(4, 3, "<f8")
        self.metrics_eval = self._view(5, 3, "<f8")
        self.logits = self._view(3, batch * num_classes, "<f4").view(batch, num_classes)

    def _view(self, which: int, numel: int, typestr: str) -> torch.Tensor:
        p = self.lib.b200dp_engine_buffer(self.h, which)
        return torch.as_tensor(_DevView(self, int(p), numel, typestr), device=self.device)

    def tap(self, what: int, index: int, numel: int, typestr: str = "<u2") -> torch.Tensor:
        p = self.lib.b200dp_engine_tap(self.h, what, index)
        t = torch.as_tensor(_DevView(self, int(p), numel, typestr), device=self.device)
        return t.view(torch.bfloat16) if typestr == "<u2" else t

    def load_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        p, b, nbt = self.layout.import_state_dict(sd)
        self.params.copy_(p)
        self.buffers.copy_(b)
        self.num_batches_tracked = nbt
        self.lib.b200dp_engine_invalidate(self.h)

    def state_dict(self) -> "OrderedDict[str, torch.Tensor]":
        return self.layout.export_state_dict(self.params, self.buffers, self.num_batches_tracked)

    def train_step(self, images: torch.Tensor, labels: torch.Tensor, aug: Optional[torch.Tensor], n: int) -> None:
        check(self.lib.b200dp_engine_train_step(self.h, images.data_ptr(), labels.data_ptr(),
                                                aug.data_ptr() if aug is not None else None, n,
                                                _lib.current_stream()))
        self.num_batches_tracked += 1

    def eval_step(self, images: torch.Tensor, labels: torch.Tensor, n: int) -> None:
        check(self.lib.b200dp_engine_eval_step(self.h, images.data_ptr(), labels.data_ptr(), n, _lib.current_stream()))

    def forward_preprocessed(self, x: torch.Tensor, n: int) -> None:
        """Eval forward of normalised bf16 NHWC input [n, S, S, 4]; logits land in ``self.logits``."""
        check(self.lib.b200dp_engine_forward_preprocessed(self.h, x.data_ptr(), n, _lib.current_stream()))

    def set_comm(self, fused, bucket_first: List[int], comm_blocks: int) -> None:
        """Hand the fused all-reduce + Adam to the executor: one launch per gradient bucket inside the step graph."""
        arr = (C.c_int * max(len(bucket_first), 1))(*bucket_first)
        check(self.lib.b200dp_engine_set_comm(self.h, C.byref(fused.args) if fused is not None else None, arr,
                                              len(bucket_first), comm_blocks))

    def profile_step(self, images, labels, aug, n) -> List[Tuple[str, float]]:
        names = C.create_string_buffer(1 << 16)
        ms = (C.c_float * 1024)()
        cnt = C.c_int()
        check(self.lib.b200dp_engine_profile_step(self.h, images.data_ptr(), labels.data_ptr(),
                                                  aug.data_ptr() if aug is not None else None, n, names, 1 << 16,
                                                  ms, 1024, C.byref(cnt), _lib.current_stream()))
        labels_ = names.value.decode().strip().split("\n")
        return list(zip(labels_, [ms[i] for i in range(cnt.value)]))

    def close(self) -> None:
        if self.h:
            self.lib.b200dp_engine_destroy(self.h)
            self.h = None


class NativeEngine:
    name = "native"

    def __init__(self, cfg, ctx) -> None:
        from .parallel.comm import FusedAllReduceAdam
        self.cfg, self.ctx = cfg, ctx
        self.device = ctx.device
        torch.manual_seed(cfg.seed)
        layout = ArenaLayout(build_spec(cfg.img_size, cfg.num_classes))
        hyper = AdamHyper(cfg.lr, cfg.beta1, cfg.beta2, cfg.eps, cfg.weight_decay)
        self.nccl = cfg.comm == "nccl"
        # gradients, master weights and BatchNorm running statistics live in the symmetric comm arena: the executor
        # accumulates gradients straight into it, peers read them there and store updated weights / statistics
        self.fused = FusedAllReduceAdam(ctx, layout.total, hyper, algo=cfg.comm_algo, extra_numel=layout.buffer_total)
        self.comm_algo = self.fused.algo_name
        self.core = NativeCore(self.device, cfg.resolved_batch_size(), cfg.img_size, cfg.num_classes, cfg.dropout,
                               cfg.seed + 7919 * ctx.rank, grads=self.fused.grads, params=self.fused.params,
                               buffers=self.fused.extra)
        init = build_model(cfg.num_classes, cfg.dropout, cfg.img_size, cfg.pretrained)
        self.core.load_state_dict(init.state_dict())
        del init
        self.broadcast_parameters()
        self.steps = 0
        self.in_graph = (not self.nccl) and cfg.comm_overlap
        if self.in_graph:
            starts = [int(x) for x in cfg.comm_buckets.split(",") if x.strip()]
            self.comm_buckets = starts
            self.core.set_comm(self.fused, starts, cfg.comm_blocks)
        if self.nccl:
            from .optim import FlatAdam
            self.adam = FlatAdam(layout.total, self.device, hyper)

    # ---- replica consistency ---------------------------------------------------------------------
    def broadcast_parameters(self) -> None:
        if self.ctx.initialized:
            torch.cuda.current_stream().synchronize()
            self.fused.broadcast_params(0)              # K-C3: DDP's constructor broadcast, as peer / multicast stores
            self.sync_buffers()

    def sync_buffers(self) -> None:
        if self.ctx.initialized:
            self.fused.broadcast_extra(0)               # K-C4: rank 0's running statistics are authoritative
            self.core.lib.b200dp_engine_invalidate(self.core.h)

    def reduce(self, values, op: str = "sum"):
        """Sum / max of a few doubles over ranks on the peer slot tables (K-C6 / K-C7, max-over-ranks timers)."""
        t = torch.as_tensor(list(values), dtype=torch.float64, device=self.device)
        return [float(x) for x in self.fused.reduce_f64(t, op).tolist()]

    # ---- steps -------------------------------------------------------------------------------------
    def train_step(self, batch, lr: float) -> None:
        batch.wait()
        aug = batch.params if self.cfg.augment else None
        if self.in_graph:     # the step graph contains the bucketed all-reduce + Adam; only the learning rate is fed in
            self.fused.set_lr(lr)
            self.core.train_step(batch.images, batch.labels, aug, batch.n)
            self.fused.poll_error()
            self.steps += 1
            return
        self.core.train_step(batch.images, batch.labels, aug, batch.n)
        if self.nccl:     # comparison baseline: NCCL all-reduce + separate Adam
            import torch.distributed as dist
            if self.ctx.initialized:
                dist.all_reduce(self.core.grads)
            self.adam.step_torch(self.core.params, self.core.grads, lr, grad_scale=1.0 / self.ctx.world_size)
            self.core.grads.zero_()
        else:
            self.fused.step(lr)
            self.fused.poll_error()
        self.steps += 1

    def eval_step(self, batch) -> None:
        batch.wait()
        self.core.eval_step(batch.images, batch.labels, batch.n)

    @torch.no_grad()
    def predict_logits(self, images_u8: torch.Tensor) -> torch.Tensor:
        n = images_u8.shape[0]
        labels = torch.zeros(n, dtype=torch.int64, device=self.device)
        keep = self.core.metrics_eval.clone()
        self.core.eval_step(images_u8.to(self.device).contiguous(), labels, n)
        out = self.core.logits[:n].clone()
        self.core.metrics_eval.copy_(keep)
        return out

    # ---- metrics / state -------------------------------------------------------------------------
    def read_metrics(self, reset: bool = True) -> Tuple[float, float, float]:
        m = (self.core.metrics_train + self.core.metrics_eval).tolist()
        self.fused.poll_error()           # the read above synchronised: a timed-out barrier is visible by now
        if reset:
            self.core.metrics_train.zero_()
            self.core.metrics_eval.zero_()
        return m[0], m[1], m[2]

    # Per-step metric read-back without stalling the pipeline: `post_metrics()` enqueues a device -> pinned-host copy of
    # the running (loss sum, correct, count) after the step just launched, `fetch_metrics()` waits for the OLDEST posted
    # copy and returns it.  A loop that posts every step and fetches one step behind reads every step's result while
    # the host stays one step ahead of the GPU (the reference's `loss.item()` at ref :183 stalls both every step).
    def post_metrics(self) -> None:
        if not hasattr(self, "_mring"):
            self._mring = torch.zeros(8, 3, dtype=torch.float64).pin_memory()
            self._mev = [torch.cuda.Event() for _ in range(8)]
            self._mhead = self._mtail = 0
        slot = self._mhead % 8
        self._mring[slot].copy_(self.core.metrics_train[:3], non_blocking=True)
        self._mev[slot].record()
        self._mhead += 1

    def fetch_metrics(self) -> Tuple[float, float, float]:
        if not hasattr(self, "_mring") or self._mtail == self._mhead:
            raise RuntimeError("fetch_metrics without a pending post_metrics")
        slot = self._mtail % 8
        self._mev[slot].synchronize()
        self._mtail += 1
        self.fused.poll_error()
        v = self._mring[slot].tolist()
        return v[0], v[1], v[2]

    def state_dict(self):
        return self.core.state_dict()

    def load_state_dict(self, sd) -> None:
        self.core.load_state_dict(sd)

    def optimizer_state(self):
        return self.adam.state_dict() if self.nccl else self.fused.state_dict()

    def load_optimizer_state(self, sd) -> None:
        (self.adam if self.nccl else self.fused).load_state_dict(sd)

    def parameter_count(self) -> int:
        return int(self.core.layout.trainable_mask().sum().item())

    def param_hash(self) -> float:
        return float(self.core.params.double().sum().item())

    def close(self) -> None:
        torch.cuda.synchronize()
        try:
            self.fused.check_error()          # a timed-out barrier must not go unnoticed, but must not leak the arena either
        finally:
            self.core.close()
            self.core.params = self.core.grads = self.core.buffers = None     # views of the arena freed next
            self.fused.close()


class NativeInference:
    """Eval-mode forward on the native kernels (BatchNorm folded into scale/shift on load).

    Two inputs: ``logits(uint8 [B,32,32,3])`` -- CIFAR-sized images through the fused resize / im2col kernel, the
    evaluation path of training -- and ``logits_preprocessed(bf16 [B,224,224,4])`` for photographs that were resized
    from their ORIGINAL resolution by ``ops.resize_normalize`` (what the reference's predict helper and Gradio app do,
    cifar10_128batch.py:157-165, GROUP03.pdf p.26-27).  Both replay a CUDA graph per batch size."""

    def __init__(self, num_classes: int, device: torch.device, state_dict: Optional[dict], batch: int = 256) -> None:
        if not available():
            raise RuntimeError("native engine unavailable")
        self.device = device
        self.core = NativeCore(device, batch, 224, num_classes, 0.0, 0)
        if state_dict is None:
            state_dict = build_model(num_classes).state_dict()
        self.core.load_state_dict(state_dict)
        self.batch = batch
        self._labels = torch.zeros(batch, dtype=torch.int64, device=device)

    @torch.no_grad()
    def logits(self, images_u8: torch.Tensor) -> torch.Tensor:
        outs = []
        for s in range(0, images_u8.shape[0], self.batch):
            chunk = images_u8[s: s + self.batch].contiguous()
            n = chunk.shape[0]
            self.core.eval_step(chunk, self._labels, n)
            outs.append(self.core.logits[:n].clone())
        return torch.cat(outs)

    @torch.no_grad()
    def logits_preprocessed(self, x: torch.Tensor, clone: bool = True) -> torch.Tensor:
        """x: normalised bf16 NHWC [B,224,224,4] on the device."""
        outs = []
        for s in range(0, x.shape[0], self.batch):
            chunk = x[s: s + self.batch].contiguous()
            n = chunk.shape[0]
            self.core.forward_preprocessed(chunk, n)
            outs.append(self.core.logits[:n].clone() if clone else self.core.logits[:n])
        return outs[0] if len(outs) == 1 else torch.cat(outs)

    def close(self) -> None:
        self.core.close()
