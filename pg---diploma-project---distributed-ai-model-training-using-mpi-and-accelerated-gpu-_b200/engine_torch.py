"""Torch engine: the model runs on ATen ops, everything around it is the framework's own.

Used for mode 1 (serial CPU, cifar10_serial_mobilenet_224.py), as the numerical oracle, and as
the on-GPU comparison baseline (SURVEY §5.8 "baseline path kept for comparison only").  Even
here the step is not the reference's loop (cifar10_mpi_mobilenet_224.py:173-185):

* parameters and gradients are views into flat arenas, so gradient sync + Adam is one pass over a
  contiguous range -- on the GPU that pass is the fused peer-memory all-reduce + Adam kernel
  (``comm='fused'``) or NCCL all-reduce + flat Adam (``comm='nccl'``, the baseline);
* loss sum / correct count / sample count accumulate in a device tensor; there is no per-step
  ``.item()`` (the reference does two host syncs per step, :183-184);
* input is raw uint8 32x32x3; resize / augmentation / normalisation run on the device.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Dict, Optional, Tuple

import torch
import torch.nn.functional as F

from .data.loader import Batch
from .data.transforms import apply_pipeline_torch
from .models.mobilenetv2 import MobileNetV2, build_model
from .optim import AdamHyper, FlatAdam


def flatten_parameters(model: torch.nn.Module, device) -> Tuple[torch.Tensor, torch.Tensor]:
    """Re-home every parameter (and its .grad) as a view into one flat fp32 vector each."""
    params = [p for p in model.parameters()]
    total = (sum(p.numel() for p in params) + 63) // 64 * 64   # 16-byte vector access in the fused kernel
    flat_p = torch.zeros(total, dtype=torch.float32, device=device)
    flat_g = torch.zeros(total, dtype=torch.float32, device=device)
    off = 0
    for p in params:
        n = p.numel()
        flat_p[off: off + n].copy_(p.detach().reshape(-1))
        p.data = flat_p[off: off + n].view(p.shape)
        p.grad = flat_g[off: off + n].view(p.shape)
        off += n
    return flat_p, flat_g


class TorchEngine:
    name = "torch"

    def __init__(self, cfg, ctx, model: Optional[MobileNetV2] = None) -> None:
        self.cfg = cfg
        self.ctx = ctx
        self.device = ctx.device
        self.cuda = self.device.type == "cuda"
        self.world = ctx.world_size
        torch.manual_seed(cfg.seed)                      # ref :58
        self.model = model or build_model(cfg.num_classes, cfg.dropout, cfg.img_size, cfg.pretrained)
        self.model.to(self.device)
        if self.cuda:
            self.model.to(memory_format=torch.channels_last)
        self.flat_p, self.flat_g = flatten_parameters(self.model, self.device)
        self.numel = self.flat_p.numel()
        hyper = AdamHyper(cfg.lr, cfg.beta1, cfg.beta2, cfg.eps, cfg.weight_decay)
        self.adam = FlatAdam(self.numel, self.device, hyper)
        self.metrics = torch.zeros(3, dtype=torch.float64, device=self.device)  # loss_sum, correct, count
        self.fused = None
        if self.cuda and cfg.comm == "fused":
            from .parallel.comm import FusedAllReduceAdam
            self.fused = FusedAllReduceAdam(ctx, self.numel, hyper, params=self.flat_p,
                                            m=self.adam.m, v=self.adam.v)
            # gradients and master weights must live in the peer-mapped arena: re-point the views
            self._rebind_grads(self.fused.grad_buffer(0))
            self._rebind_params(self.fused.params)
        self.broadcast_parameters()
        self.steps = 0

    # ---- replica consistency (K-C3, K-C4) ----------------------------------------------------
    def _rebind_grads(self, flat_g: torch.Tensor) -> None:
        off = 0
        for p in self.model.parameters():
            n = p.numel()
            p.grad = flat_g[off: off + n].view(p.shape)
            off += n
        self.flat_g = flat_g

    def _rebind_params(self, flat_p: torch.Tensor) -> None:
        off = 0
        for p in self.model.parameters():
            n = p.numel()
            p.data = flat_p[off: off + n].view(p.shape)
            off += n
        self.flat_p = flat_p

    def broadcast_parameters(self) -> None:
        """DDP's constructor broadcast (K-C3): every replica starts from rank 0's values."""
        import torch.distributed as dist
        if self.ctx.initialized:
            dist.broadcast(self.flat_p, src=0)
            self.sync_buffers()

    def sync_buffers(self) -> None:
        """Rank 0's BatchNorm running statistics are authoritative (DDP ``broadcast_buffers``, K-C4).
        They only matter when evaluating or saving, so they are synchronised there instead of
        before every forward; what any rank can observe is identical."""
        import torch.distributed as dist
        if self.ctx.initialized:
            for b in self.model.buffers():
                dist.broadcast(b, src=0)

    # ---- steps ---------------------------------------------------------------------------------
    def _inputs(self, batch: Batch, train: bool) -> torch.Tensor:
        if self.cuda:
            batch.wait()
            from . import ops
            if ops.native_available():
                x = ops.preprocess(batch.images, batch.params if train else None,
                                   out_dtype=torch.bfloat16, channels=3)      # NHWC bf16
                return x.permute(0, 3, 1, 2)                                  # NCHW view, channels_last memory
            x = apply_pipeline_torch(batch.images, batch.params if train else None, self.cfg.img_size)
            return x.contiguous(memory_format=torch.channels_last)
        if batch.x is not None:        # the loader thread already ran the pipeline (CPU mode)
            return batch.x
        return apply_pipeline_torch(batch.images, batch.params if train else None, self.cfg.img_size)

    def _forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.cuda:
            with torch.autocast("cuda", dtype=torch.bfloat16):
                return self.model(x).float()
        return self.model(x)

    def train_step(self, batch: Batch, lr: float) -> None:
        self.model.train()
        x = self._inputs(batch, True)
        logits = self._forward(x)
        loss = F.cross_entropy(logits, batch.labels)              # ref :178 (mean reduction)
        loss.backward()                                           # accumulates into the flat grad views
        self.steps += 1
        if self.fused is not None:
            self.fused.step(lr)                                   # all-reduce + 1/W + Adam + grad zero, one kernel
            self._rebind_if_needed()
        else:
            if self.ctx.initialized:
                import torch.distributed as dist
                dist.all_reduce(self.flat_g)                      # baseline path (K-C5 via NCCL/gloo)
            self.adam.step_torch(self.flat_p, self.flat_g, lr, grad_scale=1.0 / self.world)
            self.flat_g.zero_()
        with torch.no_grad():
            n = batch.n
            self.metrics[0] += loss.detach().double() * n
            self.metrics[1] += (logits.argmax(1) == batch.labels).sum()
            self.metrics[2] += n

    def _rebind_if_needed(self) -> None:
        g = self.fused.current_grad_buffer()
        if g.data_ptr() != self.flat_g.data_ptr():
            self._rebind_grads(g)

    @torch.no_grad()
    def eval_step(self, batch: Batch) -> None:
        self.model.eval()
        x = self._inputs(batch, False)
        logits = self._forward(x)
        loss = F.cross_entropy(logits, batch.labels)
        self.metrics[0] += loss.double() * batch.n
        self.metrics[1] += (logits.argmax(1) == batch.labels).sum()
        self.metrics[2] += batch.n

    @torch.no_grad()
    def predict_logits(self, images_u8: torch.Tensor) -> torch.Tensor:
        self.model.eval()
        b = Batch(images_u8.to(self.device), torch.zeros(images_u8.shape[0], dtype=torch.int64, device=self.device),
                  torch.zeros(images_u8.shape[0], 12), images_u8.shape[0])
        return self._forward(self._inputs(b, False))

    # ---- metrics / state -------------------------------------------------------------------
    def read_metrics(self, reset: bool = True) -> Tuple[float, float, float]:
        vals = self.metrics.tolist()          # the one host sync per epoch
        if reset:
            self.metrics.zero_()
        return vals[0], vals[1], vals[2]

    def state_dict(self) -> "OrderedDict[str, torch.Tensor]":
        return OrderedDict((k, v.detach().cpu().clone()) for k, v in self.model.state_dict().items())

    def load_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        own = self.model.state_dict()
        with torch.no_grad():
            for k, v in own.items():
                v.copy_(sd[k])

    def optimizer_state(self) -> Dict[str, object]:
        return self.adam.state_dict() if self.fused is None else self.fused.state_dict()

    def load_optimizer_state(self, sd: Dict[str, object]) -> None:
        if self.fused is None:
            self.adam.load_state_dict(sd)
        else:
            self.fused.load_state_dict(sd)

    def parameter_count(self) -> int:
        return sum(p.numel() for p in self.model.parameters())

    def param_hash(self) -> float:
        """Cheap cross-rank equality probe for the replica-consistency check."""
        return float(self.flat_p.double().sum().item())

    def close(self) -> None:
        if self.fused is not None:
            self.fused.close()
