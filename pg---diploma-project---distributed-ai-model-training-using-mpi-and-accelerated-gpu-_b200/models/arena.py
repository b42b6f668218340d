"""Flat parameter arena <-> torchvision ``state_dict`` (hard part 9).

The native engine keeps every trainable tensor in ONE flat index space shared by five arrays
(fp32 master weights, Adam m, Adam v, fp32 gradients x2 buffers, bf16 working weights), laid
out in the order and shape the kernels consume:

* pointwise conv ``[Cout, Cin, 1, 1]`` -> ``[Cout_pad, Cin_pad]`` row-major (K contiguous; 24 -> 32 padding, pads stay 0)
* depthwise conv ``[C, 1, 3, 3]``      -> ``[9, C_pad]`` (tap-major, channels contiguous)
* stem conv ``[32, 3, 3, 3]``           -> ``[27, 32]`` ((ky,kx,ci)-major, Cout contiguous)
* BatchNorm weight / bias ``[C]``       -> ``[C_pad]``
* classifier ``[classes, 1280]``, ``[classes]`` unchanged

so "a gradient bucket" is an address range (the reference's DDP copies grads into buckets, K-C5)
and the fused all-reduce + Adam kernel is a single pass over ``[0, total)``.  Padding and
permutation never leak: ``export_state_dict`` returns exactly the 314-entry fp32/int64 dict that
``torchvision.models.mobilenet_v2`` + ``nn.Linear(1280, 10)`` loads strictly
(cifar10_mpi_mobilenet_224.py:240-249; GROUP03.pdf p.26), and ``import_state_dict`` is its inverse.
"""
from __future__ import annotations

from collections import OrderedDict
from dataclasses import dataclass
from typing import Dict, List, Tuple

import torch

from .spec import LAST_CHANNELS, NetSpec, build_spec, pad16

ALIGN = 64  # elements; 256 B in fp32 and 128 B in bf16, enough for TMA (16 B) and 16-byte vector access


@dataclass(frozen=True)
class Segment:
    key: str                       # state_dict key
    kind: str                      # pw | dw | stem | bn | fc_w | fc_b
    tv_shape: Tuple[int, ...]      # torchvision shape
    eng_shape: Tuple[int, ...]     # engine (padded / permuted) shape
    offset: int                    # element offset in the arena
    conv_index: int                # owning conv (-1 for classifier)

    @property
    def numel(self) -> int:
        n = 1
        for d in self.eng_shape:
            n *= d
        return n


@dataclass(frozen=True)
class BufferSegment:
    key: str                       # running_mean / running_var key
    channels: int
    channels_pad: int
    offset: int
    conv_index: int


class ArenaLayout:
    def __init__(self, spec: NetSpec | None = None) -> None:
        self.spec = spec or build_spec()
        segs: List[Segment] = []
        bufs: List[BufferSegment] = []
        off = 0
        boff = 0

        def add(key, kind, tv, eng, ci):
            nonlocal off
            s = Segment(key, kind, tuple(tv), tuple(eng), off, ci)
            segs.append(s)
            off += (s.numel + ALIGN - 1) // ALIGN * ALIGN

        for c in self.spec.convs:
            cp_out = pad16(c.cout)
            if c.kind == "pw1x1":
                add(c.conv_key + ".weight", "pw", c.weight_shape, (cp_out, pad16(c.cin)), c.index)
            elif c.kind == "dw3x3":
                add(c.conv_key + ".weight", "dw", c.weight_shape, (9, cp_out), c.index)
            else:
                add(c.conv_key + ".weight", "stem", c.weight_shape, (27, c.cout), c.index)
            add(c.bn_key + ".weight", "bn", (c.cout,), (cp_out,), c.index)
            add(c.bn_key + ".bias", "bn", (c.cout,), (cp_out,), c.index)
            for nm in ("running_mean", "running_var"):
                bufs.append(BufferSegment(f"{c.bn_key}.{nm}", c.cout, cp_out, boff, c.index))
                boff += (cp_out + ALIGN - 1) // ALIGN * ALIGN
        add("classifier.1.weight", "fc_w", (self.spec.num_classes, LAST_CHANNELS),
            (self.spec.num_classes, LAST_CHANNELS), -1)
        add("classifier.1.bias", "fc_b", (self.spec.num_classes,), (self.spec.num_classes,), -1)
        self.segments = segs
        self.buffers = bufs
        self.total = off
        self.buffer_total = boff
        self.by_key: Dict[str, Segment] = {s.key: s for s in segs}
        self.buf_by_key: Dict[str, BufferSegment] = {b.key: b for b in bufs}

    # ---- per-tensor conversions --------------------------------------------------------------
    @staticmethod
    def to_engine(seg: Segment, t: torch.Tensor) -> torch.Tensor:
        t = t.detach().to(torch.float32)
        if seg.kind == "pw":
            out = torch.zeros(seg.eng_shape, dtype=torch.float32, device=t.device)
            out[: seg.tv_shape[0], : seg.tv_shape[1]] = t.reshape(seg.tv_shape[0], seg.tv_shape[1])
            return out
        if seg.kind == "dw":
            out = torch.zeros(seg.eng_shape, dtype=torch.float32, device=t.device)
            out[:, : seg.tv_shape[0]] = t.reshape(seg.tv_shape[0], 9).t()
            return out
        if seg.kind == "stem":
            return t.permute(2, 3, 1, 0).reshape(seg.eng_shape).contiguous()
        if seg.kind == "bn":
            out = torch.zeros(seg.eng_shape, dtype=torch.float32, device=t.device)
            out[: seg.tv_shape[0]] = t
            return out
        return t.reshape(seg.eng_shape).contiguous()

    @staticmethod
    def from_engine(seg: Segment, e: torch.Tensor) -> torch.Tensor:
        e = e.reshape(seg.eng_shape)
        if seg.kind == "pw":
            return e[: seg.tv_shape[0], : seg.tv_shape[1]].reshape(seg.tv_shape).contiguous().clone()
        if seg.kind == "dw":
            return e[:, : seg.tv_shape[0]].t().reshape(seg.tv_shape).contiguous().clone()
        if seg.kind == "stem":
            co, ci = seg.tv_shape[0], seg.tv_shape[1]
            return e.reshape(3, 3, ci, co).permute(3, 2, 0, 1).contiguous().clone()
        if seg.kind == "bn":
            return e[: seg.tv_shape[0]].clone()
        return e.reshape(seg.tv_shape).clone()

    # ---- whole-dict conversions --------------------------------------------------------------
    def import_state_dict(self, sd: Dict[str, torch.Tensor], device: str | torch.device = "cpu"
                          ) -> Tuple[torch.Tensor, torch.Tensor, int]:
        """state_dict -> (params[total] fp32, buffers[buffer_total] fp32, num_batches_tracked)."""
        params = torch.zeros(self.total, dtype=torch.float32)
        for s in self.segments:
            if s.key not in sd:
                raise KeyError(f"state_dict is missing {s.key}")
            if tuple(sd[s.key].shape) != s.tv_shape:
                raise ValueError(f"{s.key}: expected shape {s.tv_shape}, got {tuple(sd[s.key].shape)}")
            params[s.offset: s.offset + s.numel] = self.to_engine(s, sd[s.key].cpu()).reshape(-1)
        buffers = torch.zeros(self.buffer_total, dtype=torch.float32)
        for b in self.buffers:
            v = sd[b.key].detach().to(torch.float32).cpu()
            if b.key.endswith("running_var"):
                buffers[b.offset: b.offset + b.channels_pad] = 1.0
            buffers[b.offset: b.offset + b.channels] = v
        nbt = 0
        for k, v in sd.items():
            if k.endswith("num_batches_tracked"):
                nbt = int(v)
                break
        return params.to(device), buffers.to(device), nbt

    def export_state_dict(self, params: torch.Tensor, buffers: torch.Tensor,
                          num_batches_tracked: int = 0) -> "OrderedDict[str, torch.Tensor]":
        """(params, buffers) -> the 314-entry CPU fp32/int64 dict in torchvision order."""
        params = params.detach().to("cpu", torch.float32)
        buffers = buffers.detach().to("cpu", torch.float32)
        out: "OrderedDict[str, torch.Tensor]" = OrderedDict()
        for c in self.spec.convs:
            for suffix in (c.conv_key + ".weight", c.bn_key + ".weight", c.bn_key + ".bias"):
                s = self.by_key[suffix]
                out[suffix] = self.from_engine(s, params[s.offset: s.offset + s.numel])
            for nm in ("running_mean", "running_var"):
                b = self.buf_by_key[f"{c.bn_key}.{nm}"]
                out[b.key] = buffers[b.offset: b.offset + b.channels].clone()
            out[c.bn_key + ".num_batches_tracked"] = torch.tensor(int(num_batches_tracked), dtype=torch.int64)
        for k in ("classifier.1.weight", "classifier.1.bias"):
            s = self.by_key[k]
            out[k] = self.from_engine(s, params[s.offset: s.offset + s.numel])
        return out

    def trainable_mask(self) -> torch.Tensor:
        """1.0 where an arena slot holds a real parameter, 0.0 on alignment / channel padding."""
        m = torch.zeros(self.total, dtype=torch.float32)
        for s in self.segments:
            ones = torch.ones(s.tv_shape, dtype=torch.float32)
            m[s.offset: s.offset + s.numel] = self.to_engine(s, ones).reshape(-1)
        return m

    def offsets_table(self) -> List[int]:
        """Flat int table handed to the native executor: for each conv (w, gamma, beta, rmean, rvar)
        element offsets, then (fc_w, fc_b)."""
        t: List[int] = []
        for c in self.spec.convs:
            t.append(self.by_key[c.conv_key + ".weight"].offset)
            t.append(self.by_key[c.bn_key + ".weight"].offset)
            t.append(self.by_key[c.bn_key + ".bias"].offset)
            t.append(self.buf_by_key[c.bn_key + ".running_mean"].offset)
            t.append(self.buf_by_key[c.bn_key + ".running_var"].offset)
        t.append(self.by_key["classifier.1.weight"].offset)
        t.append(self.by_key["classifier.1.bias"].offset)
        return t
