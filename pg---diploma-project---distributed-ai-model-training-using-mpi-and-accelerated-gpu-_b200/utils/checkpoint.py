"""Checkpoint save / resume.

The reference keeps ``best_state = deepcopy(state_dict())`` in memory whenever test accuracy
improves (cifar10_mpi_mobilenet_224.py:159-160, 238-240) and writes it once, after the last epoch,
from rank 0 (:243-250) -- a crash loses everything and there is no resume.  Kept: the file names
and the exact content (a 314-entry torchvision-layout ``OrderedDict`` of CPU tensors, no
``module.`` prefix).  Added: the best file is (atomically) rewritten whenever it improves, and an
optional sidecar ``<name>.resume`` holds Adam m/v/step, epoch, best accuracy and RNG state so a run
can continue; the torchvision-format file is never polluted with those.
"""
from __future__ import annotations

import os
import tempfile
from collections import OrderedDict
from typing import Any, Dict, Optional

import torch


def atomic_torch_save(obj: Any, path: str) -> None:
    d = os.path.dirname(os.path.abspath(path)) or "."
    os.makedirs(d, exist_ok=True)
    fd, tmp = tempfile.mkstemp(prefix="ckpt_", suffix=".part", dir=d)
    try:
        with os.fdopen(fd, "wb") as f:
            torch.save(obj, f)
        os.replace(tmp, path)
    finally:
        if os.path.exists(tmp):
            os.remove(tmp)


def to_cpu_state_dict(sd: Dict[str, torch.Tensor]) -> "OrderedDict[str, torch.Tensor]":
    out: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    for k, v in sd.items():
        if k.startswith("module."):
            k = k[len("module."):]
        out[k] = v.detach().to("cpu").clone()
    return out


def save_model(sd: Dict[str, torch.Tensor], path: str) -> None:
    atomic_torch_save(to_cpu_state_dict(sd), path)


def load_model(path: str, map_location: str = "cpu") -> "OrderedDict[str, torch.Tensor]":
    """What the reference's Gradio loader does: ``torch.load(path, map_location=device)`` (GROUP03.pdf p.26)."""
    return torch.load(path, map_location=map_location, weights_only=True)


def save_resume(path: str, *, epoch: int, best_acc: float, optimizer: Dict[str, Any],
                model: Dict[str, torch.Tensor], best_model: Optional[Dict[str, torch.Tensor]] = None,
                extra: Optional[Dict[str, Any]] = None) -> None:
    atomic_torch_save({"format": "b200dp-resume-v1", "epoch": int(epoch), "best_acc": float(best_acc),
                       "optimizer": optimizer, "model": to_cpu_state_dict(model),
                       "best_model": to_cpu_state_dict(best_model) if best_model is not None else None,
                       "extra": extra or {}}, path)


def load_resume(path: str) -> Dict[str, Any]:
    obj = torch.load(path, map_location="cpu", weights_only=True)      # tensors, numbers, strings, dicts only
    if not isinstance(obj, dict) or obj.get("format") != "b200dp-resume-v1":
        raise ValueError(f"{path} is not a b200dp resume file")
    return obj
