"""Inference entry points.

* ``predict_cifar10_image(path, topk=3, conf_threshold=0.5)`` -- the helper at
  cifar10_128batch.py:159-188: open an image, test transform, batch-1 forward, softmax, top-k,
  print, flag "uncertain" below the threshold, return ``[(label, confidence)]``.
* ``Predictor`` -- what the Gradio app does at import time (GROUP03.pdf p.26): rebuild the model,
  ``torch.load`` the checkpoint, strict ``load_state_dict``, ``eval()``; then ``predict(PIL image)
  -> {class: prob}`` for the top 3.

Differences on purpose: ImageNet normalisation (the Gradio app uses CIFAR statistics although the
model was trained with ImageNet ones -- SURVEY §2.1 quirks); any image size is accepted (it is
resampled to the 32x32 source grid the training pipeline upsamples from when larger, matching how
CIFAR images reach the network).  On a GPU the forward runs on the native engine in eval mode
(BatchNorm folded into scale/shift), else on ATen ops.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple, Union

import torch

from . import CIFAR10_CLASSES
from .models.mobilenetv2 import MobileNetV2
from .utils.checkpoint import load_model


def _to_u8_32(img) -> torch.Tensor:
    """PIL image / ndarray / tensor -> uint8 [32,32,3] (RGB)."""
    import numpy as np
    try:
        from PIL import Image
    except Exception:  # pragma: no cover
        Image = None
    if Image is not None and isinstance(img, Image.Image):
        img = img.convert("RGB")                                  # ref 128batch:164
        if img.size != (32, 32):
            img = img.resize((32, 32), Image.BILINEAR)
        arr = np.asarray(img, dtype=np.uint8)
    else:
        arr = np.asarray(img)
        if arr.ndim == 2:
            arr = np.stack([arr] * 3, axis=-1)
        if arr.shape[:2] != (32, 32):
            t = torch.from_numpy(arr.astype("float32")).permute(2, 0, 1)[None]
            t = torch.nn.functional.interpolate(t, size=(32, 32), mode="bilinear", align_corners=False)
            arr = t[0].permute(1, 2, 0).round().clamp(0, 255).to(torch.uint8).numpy()
        arr = arr.astype(np.uint8)[..., :3]
    return torch.from_numpy(np.ascontiguousarray(arr))


class Predictor:
    def __init__(self, checkpoint: Optional[str] = None, device: Optional[str] = None,
                 classes: Sequence[str] = CIFAR10_CLASSES, state_dict: Optional[dict] = None,
                 engine: str = "auto") -> None:
        self.classes = tuple(classes)
        self.device = torch.device(device or ("cuda" if torch.cuda.is_available() else "cpu"))
        sd = state_dict if state_dict is not None else (load_model(checkpoint) if checkpoint else None)
        self._native = None
        if self.device.type == "cuda" and engine in ("auto", "native"):
            try:
                from .engine_native import NativeInference
                self._native = NativeInference(len(self.classes), self.device, sd)
            except Exception:
                if engine == "native":
                    raise
        if self._native is None:
            self.model = MobileNetV2(num_classes=len(self.classes))
            if sd is not None:
                self.model.load_state_dict(sd, strict=True)           # GROUP03.pdf p.26
            self.model.to(self.device).eval()

    @torch.no_grad()
    def logits(self, images_u8: torch.Tensor) -> torch.Tensor:
        """uint8 [B,32,32,3] -> fp32 logits [B, classes]."""
        if self._native is not None:
            return self._native.logits(images_u8.to(self.device))
        from .data.transforms import apply_pipeline_torch
        x = apply_pipeline_torch(images_u8.to(self.device), None)
        return self.model(x).float()

    @torch.no_grad()
    def probabilities(self, img) -> torch.Tensor:
        x = _to_u8_32(img)[None]
        return torch.softmax(self.logits(x)[0], dim=0).cpu()         # ref 128batch:170

    def topk(self, img, k: int = 3) -> List[Tuple[str, float]]:
        p = self.probabilities(img)
        conf, idx = torch.topk(p, k=min(k, p.numel()))               # ref 128batch:172
        return [(self.classes[i], float(c)) for c, i in zip(conf.tolist(), idx.tolist())]

    def predict(self, img) -> Dict[str, float]:
        """Gradio contract: ``predict(PIL.Image) -> {class: prob}`` for the top 3 (GROUP03.pdf p.27)."""
        return {name: conf for name, conf in self.topk(img, 3)}


def predict_cifar10_image(img_path: Union[str, "object"], topk: int = 3, conf_threshold: float = 0.5,
                          predictor: Optional[Predictor] = None, checkpoint: Optional[str] = None,
                          quiet: bool = False) -> List[Tuple[str, float]]:
    """Top-k prediction for an image file (cifar10_128batch.py:159-188)."""
    from PIL import Image
    predictor = predictor or Predictor(checkpoint)
    img = Image.open(img_path).convert("RGB") if isinstance(img_path, (str, bytes)) else img_path
    results = predictor.topk(img, topk)
    best_label, best_conf = results[0]
    if not quiet:
        print(f"\nImage: {img_path}")
        print("Top predictions:")
        for label, conf in results:
            print(f"  {label:10s}: {conf:.3f}")
        if best_conf < conf_threshold:
            print(f"Model is uncertain (max confidence {best_conf:.3f} < {conf_threshold}).")
        else:
            print(f"Predicted: {best_label} (confidence {best_conf:.3f})")
    return results


def main(argv: Optional[Sequence[str]] = None) -> int:
    import argparse
    ap = argparse.ArgumentParser(prog="b200dp.inference", description="Top-k CIFAR-10 prediction for image files")
    ap.add_argument("images", nargs="+")
    ap.add_argument("--checkpoint", default="best_mobilenetv2_cifar10_224.pth")
    ap.add_argument("--topk", type=int, default=3)
    ap.add_argument("--threshold", type=float, default=0.5)
    ap.add_argument("--device", default=None)
    ns = ap.parse_args(argv)
    pred = Predictor(ns.checkpoint, ns.device)
    for path in ns.images:
        predict_cifar10_image(path, ns.topk, ns.threshold, predictor=pred)
    return 0


if __name__ == "__main__":
    import sys
    sys.exit(main())
