"""Inference entry points.

* ``predict_cifar10_image(path, topk=3, conf_threshold=0.5)`` -- the helper at
  cifar10_128batch.py:159-188: open an image, test transform, batch-1 forward, softmax, top-k,
  print, flag "uncertain" below the threshold, return ``[(label, confidence)]``.
* ``Predictor`` -- what the Gradio app does at import time (GROUP03.pdf p.26): rebuild the model,
  ``torch.load`` the checkpoint, strict ``load_state_dict``, ``eval()``; then ``predict(PIL image)
  -> {class: prob}`` for the top 3.

The image goes through the reference's inference transform: ``Resize((224, 224))`` of the ORIGINAL
image (PIL's anti-aliased bilinear), ``ToTensor``, ``Normalize`` (cifar10_128batch.py:42-47, :157).
On a GPU that is one kernel (``ops.resize_normalize``), the forward runs on the native engine in
eval mode (BatchNorm folded into scale/shift) and softmax + top-k is the ``softmax_topk`` kernel;
on the CPU it is PIL + ATen, i.e. the reference's own path.

Difference on purpose: ImageNet normalisation everywhere (the Gradio app uses CIFAR statistics
although the model was trained with ImageNet ones -- SURVEY §2.1 quirks).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple, Union

import torch

from . import CIFAR10_CLASSES
from .models.mobilenetv2 import MobileNetV2
from .utils.checkpoint import load_model


def _to_u8_hwc(img) -> torch.Tensor:
    """PIL image / ndarray / tensor -> uint8 [H,W,3] (RGB) at its own resolution."""
    import numpy as np
    try:
        from PIL import Image
    except Exception:  # pragma: no cover
        Image = None
    if Image is not None and isinstance(img, Image.Image):
        arr = np.asarray(img.convert("RGB"), dtype=np.uint8)      # ref 128batch:164
    else:
        arr = np.asarray(img.cpu() if isinstance(img, torch.Tensor) else img)
        if arr.ndim == 2:
            arr = np.stack([arr] * 3, axis=-1)
        arr = arr.astype(np.uint8)[..., :3]
    return torch.from_numpy(np.ascontiguousarray(arr))


_MEAN = torch.tensor([0.485, 0.456, 0.406]).view(3, 1, 1)
_STD = torch.tensor([0.229, 0.224, 0.225]).view(3, 1, 1)


def reference_transform(img_hwc_u8: torch.Tensor, size: int = 224) -> torch.Tensor:
    """The reference's test transform on the CPU: PIL ``Resize((size, size))`` + ToTensor + Normalize -> fp32 [3,S,S]."""
    from PIL import Image
    pil = Image.fromarray(img_hwc_u8.numpy()).resize((size, size), Image.BILINEAR)
    import numpy as np
    x = torch.from_numpy(np.asarray(pil, dtype=np.uint8).copy()).permute(2, 0, 1).float() / 255.0
    return (x - _MEAN) / _STD


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
        """uint8 [B,32,32,3] (CIFAR-sized) -> fp32 logits [B, classes]: the evaluation path of training."""
        if self._native is not None:
            return self._native.logits(images_u8.to(self.device))
        from .data.transforms import apply_pipeline_torch
        x = apply_pipeline_torch(images_u8.to(self.device), None)
        return self.model(x).float()

    @torch.no_grad()
    def image_logits(self, img) -> torch.Tensor:
        """One image of any size -> fp32 logits [classes], through Resize((224, 224)) of the original."""
        hwc = _to_u8_hwc(img)
        if self._native is not None:
            from . import ops
            x = ops.resize_normalize(hwc[None].to(self.device), 224)
            return self._native.logits_preprocessed(x)[0]
        x = reference_transform(hwc)[None].to(self.device)
        return self.model(x).float()[0]

    @torch.no_grad()
    def probabilities(self, img) -> torch.Tensor:
        lg = self.image_logits(img)
        if self._native is not None:
            from . import ops
            return ops.softmax_topk(lg[None].contiguous(), 1)[0][0].cpu()
        return torch.softmax(lg, dim=0).cpu()                        # ref 128batch:170

    def topk(self, img, k: int = 3) -> List[Tuple[str, float]]:
        lg = self.image_logits(img)
        k = min(k, lg.numel())
        if self._native is not None:                                 # softmax + top-k on the device (K15)
            from . import ops
            _, conf, idx = ops.softmax_topk(lg[None].contiguous(), k)
            conf, idx = conf[0].cpu(), idx[0].cpu()
        else:
            conf, idx = torch.topk(torch.softmax(lg, dim=0).cpu(), k=k)     # ref 128batch:170-172
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
    if not quiet:                                                    # the reference's printout, ref 128batch:176-186
        print("Top-{} predictions:".format(topk))
        for label, conf in results:
            print(f"  {label}: {conf:.3f}")
        if best_conf < conf_threshold:
            print(f"Prediction uncertain (best_conf={best_conf:.3f} < {conf_threshold})")
        else:
            print(f"Predicted: {best_label} confidence: {best_conf:.3f}")
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
