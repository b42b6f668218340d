"""MobileNetV2 with a replaceable classifier head, built from ``spec.py`` with ATen ops.

This is the *torch engine*: it serves mode 1 (serial CPU, cifar10_serial_mobilenet_224.py), is
the fp32 numerical oracle for every native kernel, and is the comparison baseline on the GPU.
The module tree reproduces torchvision's names exactly (``features.N.conv.K...``,
``classifier.1``), so ``state_dict()`` has the 314 entries of SURVEY §2.5 and loads strictly
into ``torchvision.models.mobilenet_v2`` + ``nn.Linear(1280, 10)`` and back
(reference: model build cifar10_mpi_mobilenet_224.py:137-140; Gradio loader GROUP03.pdf p.26).
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import nn

from .spec import BN_EPS, BN_MOMENTUM, LAST_CHANNELS, NetSpec, build_spec


def _conv_bn_act(cin: int, cout: int, k: int, stride: int, groups: int) -> nn.Sequential:
    return nn.Sequential(
        nn.Conv2d(cin, cout, k, stride, (k - 1) // 2, groups=groups, bias=False),
        nn.BatchNorm2d(cout, eps=BN_EPS, momentum=BN_MOMENTUM),
        nn.ReLU6(inplace=True),
    )


class InvertedResidual(nn.Module):
    def __init__(self, cin: int, cout: int, stride: int, expand: int) -> None:
        super().__init__()
        hidden = cin * expand
        self.use_res = stride == 1 and cin == cout
        layers = []
        if expand != 1:
            layers.append(_conv_bn_act(cin, hidden, 1, 1, 1))
        layers.append(_conv_bn_act(hidden, hidden, 3, stride, hidden))
        layers.append(nn.Conv2d(hidden, cout, 1, 1, 0, bias=False))
        layers.append(nn.BatchNorm2d(cout, eps=BN_EPS, momentum=BN_MOMENTUM))
        self.conv = nn.Sequential(*layers)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return x + self.conv(x) if self.use_res else self.conv(x)


class MobileNetV2(nn.Module):
    def __init__(self, num_classes: int = 10, dropout: float = 0.2, img_size: int = 224,
                 spec: Optional[NetSpec] = None) -> None:
        super().__init__()
        self.spec = spec or build_spec(img_size, num_classes)
        feats = [_conv_bn_act(3, self.spec.convs[0].cout, 3, 2, 1)]
        for b in self.spec.blocks:
            feats.append(InvertedResidual(b.cin, b.cout, b.stride, b.expand))
        feats.append(_conv_bn_act(self.spec.blocks[-1].cout, LAST_CHANNELS, 1, 1, 1))
        self.features = nn.Sequential(*feats)
        self.classifier = nn.Sequential(nn.Dropout(p=dropout), nn.Linear(LAST_CHANNELS, num_classes))
        self.reset_parameters()

    def reset_parameters(self) -> None:
        # torchvision's initialisation (the reference fine-tunes from ImageNet weights, which do not
        # exist offline; random init follows the same scheme so a from-scratch run is well-posed).
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out")
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.ones_(m.weight)
                nn.init.zeros_(m.bias)
            elif isinstance(m, nn.Linear):
                nn.init.normal_(m.weight, 0, 0.01)
                nn.init.zeros_(m.bias)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.features(x)
        x = torch.nn.functional.adaptive_avg_pool2d(x, (1, 1)).flatten(1)
        return self.classifier(x)


def build_model(num_classes: int = 10, dropout: float = 0.2, img_size: int = 224,
                pretrained: str = "", map_location: str = "cpu") -> MobileNetV2:
    """``mobilenet_v2(pretrained=True)`` + new head (ref :137-139).  ``pretrained`` is a path to a
    torchvision-format state_dict; a 1000-class head in the file is dropped, mirroring the
    reference's head replacement.  All layers stay trainable (the reference freezes nothing)."""
    model = MobileNetV2(num_classes=num_classes, dropout=dropout, img_size=img_size)
    if pretrained:
        sd = torch.load(pretrained, map_location=map_location, weights_only=True)
        own = model.state_dict()
        for k in ("classifier.1.weight", "classifier.1.bias"):
            if k in sd and sd[k].shape != own[k].shape:
                sd[k] = own[k]
        model.load_state_dict(sd, strict=True)
    return model
