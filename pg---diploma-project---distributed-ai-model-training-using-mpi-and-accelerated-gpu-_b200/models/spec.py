"""MobileNetV2 architecture table (SURVEY §2.4 "Model facts", §3.3).

One place that says what the network is: 52 convolutions (1 stem 3x3 s2, 17 depthwise 3x3,
34 pointwise 1x1), 52 BatchNorm2d, 35 ReLU6, 10 residual adds, global average pool, dropout,
Linear(1280, num_classes).  The torch-op model (``mobilenetv2.py``), the flat-arena layout
(``state_dict.py``) and the native executor (``csrc/engine.cpp`` builds the same table in
C++; ``tests/test_spec.py`` checks they agree) are all derived from it.  Key names follow
torchvision's ``mobilenet_v2`` so checkpoints interchange with the reference
(cifar10_mpi_mobilenet_224.py:137-139, :240-249).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Tuple

# (expand t, out channels c, repeats n, first stride s) -- torchvision mobilenetv2.py inverted_residual_setting
INVERTED_RESIDUAL_SETTING: Tuple[Tuple[int, int, int, int], ...] = (
    (1, 16, 1, 1), (6, 24, 2, 2), (6, 32, 3, 2), (6, 64, 4, 2),
    (6, 96, 3, 1), (6, 160, 3, 2), (6, 320, 1, 1),
)
STEM_CHANNELS = 32
LAST_CHANNELS = 1280
BN_EPS = 1e-5
BN_MOMENTUM = 0.1


def pad16(c: int) -> int:
    """Channel counts are padded to a multiple of 16 in the native layout (UMMA K=16 for bf16,
    N % 16 at M=128).  Only the 24-channel tensors are affected (24 -> 32)."""
    return (c + 15) // 16 * 16


@dataclass(frozen=True)
class ConvSpec:
    index: int            # position in forward order, 0..51
    role: str             # stem | expand | dw | project | last
    kind: str             # conv3x3 | dw3x3 | pw1x1
    cin: int
    cout: int
    stride: int
    hin: int              # input spatial size (square)
    hout: int
    relu6: bool           # ReLU6 after the BatchNorm (False for the linear bottleneck projection)
    conv_key: str         # state_dict prefix of the conv weight, e.g. "features.2.conv.0.0"
    bn_key: str           # state_dict prefix of the BatchNorm, e.g. "features.2.conv.0.1"
    block: int            # features.<block>

    @property
    def weight_shape(self) -> Tuple[int, int, int, int]:
        if self.kind == "dw3x3":
            return (self.cout, 1, 3, 3)
        if self.kind == "conv3x3":
            return (self.cout, self.cin, 3, 3)
        return (self.cout, self.cin, 1, 1)

    @property
    def macs_per_image(self) -> int:
        k = 9 if self.kind != "pw1x1" else 1
        cin = 1 if self.kind == "dw3x3" else self.cin
        return self.hout * self.hout * self.cout * cin * k


@dataclass(frozen=True)
class BlockSpec:
    index: int            # features.<index>, 1..17
    cin: int
    cout: int
    stride: int
    expand: int
    hin: int
    hout: int
    use_res: bool
    convs: Tuple[int, ...]   # indices into the conv list (expand?, dw, project)

    @property
    def hidden(self) -> int:
        return self.cin * self.expand


@dataclass(frozen=True)
class NetSpec:
    img_size: int
    num_classes: int
    convs: Tuple[ConvSpec, ...]
    blocks: Tuple[BlockSpec, ...]

    @property
    def feature_hw(self) -> int:
        return self.convs[-1].hout

    def param_count(self) -> int:
        n = 0
        for c in self.convs:
            w = c.weight_shape
            n += w[0] * w[1] * w[2] * w[3] + 2 * c.cout
        return n + LAST_CHANNELS * self.num_classes + self.num_classes

    def macs_per_image(self) -> int:
        return sum(c.macs_per_image for c in self.convs) + LAST_CHANNELS * self.num_classes


def _out_hw(h: int, stride: int) -> int:
    return (h + 2 - 3) // stride + 1   # 3x3, pad 1


def build_spec(img_size: int = 224, num_classes: int = 10) -> NetSpec:
    convs: List[ConvSpec] = []
    blocks: List[BlockSpec] = []
    h = img_size
    ho = _out_hw(h, 2)
    convs.append(ConvSpec(0, "stem", "conv3x3", 3, STEM_CHANNELS, 2, h, ho, True,
                          "features.0.0", "features.0.1", 0))
    h = ho
    cin = STEM_CHANNELS
    feat = 1
    for t, c, n, s in INVERTED_RESIDUAL_SETTING:
        for i in range(n):
            stride = s if i == 0 else 1
            hidden = cin * t
            ho = _out_hw(h, stride)
            ids = []
            base = f"features.{feat}.conv"
            j = 0
            if t != 1:
                convs.append(ConvSpec(len(convs), "expand", "pw1x1", cin, hidden, 1, h, h, True,
                                      f"{base}.{j}.0", f"{base}.{j}.1", feat))
                ids.append(len(convs) - 1)
                j += 1
            convs.append(ConvSpec(len(convs), "dw", "dw3x3", hidden, hidden, stride, h, ho, True,
                                  f"{base}.{j}.0", f"{base}.{j}.1", feat))
            ids.append(len(convs) - 1)
            j += 1
            convs.append(ConvSpec(len(convs), "project", "pw1x1", hidden, c, 1, ho, ho, False,
                                  f"{base}.{j}", f"{base}.{j + 1}", feat))
            ids.append(len(convs) - 1)
            blocks.append(BlockSpec(feat, cin, c, stride, t, h, ho, stride == 1 and cin == c, tuple(ids)))
            cin = c
            h = ho
            feat += 1
    convs.append(ConvSpec(len(convs), "last", "pw1x1", cin, LAST_CHANNELS, 1, h, h, True,
                          f"features.{feat}.0", f"features.{feat}.1", feat))
    return NetSpec(img_size, num_classes, tuple(convs), tuple(blocks))


def state_dict_keys(spec: NetSpec) -> List[Tuple[str, Tuple[int, ...], str]]:
    """(key, shape, dtype) in torchvision ``state_dict()`` order: 158 parameters + 156 buffers = 314 entries."""
    out: List[Tuple[str, Tuple[int, ...], str]] = []
    for c in spec.convs:
        out.append((c.conv_key + ".weight", c.weight_shape, "float32"))
        out.append((c.bn_key + ".weight", (c.cout,), "float32"))
        out.append((c.bn_key + ".bias", (c.cout,), "float32"))
        out.append((c.bn_key + ".running_mean", (c.cout,), "float32"))
        out.append((c.bn_key + ".running_var", (c.cout,), "float32"))
        out.append((c.bn_key + ".num_batches_tracked", (), "int64"))
    out.append(("classifier.1.weight", (spec.num_classes, LAST_CHANNELS), "float32"))
    out.append(("classifier.1.bias", (spec.num_classes,), "float32"))
    return out


def parameter_keys(spec: NetSpec) -> List[Tuple[str, Tuple[int, ...]]]:
    """Trainable tensors in ``model.parameters()`` order (158 tensors)."""
    return [(k, s) for k, s, _ in state_dict_keys(spec)
            if not k.endswith(("running_mean", "running_var", "num_batches_tracked"))]
