"""CPU image pipeline vs the reference's own CPU path, torchvision/PIL."""
import math

import numpy as np
import pytest
import torch

from b200dp.data import transforms as T
from b200dp.data.datasets import SyntheticCIFAR10

tv = pytest.importorskip("torchvision")
from PIL import Image  # noqa: E402
import torchvision.transforms as TT  # noqa: E402
import torchvision.transforms.functional as TF  # noqa: E402


@pytest.fixture(scope="module")
def images():
    return SyntheticCIFAR10(6, train=True).images


def _norm(pil):
    t = TF.to_tensor(pil)
    return TF.normalize(t, [0.485, 0.456, 0.406], [0.229, 0.224, 0.225])


def test_test_transform_matches_torchvision(images):
    ours = T.apply_pipeline_torch(images, None)
    ref_tf = TT.Compose([TT.Resize((224, 224)), TT.ToTensor(),
                         TT.Normalize([0.485, 0.456, 0.406], [0.229, 0.224, 0.225])])   # ref :84-89
    for i in range(images.shape[0]):
        ref = ref_tf(Image.fromarray(images[i].numpy()))
        diff = (ours[i] - ref).abs() * 0.226 * 255      # back to 8-bit levels
        assert diff.max() <= 1.01, diff.max()           # PIL rounds twice (two passes), we round once
        assert (diff > 0.5).float().mean() < 0.25   # 1-level differences only (PIL rounds after each pass)


def test_geometry_matches_pil_with_fixed_params(images):
    """crop box + flip + rotation with hand-picked parameters vs the PIL functional ops."""
    B = images.shape[0]
    p = T.identity_params(B)
    p[:, T.P_ENABLE] = 1
    boxes = [(10, 20, 180, 200), (0, 0, 224, 224), (30, 5, 150, 190), (3, 40, 200, 160), (12, 12, 199, 201), (0, 24, 224, 188)]
    angles = [0.0, 12.5, -15.0, 7.0, -3.3, 14.9]
    for i, (top, left, h, w) in enumerate(boxes):
        p[i, T.P_TOP], p[i, T.P_LEFT], p[i, T.P_H], p[i, T.P_W] = top, left, h, w
        p[i, T.P_FLIP] = i % 2
        p[i, T.P_ANGLE] = angles[i]
    ours = T.apply_pipeline_torch(images, p)
    for i, (top, left, h, w) in enumerate(boxes):
        pil = Image.fromarray(images[i].numpy()).resize((224, 224), Image.BILINEAR)
        pil = TF.resized_crop(pil, top, left, h, w, [224, 224])
        if i % 2:
            pil = TF.hflip(pil)
        pil = TF.rotate(pil, angles[i])
        ref = _norm(pil)
        diff = (ours[i] - ref).abs() * 0.226 * 255
        # nearest-neighbour rotation can pick the adjacent pixel at exact .5 boundaries
        assert (diff > 2.5).float().mean() < 0.01, (i, float((diff > 2.5).float().mean()))


@pytest.mark.parametrize("op", ["brightness", "contrast", "saturation", "hue"])
def test_color_ops_match_pil(images, op):
    B = images.shape[0]
    p = T.identity_params(B)
    p[:, T.P_ENABLE] = 1
    col = {"brightness": T.P_BRIGHT, "contrast": T.P_CONTRAST, "saturation": T.P_SAT, "hue": T.P_HUE}[op]
    vals = [0.7, 1.3, 1.0, 0.85, 1.2, 0.95] if op != "hue" else [-0.1, 0.1, 0.0, 0.05, -0.03, 0.08]
    p[:, col] = torch.tensor(vals)
    ours = T.apply_pipeline_torch(images, p)
    fn = {"brightness": TF.adjust_brightness, "contrast": TF.adjust_contrast,
          "saturation": TF.adjust_saturation, "hue": TF.adjust_hue}[op]
    tol = 8.0 if op == "hue" else 3.0       # PIL's HSV round trip is 8-bit per channel
    for i in range(B):
        pil = Image.fromarray(images[i].numpy()).resize((224, 224), Image.BILINEAR)
        ref = _norm(fn(pil, vals[i]))
        diff = (ours[i] - ref).abs() * 0.226 * 255
        assert diff.mean() < 1.0 and (diff > tol).float().mean() < 0.01, (op, i, float(diff.mean()), float(diff.max()))


def test_sampled_params_are_in_reference_ranges():
    g = torch.Generator().manual_seed(0)
    p = T.sample_aug_params(4096, g)
    area = p[:, T.P_H] * p[:, T.P_W] / (224 * 224)
    assert area.min() >= 0.69 and area.max() <= 1.0 + 1e-6             # scale=(0.7, 1.0), ref :74
    ratio = p[:, T.P_W] / p[:, T.P_H]
    assert ratio.min() >= 0.73 and ratio.max() <= 1.37                 # ratio 3/4..4/3
    assert (p[:, T.P_TOP] + p[:, T.P_H] <= 224).all() and (p[:, T.P_LEFT] + p[:, T.P_W] <= 224).all()
    assert 0.4 < p[:, T.P_FLIP].mean() < 0.6
    for c in (T.P_BRIGHT, T.P_CONTRAST, T.P_SAT):
        assert p[:, c].min() >= 0.7 and p[:, c].max() <= 1.3           # ref :76
    assert p[:, T.P_HUE].abs().max() <= 0.1 and p[:, T.P_ANGLE].abs().max() <= 15.0   # ref :77-78
    assert set(p[:, T.P_ORDER].long().tolist()) == set(range(24))


def test_identity_params_equal_test_transform(images):
    a = T.apply_pipeline_torch(images, None)
    p = T.identity_params(images.shape[0]); p[:, T.P_ENABLE] = 1
    b = T.apply_pipeline_torch(images, p)
    assert (a - b).abs().max() < 1e-4
