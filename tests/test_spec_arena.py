"""Architecture table and flat-arena <-> torchvision state_dict."""
import pytest
import torch

from b200dp.models.arena import ArenaLayout
from b200dp.models.mobilenetv2 import MobileNetV2
from b200dp.models.spec import build_spec, parameter_keys, state_dict_keys


def test_spec_counts():
    sp = build_spec()
    assert len(sp.convs) == 52
    assert sum(c.kind == "dw3x3" for c in sp.convs) == 17
    assert sum(c.kind == "pw1x1" for c in sp.convs) == 34
    assert sum(b.use_res for b in sp.blocks) == 10
    assert sum(c.relu6 for c in sp.convs) == 35
    assert sp.param_count() == 2_236_682            # cifar_mpi_gpu128_26188.out:30
    assert abs(sp.macs_per_image() / 1e6 - 299.5) < 0.1
    assert sp.feature_hw == 7


def test_state_dict_keys_match_torchvision():
    tv = pytest.importorskip("torchvision")
    ref = tv.models.mobilenet_v2(weights=None)
    ref.classifier[1] = torch.nn.Linear(ref.classifier[1].in_features, 10)
    sp = build_spec()
    ours = state_dict_keys(sp)
    theirs = [(k, tuple(v.shape), str(v.dtype).replace("torch.", "")) for k, v in ref.state_dict().items()]
    assert len(ours) == 314
    assert ours == theirs
    assert [k for k, _ in parameter_keys(sp)] == [k for k, _ in ref.named_parameters()]


def test_own_model_strict_loads_into_torchvision_and_back():
    tv = pytest.importorskip("torchvision")
    ours = MobileNetV2(num_classes=10)
    ref = tv.models.mobilenet_v2(weights=None)
    ref.classifier[1] = torch.nn.Linear(1280, 10)
    ref.load_state_dict(ours.state_dict(), strict=True)
    ours.load_state_dict(ref.state_dict(), strict=True)
    ours.eval(); ref.eval()
    x = torch.randn(2, 3, 64, 64)
    with torch.no_grad():
        assert torch.allclose(ours(x), ref(x), atol=1e-5)


def test_arena_roundtrip_identity():
    torch.manual_seed(0)
    m = MobileNetV2(num_classes=10)
    for b in m.buffers():
        if b.dtype.is_floating_point:
            b.uniform_(0.5, 1.5)
    sd = m.state_dict()
    L = ArenaLayout()
    params, buffers, nbt = L.import_state_dict(sd)
    assert params.numel() == L.total and L.total % 64 == 0
    out = L.export_state_dict(params, buffers, 7)
    assert list(out.keys()) == list(sd.keys())
    for k, v in sd.items():
        if k.endswith("num_batches_tracked"):
            assert int(out[k]) == 7 and out[k].dtype == torch.int64
        else:
            assert out[k].dtype == torch.float32 and out[k].shape == v.shape
            assert torch.equal(out[k], v), k
    m2 = MobileNetV2(num_classes=10)
    m2.load_state_dict(out, strict=True)


def test_arena_padding_is_zero_and_mask_counts_params():
    L = ArenaLayout()
    m = MobileNetV2(num_classes=10)
    params, _, _ = L.import_state_dict(m.state_dict())
    mask = L.trainable_mask()
    assert int(mask.sum().item()) == 2_236_682
    assert torch.all(params[mask == 0] == 0)
    for s in L.segments:
        assert s.offset % 64 == 0
    assert len(L.offsets_table()) == 52 * 5 + 2
