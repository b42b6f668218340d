"""Per-layer diagnostic: native executor vs PyTorch.

fp32 PyTorch at random init is a chaotic reference (a 0.2 % bf16 rounding grows ~1.3x per block), so the
gradient check is *teacher forced*: the torch forward is evaluated at the native activations through a
straight-through substitution  y = native + (torch - torch.detach()),  which makes every Jacobian of the
torch backward be taken at the same point as the native backward.
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from b200dp import ops, engine_native as en
from b200dp.models.mobilenetv2 import MobileNetV2
from b200dp.data.datasets import SyntheticCIFAR10
from b200dp.testing import teacher_forced_backward, cos

B = int(os.environ.get("DIAG_B", "16"))
torch.manual_seed(0)
ref = MobileNetV2(num_classes=10, dropout=0.0).cuda()
with torch.no_grad():
    for m in ref.modules():
        if isinstance(m, torch.nn.BatchNorm2d):
            m.weight.uniform_(0.5, 1.5); m.bias.uniform_(-0.2, 0.2)
core = en.NativeCore(torch.device("cuda", 0), B, 224, 10, 0.0, 123, use_graphs=False)
core.load_state_dict(ref.state_dict())
ds = SyntheticCIFAR10(B, True)
images, labels = ds.images.cuda(), ds.labels.cuda()
x = ops.preprocess(images, None, channels=3).permute(0, 3, 1, 2).float()
core.train_step(images, labels, None, B); torch.cuda.synchronize()
loss, fwd_err = teacher_forced_backward(ref, core, x, labels, B)
print("loss native", core.metrics_train[0].item() / B, "teacher-forced torch", loss)
print("--- per-layer forward error (torch conv on the native input vs native output)")
for name, e in fwd_err:
    print(f"{name:34s} relerr={e:.5f}")
print("--- gradients (teacher forced)")
g = core.layout.export_state_dict(core.grads, core.buffers)
for k, p in ref.named_parameters():
    a, b = g[k].cuda(), p.grad
    print(f"{k:40s} cos={cos(a, b):.4f} ratio={(a.norm() / (b.norm() + 1e-20)).item():.3f} |ref|={b.norm().item():.3e}")
# context: how far is torch's own bf16 autocast from fp32 on this random-init network?
ref32 = MobileNetV2(num_classes=10, dropout=0.0).cuda(); ref32.load_state_dict(ref.state_dict())
ref.zero_grad(); ref.train(); ref32.train()
F.cross_entropy(ref32(x), labels).backward()
with torch.autocast("cuda", dtype=torch.bfloat16):
    l16 = F.cross_entropy(ref(x).float(), labels)
l16.backward()
cs = [cos(p.grad, q.grad) for p, q in zip(ref.parameters(), ref32.parameters()) if q.grad.norm() > 1e-4]
print(f"context: torch bf16-autocast vs fp32 gradient cosine: min={min(cs):.3f} median={sorted(cs)[len(cs)//2]:.3f}")
