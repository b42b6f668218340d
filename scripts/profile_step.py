"""Per-kernel timing of one native training step (CUDA events between launches, no graph)."""
import os, sys, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200dp import engine_native as en
from b200dp.data.datasets import SyntheticCIFAR10
from b200dp.data.transforms import sample_aug_params

B = int(os.environ.get("PROF_B", "128"))
core = en.NativeCore(torch.device("cuda", 0), B, 224, 10, 0.2, 1, use_graphs=False)
from b200dp.models.mobilenetv2 import MobileNetV2
core.load_state_dict(MobileNetV2(10).state_dict())
ds = SyntheticCIFAR10(B, True)
images, labels, aug = ds.images.cuda(), ds.labels.cuda(), sample_aug_params(B).cuda()
for _ in range(3):
    core.grads.zero_(); core.train_step(images, labels, aug, B)
torch.cuda.synchronize()
core.grads.zero_()
rows = core.profile_step(images, labels, aug, B)
spec = core.layout.spec
total = sum(ms for _, ms in rows)
print(f"total {total:.3f} ms over {len(rows)} marks (B={B})")
agg = collections.OrderedDict()
for name, ms in rows:
    key = name.split("#")[0]
    agg.setdefault(key, [0.0, 0]); agg[key][0] += ms; agg[key][1] += 1
print("--- by kind")
for k, (ms, n) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    print(f"{k:22s} {ms:8.3f} ms  {100 * ms / total:5.1f}%  x{n}")
print("--- all marks")
for name, ms in rows:
    extra = ""
    if "#" in name:
        c = spec.convs[int(name.split("#")[1])]
        extra = f"{c.role:8s} {c.cin:4d}->{c.cout:4d} {c.hin:3d}->{c.hout:3d}"
    print(f"{name:24s} {ms:7.3f}  {extra}")
