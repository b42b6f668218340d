"""Two un-graphed native training steps at batch 128 (target for `ncu --metrics gpu__time_duration.sum`)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200dp import engine_native as en
from b200dp.models.mobilenetv2 import MobileNetV2
from b200dp.data.datasets import SyntheticCIFAR10
from b200dp.data.transforms import sample_aug_params
B = int(os.environ.get("PROF_B", "128"))
core = en.NativeCore(torch.device("cuda", 0), B, 224, 10, 0.2, 1, use_graphs=False)
core.load_state_dict(MobileNetV2(10).state_dict())
ds = SyntheticCIFAR10(B, True)
images, labels, aug = ds.images.cuda(), ds.labels.cuda(), sample_aug_params(B).cuda()
for _ in range(int(os.environ.get("STEPS", "2"))):
    core.grads.zero_()
    core.train_step(images, labels, aug, B)
torch.cuda.synchronize()
print("done")
