"""Three un-graphed native training steps at batch 128 (every kernel is its own launch), for `ncu` launch lists:
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
    --log-file gpurun_out/launches.csv python scripts/one_step.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200dp import engine_native as en
from b200dp.data.datasets import SyntheticCIFAR10
from b200dp.data.transforms import sample_aug_params
from b200dp.models.mobilenetv2 import MobileNetV2
from b200dp.optim import AdamHyper
from b200dp.parallel.bootstrap import DistContext, RankInfo
from b200dp.parallel.comm import FusedAllReduceAdam

B = int(os.environ.get("PROF_B", "128"))
dev = torch.device("cuda", 0)
ctx = DistContext(RankInfo(0, 1, 0, "single"), dev, "nccl", False)
from b200dp.models.arena import ArenaLayout
L = ArenaLayout()
fused = FusedAllReduceAdam(ctx, L.total, AdamHyper(1e-4), extra_numel=L.buffer_total)
core = en.NativeCore(dev, B, 224, 10, 0.2, 1, use_graphs=False, grads=fused.grads, params=fused.params, buffers=fused.extra)
core.load_state_dict(MobileNetV2(10).state_dict())
core.set_comm(fused, [], 148)
ds = SyntheticCIFAR10(B, True)
images, labels, aug = ds.images.cuda(), ds.labels.cuda(), sample_aug_params(B).cuda()
for _ in range(int(os.environ.get("PROF_STEPS", "3"))):
    core.train_step(images, labels, aug, B)
torch.cuda.synchronize()
print("done", core.metrics_train.tolist())
