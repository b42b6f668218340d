"""Step-by-step loss of the first training steps: native engine vs torch engine (fused / torch Adam)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200dp.config import TrainConfig
from b200dp.data.datasets import SyntheticCIFAR10
from b200dp.data.loader import Batch
from b200dp.data.transforms import identity_params, sample_aug_params
from b200dp.parallel.bootstrap import DistContext, RankInfo
from b200dp.trainer import make_engine

B, STEPS = 128, int(os.environ.get("STEPS", "40"))
ds = SyntheticCIFAR10(B * STEPS, True)
ctx = DistContext(RankInfo(0, 1, 0, "single"), torch.device("cuda", 0), "nccl", False)
res = {}
for name, kw in (("native", dict(engine="native")), ("torch+fused", dict(engine="torch")), ("torch+flatadam", dict(engine="torch", comm="nccl"))):
    cfg = TrainConfig(mode="gpu", batch_size=B, lr=float(os.environ.get("LR", "1e-4")), dropout=float(os.environ.get("DROPOUT", "0")), augment=os.environ.get("AUG", "0") == "1", **kw)
    eng = make_engine(cfg, ctx)
    losses = []
    for i in range(STEPS):
        sl = slice(i * B, (i + 1) * B)
        g = torch.Generator().manual_seed(i)
        par = sample_aug_params(B, g) if cfg.augment else identity_params(B)
        b = Batch(ds.images[sl].cuda(), ds.labels[sl].cuda(), par.cuda(), B)
        eng.train_step(b, cfg.lr)
        l, c, n = eng.read_metrics()
        losses.append(l / n)
    res[name] = losses
    eng.close()
for i in range(0, STEPS, max(1, STEPS // 16)):
    print(i, "  ".join(f"{k}={v[i]:.4f}" for k, v in res.items()))
