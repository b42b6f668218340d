"""Accuracy-trajectory check: the same short fine-tuning run on the native engine and on the torch engine
(cuDNN/ATen bf16) -- same synthetic learnable CIFAR-shaped data, same seed, same hyper-parameters."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200dp.config import TrainConfig
from b200dp.trainer import Trainer
from b200dp.utils.logging import RankLogger

out = {}
for engine in os.environ.get("ENGINES", "native,torch").split(","):
    cfg = TrainConfig(mode="gpu", engine=engine, epochs=int(os.environ.get("EPOCHS", "4")), lr=float(os.environ.get("LR", "1e-3")), lr_step_size=int(os.environ.get("LRSTEP", "3")), seed=int(os.environ.get("SEED", "42")),
                      train_samples=int(os.environ.get("TRAIN", "12800")), test_samples=2560, dataset="synthetic",
                      out_dir="/tmp/traj_" + engine, augment=os.environ.get("AUG", "1") == "1")
    os.makedirs(cfg.out_dir, exist_ok=True)
    t = Trainer(cfg)
    t0 = time.time()
    res = t.fit()
    t.close()
    out[engine] = {"history": res["history"], "best_acc": res["best_acc"], "seconds": time.time() - t0}
print(json.dumps(out))
