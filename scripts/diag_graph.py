import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200dp import engine_native as en
from b200dp.models.mobilenetv2 import MobileNetV2
from b200dp.data.datasets import SyntheticCIFAR10
B = 8
torch.manual_seed(0)
sd = MobileNetV2(10, dropout=0.0).state_dict()
def mk(graphs):
    c = en.NativeCore(torch.device("cuda", 0), B, 224, 10, 0.0, 123, use_graphs=graphs)
    c.load_state_dict(sd); return c
ds = SyntheticCIFAR10(B, True); images, labels = ds.images.cuda(), ds.labels.cuda()
a, b, g = mk(False), mk(False), mk(True)
for it in range(3):
    for c in (a, b, g):
        c.grads.zero_(); c.train_step(images, labels, None, B)
    torch.cuda.synchronize()
    rel = lambda x, y: ((x - y).norm() / x.norm()).item()
    print(f"iter {it}: direct-vs-direct {rel(a.grads, b.grads):.4e}  graph-vs-direct {rel(a.grads, g.grads):.4e}",
          "loss", a.metrics_train[0].item(), b.metrics_train[0].item(), g.metrics_train[0].item())
    L = a.layout
    if it == 2:
        for s in L.segments[:12] + L.segments[-8:]:
            x, y = a.grads[s.offset:s.offset + s.numel], g.grads[s.offset:s.offset + s.numel]
            z = b.grads[s.offset:s.offset + s.numel]
            print(f"  {s.key:36s} d-d {rel(x, z):.3e} g-d {rel(x, y):.3e}")
