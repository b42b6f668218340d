"""Strong torch baseline (BASELINE.md §3 variant b): torchvision MobileNetV2 + 10-class head, bf16 autocast,
channels_last, device-resident synthetic 224x224 batches, DDP + NCCL + torch.optim.Adam, per-rank batch 128.
Prints one JSON line from rank 0.  Launch: python baseline/torch_ddp_baseline.py  |  torchrun ... --nproc-per-node N.
"""
import json
import os
import sys

import torch
import torch.distributed as dist
import torch.nn as nn
import torchvision


def main():
    steps = int(os.environ.get("STEPS", "30")); warm = int(os.environ.get("WARMUP", "5")); B = 128
    fp32 = os.environ.get("FP32", "0") == "1"
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    model = torchvision.models.mobilenet_v2(weights=None)
    model.classifier[1] = nn.Linear(model.classifier[1].in_features, 10)
    model = model.cuda()
    if not fp32:
        model = model.to(memory_format=torch.channels_last)
    if world > 1:
        model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local])
    opt = torch.optim.Adam(model.parameters(), lr=1e-4)
    crit = nn.CrossEntropyLoss()
    xs = [torch.randn(B, 3, 224, 224, device="cuda") for _ in range(4)]
    if not fp32:
        xs = [x.contiguous(memory_format=torch.channels_last) for x in xs]
    ys = [torch.randint(0, 10, (B,), device="cuda") for _ in range(4)]

    def step(i):
        opt.zero_grad()
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=not fp32):
            out = model(xs[i % 4])
            loss = crit(out.float(), ys[i % 4])
        loss.backward()
        opt.step()

    for i in range(warm):
        step(i)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        step(i)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(json.dumps({"impl": "torch_ddp_strong" if not fp32 else "torch_ddp_fp32", "n_gpus": world,
                          "value": world * steps * B / (ms.item() / 1e3), "unit": "images/sec",
                          "ms_per_step": ms.item() / steps, "dtype": "fp32" if fp32 else "bf16 autocast, channels_last"}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    sys.exit(main())
