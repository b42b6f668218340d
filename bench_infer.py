#!/usr/bin/env python
"""Inference benchmark (BASELINE.json config 5): batch-1 latency and batch-256 throughput on one B200.

    python bench_infer.py [--calls 1000] [--batches 50]

The path measured is the reference's predict helper (cifar10_128batch.py:159-188; Gradio app GROUP03.pdf p.26-27):
image -> Resize((224,224)) + ToTensor + Normalize -> MobileNetV2 (eval) -> softmax -> top-3.
Arms (same box, random-init weights, one 500x375 synthetic photograph):
  * native        -- ops.resize_normalize + NativeInference (CUDA-graph replay) + ops.softmax_topk; batch-1 latency is
                     host wall clock per call including the 562 KB H2D copy of the uint8 image and the D2H read of top-3;
  * native-device -- the same without the image upload (image already on the device): the kernel-side latency;
  * torchvision   -- torchvision.models.mobilenet_v2 eval on the GPU, fp32 and bf16 autocast channels_last, input
                     resized by PIL on the CPU as the reference does (latency), or already resident (throughput).
Prints one JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pct(xs, q):
    xs = sorted(xs)
    return xs[min(len(xs) - 1, int(q * len(xs)))]


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--calls", type=int, default=1000)
    ap.add_argument("--batches", type=int, default=50)
    ns = ap.parse_args()
    import numpy as np
    import torch
    import torchvision
    from PIL import Image

    from b200dp import ops
    from b200dp.engine_native import NativeInference
    from b200dp.inference import reference_transform

    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    tv = torchvision.models.mobilenet_v2(weights=None)
    tv.classifier[1] = torch.nn.Linear(tv.last_channel, 10)
    tv.eval().to(dev)
    sd = {k: v.cpu() for k, v in tv.state_dict().items()}
    rng = np.random.RandomState(0)
    photo = rng.randint(0, 256, size=(375, 500, 3), dtype=np.uint8)
    photo_t = torch.from_numpy(photo)
    photo_pin = photo_t.pin_memory()
    out = {"metric": "MobileNetV2@224 inference: batch-1 latency (ms) and batch-256 images/sec", "n_gpus": 1,
           "data": "synthetic 500x375 photograph, random-init weights", "calls": ns.calls}

    # ---------------- batch 1 latency ----------------
    n1 = NativeInference(10, dev, sd, batch=1)
    photo_dev = photo_t.to(dev)

    def native_call(upload: bool):
        src = photo_pin.to(dev, non_blocking=True) if upload else photo_dev
        x = ops.resize_normalize(src[None], 224)
        lg = n1.logits_preprocessed(x, clone=False)
        _, conf, idx = ops.softmax_topk(lg, 3)
        return conf.cpu(), idx.cpu()          # D2H + sync

    def tv_call(dtype):
        x = reference_transform(photo_t)[None].to(dev)           # PIL resize on the CPU, as the reference does
        with torch.no_grad():
            if dtype == "bf16":
                with torch.autocast("cuda", dtype=torch.bfloat16):
                    lg = tv(x.contiguous(memory_format=torch.channels_last)).float()
            else:
                lg = tv(x)
            conf, idx = torch.topk(torch.softmax(lg, 1)[0], 3)
        return conf.cpu(), idx.cpu()

    def lat(fn, calls):
        for _ in range(20):
            fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(calls):
            t0 = time.perf_counter()
            fn()
            ts.append((time.perf_counter() - t0) * 1e3)
        return {"p50_ms": pct(ts, 0.5), "p99_ms": pct(ts, 0.99), "mean_ms": sum(ts) / len(ts)}

    out["batch1"] = {
        "native": lat(lambda: native_call(True), ns.calls),
        "native_device_input": lat(lambda: native_call(False), ns.calls),
        "torchvision_fp32": lat(lambda: tv_call("fp32"), max(ns.calls // 5, 50)),
        "torchvision_bf16": lat(lambda: tv_call("bf16"), max(ns.calls // 5, 50)),
    }
    n1.close()

    # ---------------- batch 256 throughput (device-resident input, CUDA events) ----------------
    B = 256
    n256 = NativeInference(10, dev, sd, batch=B)
    photos = torch.from_numpy(rng.randint(0, 256, size=(B, 375, 500, 3), dtype=np.uint8)).to(dev)
    x_pre = ops.resize_normalize(photos, 224)
    cifar = torch.from_numpy(rng.randint(0, 256, size=(B, 32, 32, 3), dtype=np.uint8)).to(dev)
    x_tv = x_pre[..., :3].permute(0, 3, 1, 2).float().contiguous()
    x_tv_cl = x_tv.contiguous(memory_format=torch.channels_last)

    def thr(fn, iters):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            fn()
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / iters
        return {"ms_per_batch": ms, "images_per_sec": B / ms * 1e3}

    def native_photo():
        x = ops.resize_normalize(photos, 224)
        lg = n256.logits_preprocessed(x, clone=False)
        ops.softmax_topk(lg, 3)

    def native_cifar():
        lg = n256.logits(cifar)
        ops.softmax_topk(lg, 3)

    def tv_fp32():
        with torch.no_grad():
            torch.topk(torch.softmax(tv(x_tv), 1), 3)

    def tv_bf16():
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
            torch.topk(torch.softmax(tv(x_tv_cl).float(), 1), 3)

    out["batch256"] = {
        "native_photos_500x375": thr(native_photo, ns.batches),
        "native_cifar_32x32": thr(native_cifar, ns.batches),
        "torchvision_fp32_resident_input": thr(tv_fp32, max(ns.batches // 2, 10)),
        "torchvision_bf16_channels_last_resident_input": thr(tv_bf16, max(ns.batches // 2, 10)),
    }
    n256.close()
    print(json.dumps(out), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
