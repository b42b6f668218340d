"""Datasets (§7.1 step 2).

The reference downloads CIFAR-10 through torchvision (cifar10_mpi_mobilenet_224.py:93-109).
There is no network here, so two sources produce the same in-memory form -- ``images`` uint8
``[N, 32, 32, 3]`` (HWC, as CIFAR stores them) and ``labels`` int64 ``[N]``:

* ``load_cifar10(root)`` reads the standard ``cifar-10-batches-py`` pickles if they are already
  under ``root`` (what ``download=False`` does at ref :104-109);
* ``SyntheticCIFAR10`` is a deterministic, *learnable* stand-in of the same shape: each class has
  a smooth colour template that is shifted, scaled and buried in noise per sample, so loss and
  accuracy trajectories mean something.

Keeping the dataset as one uint8 tensor (150 MB for 50k images) lets the loader gather a batch
with one ``index_select`` and move 393 KB per step over PCIe instead of the reference's 77 MB of
pre-resized fp32; resize / augmentation / normalisation run on the device.
"""
from __future__ import annotations

import os
import pickle
from dataclasses import dataclass
from typing import Optional, Tuple

import numpy as np
import torch


@dataclass
class ArrayDataset:
    images: torch.Tensor      # uint8 [N, 32, 32, 3]
    labels: torch.Tensor      # int64 [N]
    name: str = "array"

    def __len__(self) -> int:
        return int(self.images.shape[0])

    def __getitem__(self, i: int) -> Tuple[torch.Tensor, int]:
        return self.images[i], int(self.labels[i])

    def pin(self) -> "ArrayDataset":
        if torch.cuda.is_available():
            self.images = self.images.pin_memory()
            self.labels = self.labels.pin_memory()
        return self

    def to(self, device) -> "ArrayDataset":
        return ArrayDataset(self.images.to(device), self.labels.to(device), self.name)


def _class_templates(num_classes: int, seed: int) -> np.ndarray:
    """One smooth 32x32x3 pattern per class, built from a few low-frequency cosines."""
    rng = np.random.RandomState(seed)
    yy, xx = np.meshgrid(np.arange(32, dtype=np.float32), np.arange(32, dtype=np.float32), indexing="ij")
    t = np.zeros((num_classes, 32, 32, 3), dtype=np.float32)
    for k in range(num_classes):
        for ch in range(3):
            acc = np.zeros((32, 32), dtype=np.float32)
            for _ in range(3):
                fx, fy = rng.uniform(0.5, 3.0, size=2)
                ph = rng.uniform(0, 2 * np.pi)
                acc += rng.uniform(0.5, 1.0) * np.cos(2 * np.pi * (fx * xx + fy * yy) / 32.0 + ph)
            t[k, :, :, ch] = acc / 3.0
    return t


def SyntheticCIFAR10(n: int, train: bool = True, num_classes: int = 10, seed: int = 1234,
                     noise: float = 0.35) -> ArrayDataset:
    """Deterministic CIFAR-shaped data.  Train and test share class templates but not samples."""
    templates = _class_templates(num_classes, seed)
    rng = np.random.RandomState(seed + (1 if train else 2))
    labels = rng.randint(0, num_classes, size=n).astype(np.int64)
    images = np.empty((n, 32, 32, 3), dtype=np.uint8)
    chunk = 4096
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        lab = labels[s:e]
        base = templates[lab]                                  # [m,32,32,3]
        dx = rng.randint(-3, 4, size=e - s)
        dy = rng.randint(-3, 4, size=e - s)
        amp = rng.uniform(0.6, 1.0, size=(e - s, 1, 1, 1)).astype(np.float32)
        shifted = np.empty_like(base)
        for i in range(e - s):
            shifted[i] = np.roll(base[i], (int(dy[i]), int(dx[i])), axis=(0, 1))
        x = 0.5 + 0.28 * amp * shifted + noise * 0.25 * rng.standard_normal(base.shape).astype(np.float32)
        images[s:e] = np.clip(x * 255.0 + 0.5, 0, 255).astype(np.uint8)
    name = f"synthetic-cifar10-{'train' if train else 'test'}-{n}"
    return ArrayDataset(torch.from_numpy(images), torch.from_numpy(labels), name)


def cifar10_present(root: str) -> bool:
    d = os.path.join(root, "cifar-10-batches-py")
    return os.path.isfile(os.path.join(d, "data_batch_1")) and os.path.isfile(os.path.join(d, "test_batch"))


def load_cifar10(root: str, train: bool = True) -> ArrayDataset:
    """Read the python-pickle CIFAR-10 batches (no download; ref :104-109 ``download=False``)."""
    d = os.path.join(root, "cifar-10-batches-py")
    files = [f"data_batch_{i}" for i in range(1, 6)] if train else ["test_batch"]
    xs, ys = [], []
    for fn in files:
        with open(os.path.join(d, fn), "rb") as f:
            entry = pickle.load(f, encoding="latin1")
        xs.append(np.asarray(entry["data"], dtype=np.uint8))
        ys.extend(entry.get("labels", entry.get("fine_labels")))
    x = np.concatenate(xs).reshape(-1, 3, 32, 32).transpose(0, 2, 3, 1).copy()   # -> HWC
    return ArrayDataset(torch.from_numpy(x), torch.tensor(ys, dtype=torch.int64),
                        f"cifar10-{'train' if train else 'test'}")


def get_datasets(root: str = "./data", source: str = "auto", train_samples: int = 50_000,
                 test_samples: int = 10_000, seed: int = 1234) -> Tuple[ArrayDataset, ArrayDataset]:
    """Train / test pair.  ``auto`` prefers real CIFAR-10 when it is on disk."""
    if source == "cifar10" or (source == "auto" and cifar10_present(root)):
        return load_cifar10(root, True), load_cifar10(root, False)
    if source not in ("auto", "synthetic"):
        raise ValueError(f"unknown dataset source {source!r}")
    return (SyntheticCIFAR10(train_samples, True, seed=seed),
            SyntheticCIFAR10(test_samples, False, seed=seed))
