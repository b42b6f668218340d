"""Dataset partition across ranks.

Index-for-index equal to ``torch.utils.data.DistributedSampler`` as the reference uses it
(cifar10_mpi_mobilenet_224.py:119-124: ``shuffle=True`` for train, ``False`` for test,
``set_epoch(epoch)`` at :165), but it hands out whole *batches of indices* as int64 tensors so
the loader can gather a batch with one ``index_select`` on the uint8 dataset (host or device)
instead of per-sample Python work.
"""
from __future__ import annotations

import math
from typing import Iterator, List

import torch


class PartitionSampler:
    def __init__(self, dataset_len: int, num_replicas: int = 1, rank: int = 0,
                 shuffle: bool = True, seed: int = 0, drop_last: bool = False) -> None:
        if not (0 <= rank < num_replicas):
            raise ValueError(f"rank {rank} outside [0, {num_replicas})")
        self.dataset_len = int(dataset_len)
        self.num_replicas = int(num_replicas)
        self.rank = int(rank)
        self.shuffle = shuffle
        self.seed = int(seed)
        self.drop_last = drop_last
        self.epoch = 0
        if drop_last and self.dataset_len % self.num_replicas != 0:
            self.num_samples = math.ceil((self.dataset_len - self.num_replicas) / self.num_replicas)
        else:
            self.num_samples = math.ceil(self.dataset_len / self.num_replicas)
        self.total_size = self.num_samples * self.num_replicas

    def set_epoch(self, epoch: int) -> None:
        self.epoch = int(epoch)

    def indices(self) -> torch.Tensor:
        """This rank's sample indices for the current epoch (int64, length ``num_samples``)."""
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            idx = torch.randperm(self.dataset_len, generator=g)
        else:
            idx = torch.arange(self.dataset_len)
        if not self.drop_last:
            pad = self.total_size - idx.numel()
            if pad > 0:
                if pad <= idx.numel():
                    idx = torch.cat([idx, idx[:pad]])
                else:
                    reps = math.ceil(pad / idx.numel())
                    idx = torch.cat([idx, idx.repeat(reps)[:pad]])
        else:
            idx = idx[: self.total_size]
        return idx[self.rank: self.total_size: self.num_replicas].contiguous()

    def __iter__(self) -> Iterator[int]:
        return iter(self.indices().tolist())

    def __len__(self) -> int:
        return self.num_samples

    def batches(self, batch_size: int, drop_last: bool = False) -> List[torch.Tensor]:
        """Split this rank's indices into batches; the last one is ragged unless ``drop_last``
        (the reference keeps it: 391 steps with a last batch of 80 at W=1, 196/40 at W=2, 49/106 at W=8)."""
        idx = self.indices()
        out = list(torch.split(idx, batch_size))
        if drop_last and out and out[-1].numel() < batch_size:
            out.pop()
        return out

    def num_batches(self, batch_size: int, drop_last: bool = False) -> int:
        if drop_last:
            return self.num_samples // batch_size
        return math.ceil(self.num_samples / batch_size)
