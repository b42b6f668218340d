import os
import pickle

import numpy as np
import torch

from b200dp.data.datasets import SyntheticCIFAR10, get_datasets, load_cifar10
from b200dp.data.loader import BatchLoader
from b200dp.parallel.sampler import PartitionSampler


def test_synthetic_is_deterministic_and_cifar_shaped():
    a = SyntheticCIFAR10(256, True)
    b = SyntheticCIFAR10(256, True)
    c = SyntheticCIFAR10(256, False)
    assert a.images.shape == (256, 32, 32, 3) and a.images.dtype == torch.uint8
    assert a.labels.dtype == torch.int64 and int(a.labels.max()) <= 9
    assert torch.equal(a.images, b.images) and torch.equal(a.labels, b.labels)
    assert not torch.equal(a.images, c.images)


def test_cifar10_pickle_reader(tmp_path):
    d = tmp_path / "cifar-10-batches-py"
    d.mkdir()
    rng = np.random.RandomState(0)
    for name in [f"data_batch_{i}" for i in range(1, 6)] + ["test_batch"]:
        with open(d / name, "wb") as f:
            pickle.dump({"data": rng.randint(0, 255, (20, 3072), dtype=np.uint8), "labels": list(rng.randint(0, 10, 20))}, f)
    tr = load_cifar10(str(tmp_path), True)
    te = load_cifar10(str(tmp_path), False)
    assert len(tr) == 100 and len(te) == 20 and tr.images.shape[1:] == (32, 32, 3)
    tr2, _ = get_datasets(str(tmp_path), "auto")
    assert tr2.name.startswith("cifar10")


def test_loader_covers_rank_partition_with_ragged_tail():
    ds = SyntheticCIFAR10(300, True)
    s = PartitionSampler(len(ds), 2, 1, shuffle=True)
    ld = BatchLoader(ds, s, 64, torch.device("cpu"), train=True, augment=True)
    ld.set_epoch(3)
    sizes, labels = [], []
    for b in ld:
        assert b.images.shape == (b.n, 32, 32, 3) and b.params.shape == (b.n, 12)
        sizes.append(b.n); labels.append(b.labels)
    assert sizes == [64, 64, 22] and len(ld) == 3
    s.set_epoch(3)
    assert torch.equal(torch.cat(labels), ds.labels[s.indices()])
