"""Partition sampler == torch DistributedSampler, index for index."""
import pytest
import torch
from torch.utils.data.distributed import DistributedSampler

from b200dp.parallel.sampler import PartitionSampler


class _DS(torch.utils.data.Dataset):
    def __init__(self, n): self.n = n
    def __len__(self): return self.n
    def __getitem__(self, i): return i


@pytest.mark.parametrize("n", [50_000, 10_000, 1003, 7])
@pytest.mark.parametrize("world", [1, 2, 4, 8])
@pytest.mark.parametrize("shuffle", [True, False])
@pytest.mark.parametrize("drop_last", [False, True])
def test_matches_distributed_sampler(n, world, shuffle, drop_last):
    ds = _DS(n)
    for rank in {0, world - 1, world // 2}:
        ref = DistributedSampler(ds, num_replicas=world, rank=rank, shuffle=shuffle, drop_last=drop_last)
        ours = PartitionSampler(n, world, rank, shuffle=shuffle, drop_last=drop_last)
        assert len(ours) == len(ref)
        for epoch in (0, 1, 5):
            ref.set_epoch(epoch); ours.set_epoch(epoch)
            assert list(ours) == list(ref)


def test_batches_ragged_tail_like_reference():
    # 50,000 samples: W=1 -> 391 steps last 80; W=2 -> 196 / 40; W=8 -> 49 / 106
    for world, steps, last in ((1, 391, 80), (2, 196, 40), (8, 49, 106)):
        s = PartitionSampler(50_000, world, 0, shuffle=True)
        b = s.batches(128)
        assert len(b) == steps == s.num_batches(128) and b[-1].numel() == last
        assert len(s.batches(128, drop_last=True)) == s.num_batches(128, True) == steps - 1


def test_partition_is_disjoint_cover():
    world, n = 4, 1000
    seen = torch.cat([PartitionSampler(n, world, r, shuffle=True).indices() for r in range(world)])
    assert sorted(seen.tolist()) == list(range(n))
