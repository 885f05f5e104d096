"""ShardedSampler reproduces torch DistributedSampler's index math (torch/utils/data/distributed.py:107-141)."""
import pytest
import torch
from torch.utils.data.distributed import DistributedSampler

from picd_b200.parallel import ShardedSampler


class _DS(torch.utils.data.Dataset):
    def __init__(self, n):
        self.n = n

    def __len__(self):
        return self.n

    def __getitem__(self, i):
        return i


@pytest.mark.parametrize("n,world", [(10, 2), (11, 4), (7, 8), (100, 3), (1, 2)])
@pytest.mark.parametrize("shuffle", [True, False])
@pytest.mark.parametrize("drop_last", [False, True])
def test_matches_torch_distributed_sampler(n, world, shuffle, drop_last):
    if drop_last and n < world:
        pytest.skip("torch yields nothing useful here")
    for epoch in (0, 3):
        for rank in range(world):
            ours = ShardedSampler(n, world, rank, shuffle=shuffle, seed=0, drop_last=drop_last)
            theirs = DistributedSampler(_DS(n), world, rank, shuffle=shuffle, seed=0, drop_last=drop_last)
            ours.set_epoch(epoch)
            theirs.set_epoch(epoch)
            assert list(ours) == list(theirs)
            assert len(ours) == len(theirs)


def test_partition_and_padding():
    world, n = 4, 10
    shards = [list(ShardedSampler(n, world, r, shuffle=True, seed=5)) for r in range(world)]
    assert all(len(s) == 3 for s in shards)                     # padded to equal size
    flat = sorted(i for s in shards for i in s)
    assert set(flat) == set(range(n)) and len(flat) == 12        # 2 wrap-around duplicates


def test_set_epoch_reshuffles():
    s = ShardedSampler(50, 2, 0)
    s.set_epoch(0)
    a = list(s)
    s.set_epoch(1)
    assert a != list(s)
    s.set_epoch(0)
    assert a == list(s)


def test_bad_rank_rejected():
    with pytest.raises(ValueError):
        ShardedSampler(10, 2, 2)
