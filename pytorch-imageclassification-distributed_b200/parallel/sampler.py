"""Rank-sharded index sampler with ``torch.utils.data.DistributedSampler`` index math.

The reference shards with torch's DistributedSampler (train.py:113,117; defaults
shuffle=True, seed=0, drop_last=False).  ``ShardedSampler`` reproduces its indices
exactly -- randperm from ``seed + epoch``, pad by wrapping to a multiple of the world
size, then take ``rank::world`` (torch/utils/data/distributed.py:107-141) -- without
needing an initialised process group, so it is testable on one CPU process.
"""
from __future__ import annotations

import math
from typing import Iterator, Optional

import torch
from torch.utils.data import Sampler

from .dist import get_rank, get_world_size


class ShardedSampler(Sampler[int]):
    def __init__(self, dataset_len, num_replicas: Optional[int] = None, rank: Optional[int] = None,
                 shuffle: bool = True, seed: int = 0, drop_last: bool = False):
        n = dataset_len if isinstance(dataset_len, int) else len(dataset_len)
        self.n = n
        self.num_replicas = get_world_size() if num_replicas is None else num_replicas
        self.rank = get_rank() if rank is None else rank
        if not 0 <= self.rank < self.num_replicas:
            raise ValueError(f"rank {self.rank} outside [0, {self.num_replicas})")
        self.shuffle, self.seed, self.drop_last, self.epoch = shuffle, seed, drop_last, 0
        if drop_last and n % self.num_replicas:
            self.num_samples = math.ceil((n - self.num_replicas) / self.num_replicas)
        else:
            self.num_samples = math.ceil(n / self.num_replicas)
        self.total_size = self.num_samples * self.num_replicas

    def set_epoch(self, epoch: int) -> None:
        """Called once per epoch (train.py:164) so every epoch reshuffles identically on all ranks."""
        self.epoch = epoch

    def global_order(self):
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            idx = torch.randperm(self.n, generator=g).tolist()
        else:
            idx = list(range(self.n))
        if not self.drop_last:
            pad = self.total_size - len(idx)
            if pad > 0:
                idx += (idx * math.ceil(pad / max(len(idx), 1)))[:pad]
        else:
            idx = idx[: self.total_size]
        return idx

    def __iter__(self) -> Iterator[int]:
        return iter(self.global_order()[self.rank: self.total_size: self.num_replicas])

    def __len__(self) -> int:
        return self.num_samples
