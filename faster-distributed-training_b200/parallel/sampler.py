"""Rank-sharded epoch sampler.

Same contract as ``torch.utils.data.DistributedSampler`` as the reference uses it
(``resnet50_test.py:330-334``, ``transformer_test.py:107-117``): a per-epoch permutation seeded by
``seed + epoch``, padded to a multiple of the world size, strided by rank.  ``set_epoch`` must be
called each epoch; the trainers here always do (the reference forgets it for ResNet — Q7).
"""
from __future__ import annotations

import math
from typing import Iterator, Optional

import torch
from torch.utils.data import Sampler


class ShardedSampler(Sampler[int]):
    def __init__(self, dataset_len: int, num_replicas: int = 1, rank: int = 0, shuffle: bool = True,
                 seed: int = 0, drop_last: bool = False):
        if not 0 <= rank < num_replicas:
            raise ValueError("rank out of range")
        self.n, self.num_replicas, self.rank = dataset_len, num_replicas, rank
        self.shuffle, self.seed, self.drop_last, self.epoch = shuffle, seed, drop_last, 0
        if drop_last and self.n % num_replicas:
            self.num_samples = self.n // num_replicas
        else:
            self.num_samples = math.ceil(self.n / num_replicas)
        self.total_size = self.num_samples * num_replicas

    def set_epoch(self, epoch: int) -> None:
        self.epoch = epoch

    def indices(self) -> torch.Tensor:
        if self.shuffle:
            g = torch.Generator().manual_seed(self.seed + self.epoch)
            idx = torch.randperm(self.n, generator=g)
        else:
            idx = torch.arange(self.n)
        if self.total_size > idx.numel():
            reps = math.ceil(self.total_size / max(idx.numel(), 1))
            idx = idx.repeat(reps)[: self.total_size]
        else:
            idx = idx[: self.total_size]
        return idx[self.rank:self.total_size:self.num_replicas]

    def __iter__(self) -> Iterator[int]:
        return iter(self.indices().tolist())

    def __len__(self) -> int:
        return self.num_samples
