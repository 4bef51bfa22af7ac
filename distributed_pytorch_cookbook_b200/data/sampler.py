"""Rank-strided sampler.

Index-for-index equal to ``torch.utils.data.DistributedSampler`` (the reference uses it
with defaults, main-ddp.py:83-84: pad to a multiple of world by wrapping, rank-strided
slice, seed 0 + epoch), plus the ``set_epoch`` call the reference forgets is
made by our trainer so every epoch reshuffles.
"""
from __future__ import annotations

import math
from typing import Iterator

import torch
from torch.utils.data import Sampler


class ShardedSampler(Sampler[int]):
    def __init__(self, dataset_len: int, num_replicas: int, rank: int, shuffle: bool = True,
                 seed: int = 0, drop_last: bool = False):
        if hasattr(dataset_len, "__len__"):
            dataset_len = len(dataset_len)
        self.n = int(dataset_len)
        self.num_replicas, self.rank = num_replicas, rank
        self.shuffle, self.seed, self.drop_last = shuffle, seed, drop_last
        self.epoch = 0
        if drop_last and self.n % num_replicas != 0:
            self.num_samples = math.ceil((self.n - num_replicas) / num_replicas)
        else:
            self.num_samples = math.ceil(self.n / num_replicas)
        self.total_size = self.num_samples * num_replicas

    def set_epoch(self, epoch: int):
        self.epoch = epoch

    def __len__(self) -> int:
        return self.num_samples

    def __iter__(self) -> Iterator[int]:
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
        return iter(idx[self.rank: self.total_size: self.num_replicas])
