"""DistributedSampler-equivalent shard sampler (D8 in SURVEY.md).

Reference use: ``DistributedSampler(train_dataset, num_replicas=world_size, rank=rank)`` at
mnist_distributed.py:73-75 with the defaults ``shuffle=True, seed=0, drop_last=False`` and
``set_epoch`` never called (so every epoch replays the same permutation).  The index stream
is bit-identical to torch's: a CPU ``torch.randperm`` seeded with ``seed + epoch``, wrapped
padding up to ``num_replicas * ceil(len / num_replicas)``, then a rank-strided slice.
"""
from __future__ import annotations

import math
from typing import Iterator, Optional, Sized

import torch


class DistributedSampler:
    def __init__(self, dataset: Sized, num_replicas: Optional[int] = None, rank: Optional[int] = None,
                 shuffle: bool = True, seed: int = 0, drop_last: bool = False):
        if num_replicas is None or rank is None:
            from .. import distributed as dist
            if not dist.is_initialized():
                raise RuntimeError("DistributedSampler needs num_replicas/rank or an initialised process group")
            num_replicas = dist.get_world_size() if num_replicas is None else num_replicas
            rank = dist.get_rank() if rank is None else rank
        if not (0 <= rank < num_replicas):
            raise ValueError(f"invalid rank {rank}: it must lie in [0, {num_replicas - 1}]")
        self.dataset, self.num_replicas, self.rank = dataset, int(num_replicas), int(rank)
        self.shuffle, self.seed, self.drop_last, self.epoch = shuffle, int(seed), drop_last, 0
        n = len(dataset)
        if drop_last and n % self.num_replicas != 0:
            self.num_samples = math.ceil((n - self.num_replicas) / self.num_replicas)
        else:
            self.num_samples = math.ceil(n / self.num_replicas)
        self.total_size = self.num_samples * self.num_replicas

    def indices(self) -> torch.Tensor:
        n = len(self.dataset)
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            idx = torch.randperm(n, generator=g)
        else:
            idx = torch.arange(n)
        if not self.drop_last:
            pad = self.total_size - n
            if pad > 0:
                reps = math.ceil(pad / max(n, 1))
                idx = torch.cat([idx, idx.repeat(reps)[:pad]])
        else:
            idx = idx[: self.total_size]
        return idx[self.rank: self.total_size: self.num_replicas]

    def __iter__(self) -> Iterator[int]:
        return iter(self.indices().tolist())

    def __len__(self) -> int:
        return self.num_samples

    def set_epoch(self, epoch: int) -> None:
        self.epoch = int(epoch)
