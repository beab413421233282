"""Rank-sharded index sampler.

Same observable behaviour as the ``DistributedSampler(train_dataset, num_replicas=world_size,
rank=rank)`` objects built at ref:main.py:109,115: every rank draws ``ceil(len/replicas)`` indices
(or ``floor`` with ``drop_last``), the epoch's permutation comes from a generator seeded with
``seed + epoch``, the index list is padded by wrapping round to its head so it divides evenly, and
rank ``r`` keeps positions ``r, r+R, r+2R, ...``.  Indices are bit-identical to the library sampler
(tests/test_sampler.py checks a grid of configurations).
"""
from __future__ import annotations

from typing import Iterator, Optional

import torch


class DistributedSampler:
    def __init__(self, dataset, num_replicas: Optional[int] = None, rank: Optional[int] = None, shuffle: bool = True,
                 seed: int = 0, drop_last: bool = False) -> None:
        if num_replicas is None or rank is None:
            from . import distributed as dist
            if not dist.is_initialized():
                raise RuntimeError("DistributedSampler needs num_replicas and rank, or an initialised process group")
            num_replicas = dist.get_world_size() if num_replicas is None else num_replicas
            rank = dist.get_rank() if rank is None else rank
        if not 0 <= rank < num_replicas:
            raise ValueError(f"Invalid rank {rank}, rank should be in the interval [0, {num_replicas - 1}]")
        self.dataset, self.num_replicas, self.rank = dataset, int(num_replicas), int(rank)
        self.shuffle, self.seed, self.drop_last, self.epoch = shuffle, seed, drop_last, 0
        n = len(dataset)
        if drop_last and n % self.num_replicas:
            self.num_samples = n // self.num_replicas
        else:
            self.num_samples = -(-n // self.num_replicas)
        self.total_size = self.num_samples * self.num_replicas

    def _global_order(self) -> list:
        n = len(self.dataset)
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            order = torch.randperm(n, generator=g).tolist()
        else:
            order = list(range(n))
        if self.drop_last:
            return order[: self.total_size]
        short = self.total_size - len(order)
        if short > 0 and order:
            reps = -(-short // len(order))
            order = order + (order * reps)[:short]
        return order

    def __iter__(self) -> Iterator[int]:
        order = self._global_order()
        return iter(order[self.rank: self.total_size: self.num_replicas])

    def __len__(self) -> int:
        return self.num_samples

    def set_epoch(self, epoch: int) -> None:
        """Change the permutation for the next ``__iter__``.  (The reference never calls this, so it
        replays the same order every epoch - ref:main.py:127-131; we keep that default.)"""
        self.epoch = epoch
