"""Index sharding with the exact semantics the reference gets from torch.

* distributed parts: ``DistributedSampler(dataset, num_replicas=W, rank=r)`` with its defaults
  (shuffle, seed 0, ``drop_last=False`` -> pad by wrapping to a multiple of W, then take every
  W-th index starting at ``rank``); ``set_epoch`` is never called by the reference
  (ref ``Part 2a/main.py:38``).
* Part 1: ``shuffle=True`` -> a ``RandomSampler`` that draws its seed from the global torch RNG
  at the start of every epoch (ref ``Part 1/main.py:96-101``).

Implemented directly on ``torch.randperm`` so the index streams are bit-identical to torch's
samplers without instantiating them (tests compare against the real ones).
"""
from __future__ import annotations

import math
from typing import Iterator, List, Optional

import torch


class ShardedSampler:
    def __init__(self, dataset_len: int, num_replicas: int = 1, rank: int = 0,
                 shuffle: bool = True, seed: int = 0, drop_last: bool = False) -> None:
        if not 0 <= rank < num_replicas:
            raise ValueError(f"rank {rank} outside [0, {num_replicas})")
        self.n = int(dataset_len)
        self.world = int(num_replicas)
        self.rank = int(rank)
        self.shuffle = shuffle
        self.seed = seed
        self.drop_last = drop_last
        self.epoch = 0
        if drop_last and self.n % self.world:
            self.num_samples = math.ceil((self.n - self.world) / self.world)
        else:
            self.num_samples = math.ceil(self.n / self.world)
        self.total_size = self.num_samples * self.world

    def set_epoch(self, epoch: int) -> None:
        self.epoch = int(epoch)

    def indices(self) -> List[int]:
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            order = torch.randperm(self.n, generator=g).tolist()
        else:
            order = list(range(self.n))
        if not self.drop_last:
            pad = self.total_size - len(order)
            if pad > 0:
                reps = math.ceil(pad / max(len(order), 1))
                order += (order * reps)[:pad]
        else:
            order = order[: self.total_size]
        return order[self.rank: self.total_size: self.world]

    def __iter__(self) -> Iterator[int]:
        return iter(self.indices())

    def __len__(self) -> int:
        return self.num_samples


class GlobalShuffleSampler:
    """Part 1's ``shuffle=True``: a fresh permutation per epoch seeded from the global RNG."""

    def __init__(self, dataset_len: int, generator: Optional[torch.Generator] = None) -> None:
        self.n = int(dataset_len)
        self.generator = generator

    def indices(self) -> List[int]:
        g = self.generator
        if g is None:
            seed = int(torch.empty((), dtype=torch.int64).random_().item())
            g = torch.Generator()
            g.manual_seed(seed)
        return torch.randperm(self.n, generator=g).tolist()

    def __iter__(self) -> Iterator[int]:
        return iter(self.indices())

    def __len__(self) -> int:
        return self.n


class SequentialSampler:
    def __init__(self, dataset_len: int) -> None:
        self.n = int(dataset_len)

    def indices(self) -> List[int]:
        return list(range(self.n))

    def __iter__(self) -> Iterator[int]:
        return iter(range(self.n))

    def __len__(self) -> int:
        return self.n


def batches_per_epoch(num_samples: int, batch_size: int, drop_last: bool = False) -> int:
    """``len(DataLoader)``: the reference keeps the short last batch."""
    if drop_last:
        return num_samples // batch_size
    return math.ceil(num_samples / batch_size)
