"""S1 — DistributedSampler with set_epoch, bit-identical to torch.utils.data.DistributedSampler.

Mirrors the call sites reference multi-GPU-training-torch.py:80-83 (construction with
shuffle=True, num_replicas, rank) and :175-177 (set_epoch before each epoch's iterator), and the
index arithmetic of torch/utils/data/distributed.py:107-141: permutation from a CPU
torch.Generator seeded with seed+epoch, pad by wrap-around to a multiple of the world size (or
drop the tail), then take every world-size-th index starting at rank. The permutation itself is
torch.randperm (mt19937 + torch's shuffle) so the stream of indices is the reference's, bit for bit.
"""
from __future__ import annotations

import math
from typing import Iterator, Optional

import torch
import torch.distributed as dist
from torch.utils.data import Sampler


class DistributedSampler(Sampler):
    def __init__(self, dataset, num_replicas: Optional[int] = None, rank: Optional[int] = None,
                 shuffle: bool = True, seed: int = 0, drop_last: bool = False) -> None:
        if num_replicas is None:
            if not dist.is_available() or not dist.is_initialized():
                raise RuntimeError("Requires distributed package to be available and initialised")
            num_replicas = dist.get_world_size()
        if rank is None:
            if not dist.is_available() or not dist.is_initialized():
                raise RuntimeError("Requires distributed package to be available and initialised")
            rank = dist.get_rank()
        if rank >= num_replicas or rank < 0:
            raise ValueError(f"Invalid rank {rank}, rank should be in the interval [0, {num_replicas - 1}]")
        self.dataset = dataset
        self.num_replicas = int(num_replicas)
        self.rank = int(rank)
        self.epoch = 0
        self.drop_last = bool(drop_last)
        n = len(self.dataset)
        if self.drop_last and n % self.num_replicas != 0:
            # the tail that does not divide evenly is dropped
            self.num_samples = math.ceil((n - self.num_replicas) / self.num_replicas)
        else:
            self.num_samples = math.ceil(n / self.num_replicas)
        self.total_size = self.num_samples * self.num_replicas
        self.shuffle = bool(shuffle)
        self.seed = int(seed)

    def _global_order(self) -> list:
        n = len(self.dataset)
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            order = torch.randperm(n, generator=g).tolist()
        else:
            order = list(range(n))
        if not self.drop_last:
            pad = self.total_size - len(order)
            if pad > 0:
                if pad <= len(order):
                    order += order[:pad]
                else:
                    order += (order * math.ceil(pad / len(order)))[:pad]
        else:
            order = order[: self.total_size]
        assert len(order) == self.total_size
        return order

    def __iter__(self) -> Iterator[int]:
        shard = self._global_order()[self.rank: self.total_size: self.num_replicas]
        assert len(shard) == self.num_samples
        return iter(shard)

    def __len__(self) -> int:
        return self.num_samples

    def set_epoch(self, epoch: int) -> None:
        """Make the next iterator use the permutation of `epoch` (README.md:82-85 of the reference)."""
        self.epoch = int(epoch)
