"""DistributedSampler — shards a dataset over ranks.

Behavioural spec: torch:utils/data/distributed.py:66-157 as used by
ref:restnet_ddp.py:108,118,137 (train *and* val sharded, ``set_epoch`` per epoch).
Index arithmetic is re-derived here so it can also run vectorised (the reference
materialises a Python list of 1.28 M ints per epoch; we keep a tensor).
"""
from __future__ import annotations

import math
from typing import Iterator, Optional

import torch


class DistributedSampler:
    def __init__(self, dataset, num_replicas: Optional[int] = None, rank: Optional[int] = None,
                 shuffle: bool = True, seed: int = 0, drop_last: bool = False):
        if num_replicas is None or rank is None:
            import torch.distributed as dist
            if not (dist.is_available() and dist.is_initialized()):
                raise RuntimeError("DistributedSampler needs an initialised process group "
                                   "or explicit num_replicas/rank")
            num_replicas = dist.get_world_size() if num_replicas is None else num_replicas
            rank = dist.get_rank() if rank is None else rank
        if not 0 <= rank < num_replicas:
            raise ValueError(f"invalid rank {rank} for {num_replicas} replicas")
        self.dataset = dataset
        self.num_replicas = num_replicas
        self.rank = rank
        self.epoch = 0
        self.shuffle = shuffle
        self.seed = seed
        self.drop_last = drop_last
        n = len(dataset)
        if drop_last and n % num_replicas != 0:
            self.num_samples = math.ceil((n - num_replicas) / num_replicas)
        else:
            self.num_samples = math.ceil(n / num_replicas)
        self.total_size = self.num_samples * num_replicas

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
        self.epoch = epoch
