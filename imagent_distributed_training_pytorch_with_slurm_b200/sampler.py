"""Index stream of one rank.

Replaces ``DistributedSampler(trainset)`` / ``set_epoch`` (imagenet.py:346-347,375; algorithm
torch/utils/data/distributed.py:107-141, defaults shuffle=True, seed=0, drop_last=False): a
permutation from the torch CPU generator seeded ``seed + epoch``, padded by wrap-around to a multiple
of the world size, then strided ``rank::world``.  Index order is bit-exact with torch's sampler
(tests/golden/golden.json).
"""
import math

import torch
import torch.distributed as dist


class DistributedSampler:
    def __init__(self, dataset, num_replicas=None, rank=None, shuffle=True, seed=0):
        if num_replicas is None:
            num_replicas = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
        if rank is None:
            rank = dist.get_rank() if dist.is_available() and dist.is_initialized() else 0
        if not 0 <= rank < num_replicas:
            raise ValueError(f"invalid rank {rank} for {num_replicas} replicas")
        self.dataset_len = dataset if isinstance(dataset, int) else len(dataset)
        self.num_replicas, self.rank, self.shuffle, self.seed, self.epoch = num_replicas, rank, shuffle, seed, 0
        self.num_samples = math.ceil(self.dataset_len / num_replicas)
        self.total_size = self.num_samples * num_replicas

    def set_epoch(self, epoch):
        self.epoch = epoch

    def indices(self):
        """int64 tensor of this rank's dataset indices for the current epoch."""
        n = self.dataset_len
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            order = torch.randperm(n, generator=g)
        else:
            order = torch.arange(n)
        pad = self.total_size - n
        if pad > 0:
            order = torch.cat([order] + [order] * math.ceil(pad / n))[: self.total_size]
        return order[self.rank: self.total_size: self.num_replicas]

    def __iter__(self):
        return iter(self.indices().tolist())

    def __len__(self):
        return self.num_samples
