"""Oracle (test infrastructure): which dataset indices a rank visits in an epoch.

Follows ``DistributedSampler`` as the reference builds it (imagenet.py:346-347: default
``shuffle=True, seed=0, drop_last=False``; ``set_epoch`` at imagenet.py:375), whose algorithm is
``torch/utils/data/distributed.py:107-141``.  The permutation itself comes from torch's CPU
generator (``torch.randperm`` with a generator seeded ``seed + epoch``); that generator is the
pinned third-party piece, the padding and striding below are restated.
"""
import math

import numpy as np
import torch


def sampler_indices_oracle(dataset_len, world_size, rank, epoch=0, seed=0, shuffle=True):
    if shuffle:
        g = torch.Generator()
        g.manual_seed(seed + epoch)
        order = torch.randperm(dataset_len, generator=g).numpy()
    else:
        order = np.arange(dataset_len, dtype=np.int64)
    per_rank = math.ceil(dataset_len / world_size)
    total = per_rank * world_size
    pad = total - dataset_len
    if pad > 0:
        reps = math.ceil(pad / dataset_len)
        order = np.concatenate([order] + [order] * reps)[:total]
    mine = order[rank:total:world_size]
    assert len(mine) == per_rank
    return mine.astype(np.int64)
