"""Oracle (test infrastructure) for the host-side helpers of the training step.

numpy restatements; nothing here is imported by the product package.
"""
import numpy as np


class MeterOracle:
    """Running mean, as ``AverageMeter`` (imagenet.py:44-60).

    ``update(val, n)`` adds ``val*n`` to ``sum`` and ``n`` to ``count``.  The reference
    always passes ``n = input[0].size(0)`` (imagenet.py:142), which is the channel count 3,
    so epoch averages are unweighted means over iterations.
    """

    def __init__(self):
        self.val = 0
        self.avg = 0
        self.sum = 0
        self.count = 0

    def update(self, val, n=1):
        self.val = val
        self.sum += val * n
        self.count += n
        self.avg = self.sum / self.count


def rank_of_target(logits, target):
    """Position the target class would take in a descending sort of each row.

    Tie rule: equal logits are ordered by ascending class index (an equal logit with a lower
    index counts as ranked ahead of the target).
    """
    logits = np.asarray(logits, dtype=np.float32)
    target = np.asarray(target, dtype=np.int64)
    n, c = logits.shape
    xt = logits[np.arange(n), target][:, None]
    idx = np.arange(c)[None, :]
    ahead = (logits > xt) | ((logits == xt) & (idx < target[:, None]))
    return ahead.sum(axis=1)


def accuracy_oracle(logits, target, topk=(1,)):
    """precision@k in percent, as ``accuracy`` (imagenet.py:63-79): the share of rows whose
    target class is among the k largest logits, times 100.  Returns float32 values (the
    reference returns 1-element float32 tensors: count.float() * (100.0 / batch))."""
    r = rank_of_target(logits, target)
    n = len(r)
    out = []
    for k in topk:
        cnt = np.float32((r < k).sum())
        out.append(np.float32(cnt * np.float32(100.0 / n)))
    return out


def reduce_mean_oracle(per_rank_values):
    """``reduce_tensor`` (imagenet.py:82-87): all-reduce SUM then divide by world size, in
    the dtype of the tensor (float32)."""
    vals = [np.float32(v) for v in per_rank_values]
    s = np.float32(0.0)
    for v in vals:
        s = np.float32(s + v)
    return np.float32(s / np.float32(len(vals)))


def lr_schedule_oracle(initial_lr, epoch):
    """``adjust_learning_rate`` (imagenet.py:154-162): python-float
    ``initial_lr * 0.1 ** (epoch // 30)``; the float repr is pinned by
    ``imagent_sgd.out:274,454,634,814``."""
    return initial_lr * (0.1 ** (epoch // 30))
