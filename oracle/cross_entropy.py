"""Oracle (test infrastructure): mean cross-entropy, its logits gradient, and top-1/top-5 precision.

``nn.CrossEntropyLoss()`` with defaults (imagenet.py:323,124; ``torch/nn/functional.py:3421-3504``:
mean reduction, no label smoothing, no class weights) and ``accuracy(..., topk=(1,5))``
(imagenet.py:63-79,134).  Accumulation in float64; results returned as float32.
"""
import numpy as np

from .step_helpers import rank_of_target


def ce_topk_oracle(logits, target, grad_scale=1.0):
    """Returns (loss_mean, top1_percent, top5_percent, dlogits[N,C], per_row_lse[N])."""
    x = np.asarray(logits, dtype=np.float64)
    t = np.asarray(target, dtype=np.int64)
    n, c = x.shape
    mx = x.max(axis=1, keepdims=True)
    ex = np.exp(x - mx)
    se = ex.sum(axis=1, keepdims=True)
    lse = (mx + np.log(se))[:, 0]
    loss_rows = lse - x[np.arange(n), t]
    soft = ex / se
    soft[np.arange(n), t] -= 1.0
    dlogits = soft * (grad_scale / n)
    r = rank_of_target(np.asarray(logits, dtype=np.float32), t)
    f = np.float32
    top1 = f(f((r < 1).sum()) * f(100.0 / n))
    top5 = f(f((r < 5).sum()) * f(100.0 / n))
    return f(loss_rows.mean()), top1, top5, dlogits.astype(f), lse.astype(f)
