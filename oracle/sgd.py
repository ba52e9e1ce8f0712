"""Oracle (test infrastructure): one SGD-with-momentum update on a flat float32 array.

Follows ``torch.optim.SGD`` as the reference builds it (imagenet.py:325: lr, momentum 0.9,
weight_decay 1e-4, dampening 0, nesterov False, applied to every parameter) in the single-tensor
form ``torch/optim/sgd.py:343-380``; each numpy operation rounds to float32 like the separate torch
ops do.  ``grad_scale`` is the DDP 1/world_size pre-divide (``reducer.hpp:497-500``) when the caller
holds un-averaged sums; the product path pre-divides at the loss instead and passes 1.
"""
import numpy as np


def sgd_step_oracle(p, g, buf, lr, momentum=0.9, weight_decay=1e-4, first_step=False, grad_scale=1.0):
    p = np.asarray(p, dtype=np.float32)
    g = np.asarray(g, dtype=np.float32)
    f = np.float32
    if grad_scale != 1.0:
        g = g * f(grad_scale)
    if weight_decay != 0:
        g = g + f(weight_decay) * p
    if momentum != 0:
        if first_step or buf is None:
            buf = g.copy()
        else:
            buf = np.asarray(buf, dtype=np.float32) * f(momentum) + g
        g = buf
    p = p + f(-lr) * g
    return p.astype(np.float32), (None if buf is None else buf.astype(np.float32))
