"""Fused flat-buffer SGD.

Replaces ``SGD(model.parameters(), lr, momentum=0.9, weight_decay=1e-4)`` (imagenet.py:325) and its
``step()`` (imagenet.py:131): one CUDA kernel over the whole parameter arena applies weight decay,
momentum and the update (and refreshes the bf16 shadow weights) instead of four multi-tensor passes
(torch/optim/sgd.py:425-473).  ``param_groups[i]['lr']`` stays writable (imagenet.py:159-160).
"""
import torch
from torch.optim.optimizer import Optimizer

from . import _cabi, _nvtx


class FlatSGD(Optimizer):
    def __init__(self, params, lr, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False):
        if dampening != 0.0 or nesterov:
            raise ValueError("FlatSGD implements the reference's configuration: dampening=0, nesterov=False")
        if lr < 0 or momentum < 0 or weight_decay < 0:
            raise ValueError("lr, momentum and weight_decay must be non-negative")
        params = list(params)
        super().__init__(params, dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay,
                                      nesterov=nesterov))
        if len(self.param_groups) != 1:
            raise ValueError("FlatSGD applies one hyper-parameter set to every parameter, as the reference does")
        arenas = {id(getattr(p, "_b200_arena", None)): getattr(p, "_b200_arena", None) for p in params}
        if len(arenas) != 1 or None in arenas.values():
            raise TypeError("FlatSGD needs the parameters of one B200ResNet (flat arena); there is no generic path")
        self.arena = next(iter(arenas.values()))
        if len(params) != len(self.arena.shapes):
            raise ValueError("FlatSGD must own every parameter of the arena")

    def zero_grad(self, set_to_none=True):
        """Nothing to clear: every backward overwrites the whole gradient arena (the reference clears the grads
        before each backward, imagenet.py:127, and never accumulates).  It re-arms the guard that makes a second
        backward without ``zero_grad()``/``step()`` raise instead of silently replacing the first one's gradients."""
        self.arena.grads_pending = False
        return None

    @torch.no_grad()
    def step(self, closure=None):
        if closure is not None:
            raise ValueError("closures are not supported")
        g = self.param_groups[0]
        a = self.arena
        a.grads_pending = False
        if a.applied_step == a.steps:   # FlatDDP(exchange="peer") updated every shard inside the backward
            a.steps += 1
            return None
        with _nvtx.range("b200.step"):
            _cabi.sgd_flat(a.params, a.grads, a.momentum, a.shadow, float(g["lr"]), float(g["weight_decay"]),
                           float(g["momentum"]), 1.0, a.steps == 0)
        a.steps += 1
        return None


SGD = FlatSGD
