"""Fused cross-entropy + top-1/top-5.

``CrossEntropyLoss`` replaces ``nn.CrossEntropyLoss().cuda()`` (imagenet.py:323,124) and ``accuracy``
replaces the reference's helper (imagenet.py:63-79): one kernel yields the mean loss and both
precisions; ``accuracy`` called on the same logits/target right after the criterion (as train() does,
imagenet.py:124,134) reuses that result instead of launching again.
"""
import torch
import torch.nn as nn

from . import _cabi, _nvtx

# most recent (logits, target) pair and its result; holding the tensors keeps their memory from
# being handed to another tensor, so a pointer match means "the same batch" - as long as nobody wrote into
# them in place since, which their version counters tell (detach() shares the counter of the original)
_last = {"key": None, "out3": None, "hold": None, "versions": None}


def _key(logits, target):
    return (logits.data_ptr(), tuple(logits.shape), logits.dtype, target.data_ptr(), target.numel())


def _remember(logits, target, out3):
    hold = (logits.detach(), target.detach())
    _last["key"], _last["out3"], _last["hold"] = _key(logits, target), out3, hold
    _last["versions"] = (hold[0]._version, hold[1]._version)


def _cached(logits, target):
    """The remembered result if (logits, target) is the remembered pair and neither was modified in place."""
    if _last["key"] != _key(logits, target):
        return None
    hold = _last["hold"]
    if (hold[0]._version, hold[1]._version) != _last["versions"]:
        return None
    return _last["out3"]


class _CEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, target):
        out3, row_lse = _cabi.ce_topk_fwd(logits, target)
        ctx.save_for_backward(logits, target, row_lse)
        _remember(logits, target, out3)
        return out3[0]  # 0-dim view: loss; out3 keeps the precisions

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g):
        logits, target, row_lse = ctx.saved_tensors
        g = g.to(torch.float32).contiguous()
        return _cabi.ce_bwd(logits, target, row_lse, g, 1.0), None


class CrossEntropyLoss(nn.Module):
    """Mean cross-entropy over the batch (defaults of ``nn.CrossEntropyLoss``)."""

    def forward(self, logits, target):
        if logits.dim() != 2:
            raise ValueError("expected [N, C] logits")
        logits = logits if logits.is_contiguous() else logits.contiguous()
        with _nvtx.range("b200.loss"):
            if torch.is_grad_enabled() and logits.requires_grad:
                return _CEFn.apply(logits, target)
            out3, _ = _cabi.ce_topk_fwd(logits, target)
            _remember(logits, target, out3)
            return out3[0]

    @staticmethod
    def last_metrics():
        """float32[3] device tensor {loss, top-1 %, top-5 %} of the most recent call."""
        return _last["out3"]


def accuracy(output, target, topk=(1,)):
    """precision@k in percent as 1-element float32 tensors, for any tuple of k (imagenet.py:63-79).

    k = 1 and k = 5 (what the reference logs, imagenet.py:134) come out of the fused loss kernel; any other k is
    counted from the rank of the target class (``b200_target_rank``): ``100/N * #{n : rank_n < k}``, which is what
    ``topk(maxk) -> eq -> sum`` gives on tie-free logits."""
    topk = tuple(int(k) for k in topk)
    if any(k < 1 for k in topk):
        raise ValueError("k must be >= 1")
    with torch.no_grad():
        output = output if output.is_contiguous() else output.contiguous()
        if max(topk) > output.size(1):
            raise RuntimeError("selected index k out of range")      # what output.topk(maxk) raises in the reference
        if all(k in (1, 5) for k in topk):
            out3 = _cached(output, target)
            if out3 is None:
                out3, _ = _cabi.ce_topk_fwd(output, target)
            return [out3[1:2] if k == 1 else out3[2:3] for k in topk]
        rank = _cabi.target_rank(output, target)
        pct = 100.0 / output.size(0)
        return [(rank < k).sum().reshape(1).float().mul_(pct) for k in topk]
