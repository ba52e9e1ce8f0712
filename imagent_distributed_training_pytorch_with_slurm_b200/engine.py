"""The training / validation step loops, with the reference's names and signatures.

``train`` mirrors imagenet.py:97-151, ``validate`` :166-210, ``AverageMeter`` :44-60,
``reduce_tensor`` :82-87, ``to_python_float`` :90-94, ``adjust_learning_rate`` :154-162.

What changes underneath (the observable results do not):
  * the reference forces four host<->device syncs per iteration (three ``.item()``, imagenet.py:143-145,
    and ``torch.cuda.synchronize()``, :147).  Here the per-iteration {loss, top-1, top-5} triples stay
    on the device; every ``sync_every`` iterations (default: once per epoch) they are stacked,
    all-reduced in ONE collective (instead of 3 per iteration, imagenet.py:137-139), divided by the
    world size in float32, copied to the host once, and fed to the same meters with the same weight
    ``n = input[0].size(0)`` (= 3, the channel count, imagenet.py:142) in iteration order - so the
    epoch averages are bit-identical to per-iteration logging;
  * ``batch_time.sum`` (the "Train time" the reference prints, imagenet.py:403) is a telescoping sum
    of wall-clock deltas, so it only needs the final synchronisation to be the same number.
``sync_every=1`` restores the reference's per-iteration synchronisation exactly.
"""
import time

import torch
import torch.distributed as dist

from .loss import CrossEntropyLoss, accuracy  # noqa: F401  (accuracy re-exported under the reference's name)


class AverageMeter(object):
    """Computes and stores the average and current value (imagenet.py:44-60)."""

    def __init__(self):
        self.reset()

    def reset(self):
        self.val = 0
        self.avg = 0
        self.sum = 0
        self.count = 0

    def update(self, val, n=1):
        self.val = val
        self.sum += val * n
        self.count += n
        self.avg = self.sum / self.count


def reduce_tensor(tensor, world_size):
    """Mean of ``tensor`` over all ranks (imagenet.py:82-87); any shape, one collective."""
    rt = tensor.clone()
    if world_size > 1:
        dist.all_reduce(rt, op=dist.ReduceOp.SUM)
    rt /= world_size
    return rt


def to_python_float(t):
    return t.item() if hasattr(t, "item") else t[0]


def adjust_learning_rate(initial_lr, optimizer, epoch):
    """LR = initial LR decayed by 10 every 30 epochs (imagenet.py:154-162)."""
    lr = initial_lr * (0.1 ** (epoch // 30))
    for group in optimizer.param_groups:
        group["lr"] = lr
    return lr


def _triple(criterion, loss, logits, target):
    if isinstance(criterion, CrossEntropyLoss):
        return criterion.last_metrics()  # float32[3] straight from the fused kernel
    p1, p5 = accuracy(logits.data, target.data, topk=(1, 5))
    return torch.stack([loss.data.float().reshape(()), p1[0], p5[0]])


def _drain(pending, world_size, weight, meters):
    if not pending:
        return
    m = reduce_tensor(torch.stack(pending), world_size)
    rows = m.cpu().tolist()  # the host sync
    for lo, p1, p5 in rows:
        meters[0].update(lo, weight)
        meters[1].update(p1, weight)
        meters[2].update(p5, weight)
    pending.clear()


def _to_device(t):
    return t if t.is_cuda else t.cuda(non_blocking=True)


def train(train_loader, model, criterion, optimizer, epoch, world_size, sync_every=None):
    batch_time, data_time = AverageMeter(), AverageMeter()
    losses, top1, top5 = AverageMeter(), AverageMeter(), AverageMeter()
    model.train()
    pending, weight, iters = [], 1, 0
    start = end = time.time()
    for i, (input, target) in enumerate(train_loader):
        data_time.update(time.time() - end)
        input, target = _to_device(input), _to_device(target)

        logits = model(input)
        loss = criterion(logits, target)

        optimizer.zero_grad()
        loss.backward()
        optimizer.step()

        pending.append(_triple(criterion, loss, logits, target))
        weight = input[0].size(0)
        iters += 1
        if sync_every and iters % sync_every == 0:
            _drain(pending, world_size, weight, (losses, top1, top5))
            torch.cuda.synchronize()
            batch_time.update(time.time() - end)
        end = time.time()
    _drain(pending, world_size, weight, (losses, top1, top5))
    torch.cuda.synchronize()
    total = time.time() - start
    if batch_time.count < iters:  # spread the un-attributed wall time evenly; the sum is what is reported
        rest = iters - batch_time.count
        remaining = total - batch_time.sum
        for _ in range(rest):
            batch_time.update(remaining / rest)
    return losses.avg, top1.avg, top5.avg, batch_time.sum


def validate(val_loader, model, criterion, world_size, sync_every=None):
    batch_time = AverageMeter()
    losses, top1, top5 = AverageMeter(), AverageMeter(), AverageMeter()
    model.eval()
    pending, weight, iters = [], 1, 0
    with torch.no_grad():
        start = end = time.time()
        for i, (input, target) in enumerate(val_loader):
            input, target = _to_device(input), _to_device(target)
            logits = model(input)
            loss = criterion(logits, target)
            pending.append(_triple(criterion, loss, logits, target))
            weight = input[0].size(0)
            iters += 1
            if sync_every and iters % sync_every == 0:
                _drain(pending, world_size, weight, (losses, top1, top5))
                torch.cuda.synchronize()
                batch_time.update(time.time() - end)
            end = time.time()
        _drain(pending, world_size, weight, (losses, top1, top5))
        torch.cuda.synchronize()
        total = time.time() - start
    if batch_time.count < iters:
        rest = iters - batch_time.count
        remaining = total - batch_time.sum
        for _ in range(rest):
            batch_time.update(remaining / rest)
    return losses.avg, top1.avg, top5.avg, batch_time.sum
