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
``sync_every=1`` restores the reference's per-iteration synchronisation exactly; ``sync_every=1, lag=1``
still reads every iteration's result on the host, one iteration late, so the device never waits for
the host.  ``graph=True`` additionally replays the whole iteration (forward, loss, backward, optimizer)
as one CUDA graph after a short eager warm-up (``GraphedStep``).
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


class _MetricLog:
    """Per-iteration {loss, top-1, top-5} triples -> host meters.

    Triples are grouped (``sync_every`` iterations per group, one group per epoch when None); a
    finished group is stacked, all-reduced once, divided by the world size in float32 and copied to
    pinned host memory asynchronously, followed by an event.  The host waits only for the event of the
    group that finished ``lag`` groups earlier, so with ``lag >= 1`` the GPU never idles waiting for
    Python to read a scalar; the meters still receive every value, in order."""

    def __init__(self, world_size, meters, sync_every, lag):
        self.world_size, self.meters = world_size, meters
        self.sync_every, self.lag = sync_every, max(0, int(lag))
        self.pending, self.inflight, self.weight = [], [], 1

    def add(self, triple, weight):
        self.pending.append(triple)
        self.weight = weight
        if self.sync_every and len(self.pending) >= self.sync_every:
            self._launch()
            self._collect(self.lag)
            return True
        return False

    def _launch(self):
        if not self.pending:
            return
        m = reduce_tensor(torch.stack(self.pending), self.world_size)
        self.pending = []
        if m.is_cuda:
            host = torch.empty(m.shape, dtype=m.dtype, pin_memory=True)
            host.copy_(m, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
        else:
            host, ev = m, None
        self.inflight.append((host, ev, self.weight))

    def _collect(self, keep):
        while len(self.inflight) > keep:
            host, ev, weight = self.inflight.pop(0)
            if ev is not None:
                ev.synchronize()  # the host sync
            for lo, p1, p5 in host.tolist():
                self.meters[0].update(lo, weight)
                self.meters[1].update(p1, weight)
                self.meters[2].update(p5, weight)

    def finish(self):
        self._launch()
        self._collect(0)


class GraphedStep:
    """forward + loss + backward + optimizer step of one iteration as a CUDA graph.

    The first ``WARMUP`` iterations run eagerly (cuDNN autotuning, lazy initialisation), the next one is
    captured with ``torch.cuda.graph`` on static input/target buffers, and every later iteration of the
    same shape is a copy into those buffers plus one ``graph.replay()`` - the ~600 kernel launches of a step
    cost the host one call instead of ~10 ms of Python (the reference loop forces four host syncs per
    iteration, imagenet.py:143-147; this is the other end of that scale).  The kernels, their order and the
    side-stream / programmatic-dependent-launch structure are exactly those of the eager step, so the numbers
    are too.  Hyper-parameters are frozen into the graph: it lives for one ``train()`` call (one epoch, one
    learning rate, imagenet.py:378-381) unless the caller builds a ``GraphedStep`` itself and passes it as
    ``train(..., graph=step)`` to reuse it.  An iteration with a different batch shape (the short last batch)
    runs eagerly."""

    WARMUP = 3

    def __init__(self, model, criterion, optimizer):
        self.model, self.criterion, self.optimizer = model, criterion, optimizer
        self.seen, self.graph, self.x, self.t, self.out3 = 0, None, None, None, None
        inner = getattr(model, "module", model)
        self.arena = getattr(inner, "arena", None)
        # Other threads of the process keep working while this one captures: the loader's producer thread allocates
        # pinned host memory (which, in "global" mode, invalidates the capture whenever the two coincide - it did, once
        # the warm-up epoch was long enough for the producer to still be busy at the fourth iteration), NCCL's watchdog
        # polls events.  None of that touches the capturing stream, so capture errors are confined to this thread.
        self.capture_error_mode = "thread_local"

    def _eager(self, x, t):
        logits = self.model(x)
        loss = self.criterion(logits, t)
        self.optimizer.zero_grad()
        loss.backward()
        self.optimizer.step()
        return _triple(self.criterion, loss, logits, t)

    def __call__(self, x, t):
        if self.graph is None and self.seen < self.WARMUP:
            self.seen += 1
            return self._eager(x, t)
        if self.graph is None:
            self.x, self.t = torch.empty_like(x), torch.empty_like(t)
            for attr in ("_b200_s2d_stem",):
                if hasattr(x, attr):
                    setattr(self.x, attr, getattr(x, attr))
            self.x.copy_(x)
            self.t.copy_(t)
            torch.cuda.synchronize()
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph, capture_error_mode=self.capture_error_mode):
                self.out3 = self._eager(self.x, self.t)
            if self.arena is not None:
                self.arena.steps -= 1  # capture ran the Python of one step without executing it
        elif x.shape != self.x.shape or t.shape != self.t.shape:
            return self._eager(x, t)
        else:
            self.x.copy_(x)
            self.t.copy_(t)
        self.graph.replay()
        if self.arena is not None:
            self.arena.steps += 1      # what optimizer.step() does on the host (keys the stem-weight cache)
        return self.out3.clone()


class SegmentedGraphStep:
    """The iteration of a multi-process (FlatDDP) run as a CHAIN of CUDA graphs with the NCCL collectives
    issued eagerly between them.

    Capturing NCCL inside one whole-iteration graph invalidates the capture on this stack, so the capture is
    cut wherever the backward completes a gradient bucket: graph 0 = forward + loss + backward down to the first
    bucket boundary, then `all_reduce(bucket 0)` (eager, asynchronous on NCCL's stream), graph 1 = backward down
    to the next boundary, ... , last graph = the fused SGD step behind the all-reduce waits.  The BatchNorm buffer
    broadcast (rank 0 wins, as DDP does) is issued after graph 0 and joined before the last graph.  The network
    is driven directly (its own forward/backward executor, no autograd), so this runner needs the package's own
    model, criterion and optimizer."""

    WARMUP = 3

    def __init__(self, model, criterion, optimizer):
        from .ddp import FlatDDP
        from .optim import FlatSGD

        if not (isinstance(model, FlatDDP) and isinstance(criterion, CrossEntropyLoss) and isinstance(optimizer, FlatSGD)):
            raise TypeError("SegmentedGraphStep drives FlatDDP(B200ResNet) + CrossEntropyLoss + FlatSGD")
        if getattr(model, "exchange", "nccl") != "nccl":
            # the peer exchange keeps momentum sharded by rank and updates inside the backward; this runner records
            # NCCL all-reduces + the flat SGD kernel over the whole arena - mixing the two corrupts the replicas
            raise ValueError("SegmentedGraphStep is the NCCL path; FlatDDP(exchange='peer') runs as one GraphedStep")
        self.ddp, self.net, self.criterion, self.optimizer = model, model.module, criterion, optimizer
        self.eager = GraphedStep(model, criterion, optimizer)
        self.seen, self.plan, self.x, self.t, self.out3 = 0, None, None, None, None
        self._cap = torch.cuda.Stream()

    # sink interface used during capture: cut the graph at bucket boundaries
    def wants(self, offset):
        return self.ddp.wants(offset)

    def grad_ready(self, offset):
        d = self.ddp
        while d._next < len(d.buckets) and d.buckets[d._next][0] >= offset:
            lo, hi = d.buckets[d._next]
            self._ctx.__exit__(None, None, None)   # end of this graph ...
            self.plan.append(self._cur)
            self.plan.append((lo, hi))             # ... an eager all-reduce of the bucket goes here ...
            d._next += 1
            self._cur = torch.cuda.CUDAGraph()     # ... and the next graph begins
            self._ctx = torch.cuda.graph(self._cur, pool=self._pool, stream=self._cap,
                                         capture_error_mode="thread_local")
            self._ctx.__enter__()

    def backward_done(self):
        self.ddp._next = 0

    def _iteration(self):
        """forward + loss + backward + optimizer step on the static buffers, driving the network's own
        executor (no autograd); the current sink decides what happens at bucket boundaries."""
        from . import _cabi

        net = self.net
        logits, tape = net._forward_train(net._prep_input(self.x))
        self.out3, row_lse = _cabi.ce_topk_fwd(logits, self.t)
        dlogits = _cabi.ce_bwd(logits, self.t, row_lse, None, 1.0)
        net._backward(tape, dlogits)
        self.optimizer.step()

    def _manual_eager(self):
        """One real iteration through ``_iteration`` on the capture stream, un-captured: cuDNN / cuBLAS create
        per-stream resources lazily and must have seen this stream (and the side stream forked from it) before
        anything is captured on it."""
        cur = torch.cuda.current_stream()
        self._cap.wait_stream(cur)
        with torch.cuda.stream(self._cap), torch.no_grad():
            self._iteration()              # sink = the FlatDDP wrapper: eager all-reduces
            if self.ddp.broadcast_buffers:
                dist.broadcast(self.net.buffer_arena, 0, group=self.ddp.process_group)
        cur.wait_stream(self._cap)

    def _capture(self):
        net = self.net
        torch.cuda.synchronize()
        self.plan, self._pool = [], torch.cuda.graph_pool_handle()
        net.attach_sink(self)
        try:
            with torch.no_grad():
                self._cur = torch.cuda.CUDAGraph()
                self._ctx = torch.cuda.graph(self._cur, pool=self._pool, stream=self._cap,
                                             capture_error_mode="thread_local")
                self._ctx.__enter__()
                self._iteration()
                self._ctx.__exit__(None, None, None)
                self.plan.append(self._cur)
        finally:
            net.attach_sink(self.ddp)
        net.arena.steps -= 1
        torch.cuda.synchronize()

    def _replay(self):
        d, works, first = self.ddp, [], True
        g = self.net.arena.grads
        for item in self.plan[:-1]:
            if isinstance(item, tuple):
                works.append(dist.all_reduce(g[item[0]:item[1]], op=dist.ReduceOp.SUM, group=d.process_group,
                                             async_op=True))
            else:
                item.replay()
                if first and d.broadcast_buffers:
                    works.append(dist.broadcast(self.net.buffer_arena, 0, group=d.process_group, async_op=True))
                first = False
        for w in works:
            w.wait()
        self.plan[-1].replay()
        self.net.arena.steps += 1

    def __call__(self, x, t):
        if self.plan is None and self.seen < self.WARMUP:
            self.seen += 1
            return self.eager._eager(x, t)
        if self.x is None:  # the iteration after the eager warm-up: same code path as the capture, un-captured
            self.x, self.t = torch.empty_like(x), torch.empty_like(t)
            if hasattr(x, "_b200_s2d_stem"):
                self.x._b200_s2d_stem = x._b200_s2d_stem
            self.x.copy_(x)
            self.t.copy_(t)
            self.ddp.drain()
            self._manual_eager()
            return self.out3.clone()
        if x.shape != self.x.shape or t.shape != self.t.shape:
            return self.eager._eager(x, t)
        self.x.copy_(x)
        self.t.copy_(t)
        if self.plan is None:
            self._capture()
        self._replay()
        return self.out3.clone()


def _to_device(t):
    return t if t.is_cuda else t.cuda(non_blocking=True)


def _meter_weight(input):
    """The reference weights every meter update with ``input[0].size(0)`` (imagenet.py:142,199), which is the
    channel count of an NCHW image batch - 3 - not the batch size.  The space-to-depth stem layout of the input
    kernel carries 16 channels per cell for the same RGB image, so the reference's number is stated here."""
    return 3 if getattr(input, "_b200_s2d_stem", None) is not None else input[0].size(0)


def _raise_on_exchange_error(model):
    """FlatDDP(exchange="peer"): a rank that gave up waiting for a peer skips its updates and latches an error word;
    surface it where the host is synchronised anyway (end of an epoch) instead of training on silently."""
    check = getattr(model, "peer_error", None)
    if check is not None and check():
        raise RuntimeError("peer-memory gradient exchange timed out waiting for another rank; parameters were left "
                           "untouched from the failing iteration on - restart from the last checkpoint")


def train(train_loader, model, criterion, optimizer, epoch, world_size, sync_every=None, lag=0, graph=False):
    batch_time, data_time = AverageMeter(), AverageMeter()
    losses, top1, top5 = AverageMeter(), AverageMeter(), AverageMeter()
    model.train()
    log, iters = _MetricLog(world_size, (losses, top1, top5), sync_every, lag), 0
    graphed = None
    if graph:
        # one whole-iteration graph for a single process; with NCCL inside, the capture is invalidated on this
        # stack (torch 2.11, NCCL 2.28), so multi-process runs use a chain of graphs cut at the collectives
        # (FlatDDP(exchange="peer") is nothing but kernel launches, so it is a single graph at any world size)
        segmented = world_size > 1 and getattr(model, "exchange", "nccl") == "nccl"
        graphed = graph if isinstance(graph, (GraphedStep, SegmentedGraphStep)) else \
            (SegmentedGraphStep if segmented else GraphedStep)(model, criterion, optimizer)
    start = end = time.time()
    for i, (input, target) in enumerate(train_loader):
        data_time.update(time.time() - end)
        input, target = _to_device(input), _to_device(target)

        if graphed is not None:
            triple = graphed(input, target)
        else:
            logits = model(input)
            loss = criterion(logits, target)

            optimizer.zero_grad()
            loss.backward()
            optimizer.step()
            triple = _triple(criterion, loss, logits, target)

        iters += 1
        if log.add(triple, _meter_weight(input)):
            if lag == 0:
                torch.cuda.synchronize()  # imagenet.py:147
            batch_time.update(time.time() - end)
        end = time.time()
    log.finish()
    torch.cuda.synchronize()
    _raise_on_exchange_error(model)
    total = time.time() - start
    if batch_time.count < iters:  # spread the un-attributed wall time evenly; the sum is what is reported
        rest = iters - batch_time.count
        remaining = total - batch_time.sum
        for _ in range(rest):
            batch_time.update(remaining / rest)
    return losses.avg, top1.avg, top5.avg, batch_time.sum


def validate(val_loader, model, criterion, world_size, sync_every=None, lag=0):
    batch_time = AverageMeter()
    losses, top1, top5 = AverageMeter(), AverageMeter(), AverageMeter()
    model.eval()
    log, iters = _MetricLog(world_size, (losses, top1, top5), sync_every, lag), 0
    with torch.no_grad():
        start = end = time.time()
        for i, (input, target) in enumerate(val_loader):
            input, target = _to_device(input), _to_device(target)
            logits = model(input)
            loss = criterion(logits, target)
            iters += 1
            if log.add(_triple(criterion, loss, logits, target), _meter_weight(input)):
                if lag == 0:
                    torch.cuda.synchronize()
                batch_time.update(time.time() - end)
            end = time.time()
        log.finish()
        torch.cuda.synchronize()
        total = time.time() - start
    if batch_time.count < iters:
        rest = iters - batch_time.count
        remaining = total - batch_time.sum
        for _ in range(rest):
            batch_time.update(remaining / rest)
    return losses.avg, top1.avg, top5.avg, batch_time.sum
