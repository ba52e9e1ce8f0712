"""Oracle (test infrastructure) and timed CPU baseline: the reference's training step on host cores.

The reference has no CPU path as shipped: ``train()`` hard-codes ``.cuda()`` (imagenet.py:119-120),
the model and criterion are moved with ``.cuda()`` (``:312,323``) and every iteration ends in
``torch.cuda.synchronize()`` (``:147``).  This module restates ``train()`` (imagenet.py:97-151) and
``validate()`` (``:166-210``) step for step with only those device calls removed, driving the stock
objects the reference constructs: ``DistributedDataParallel(torchvision resnet)`` (``:312,316``),
``nn.CrossEntropyLoss()`` (``:323``), ``SGD(lr, momentum=0.9, weight_decay=1e-4)`` (``:325``), over
the gloo backend (``--backend`` flag, ``:440``).

Order inside one iteration, kept as in the reference: forward -> loss -> zero_grad -> backward
(DDP all-reduce of grads pre-divided by world size) -> optimizer.step -> accuracy(top1, top5) ->
three all-reduce means (loss, prec1, prec5) -> three ``.item()`` -> meters updated with weight
``input[0].size(0)`` (= 3 channels, imagenet.py:142) -> batch-time meter.
"""
import os
import time

import numpy as np
import torch
import torch.distributed as dist
import torch.nn as nn

from .step_helpers import MeterOracle


def _topk_percent(logits, target, ks):
    # accuracy(): imagenet.py:63-79
    with torch.no_grad():
        n = target.size(0)
        _, pred = logits.topk(max(ks), 1, True, True)
        hit = pred.t().eq(target.reshape(1, -1).expand_as(pred.t()))
        return [hit[:k].reshape(-1).float().sum(0, keepdim=True).mul_(100.0 / n) for k in ks]


def _mean_over_ranks(t, world_size):
    # reduce_tensor(): imagenet.py:82-87
    r = t.clone()
    if world_size > 1:
        dist.all_reduce(r, op=dist.ReduceOp.SUM)
    r /= world_size
    return r


def train_epoch_cpu(batches, model, criterion, optimizer, world_size, trace=None):
    """``batches`` yields (input float32 NCHW, target int64).  Returns what ``train()`` returns:
    (loss.avg, top1.avg, top5.avg, batch_time.sum).  ``trace`` (a list) receives the per-iteration
    (loss, prec1, prec5) python floats."""
    m_time, m_loss, m_top1, m_top5 = MeterOracle(), MeterOracle(), MeterOracle(), MeterOracle()
    model.train()
    end = time.time()
    for inp, tgt in batches:
        logits = model(inp)
        loss = criterion(logits, tgt)
        optimizer.zero_grad()
        loss.backward()
        optimizer.step()
        p1, p5 = _topk_percent(logits.data, tgt.data, (1, 5))
        rl = _mean_over_ranks(loss.data, world_size)
        p1 = _mean_over_ranks(p1, world_size)
        p5 = _mean_over_ranks(p5, world_size)
        w = inp[0].size(0)
        vals = (rl.item(), p1.item(), p5.item())
        m_loss.update(vals[0], w)
        m_top1.update(vals[1], w)
        m_top5.update(vals[2], w)
        if trace is not None:
            trace.append(vals)
        m_time.update(time.time() - end)
        end = time.time()
    return m_loss.avg, m_top1.avg, m_top5.avg, m_time.sum


def validate_epoch_cpu(batches, model, criterion, world_size, trace=None):
    m_time, m_loss, m_top1, m_top5 = MeterOracle(), MeterOracle(), MeterOracle(), MeterOracle()
    model.eval()
    with torch.no_grad():
        end = time.time()
        for inp, tgt in batches:
            logits = model(inp)
            loss = criterion(logits, tgt)
            p1, p5 = _topk_percent(logits.data, tgt.data, (1, 5))
            rl = _mean_over_ranks(loss.data, world_size)
            p1 = _mean_over_ranks(p1, world_size)
            p5 = _mean_over_ranks(p5, world_size)
            w = inp[0].size(0)
            vals = (rl.item(), p1.item(), p5.item())
            m_loss.update(vals[0], w)
            m_top1.update(vals[1], w)
            m_top5.update(vals[2], w)
            if trace is not None:
                trace.append(vals)
            m_time.update(time.time() - end)
            end = time.time()
    return m_loss.avg, m_top1.avg, m_top5.avg, m_time.sum


def build_reference_objects(arch, num_classes, lr, seed, world_size):
    """Model / criterion / optimizer exactly as run() builds them (imagenet.py:215,312-325),
    minus ``.cuda()``.  ``arch`` generalises the hard-coded resnet18."""
    import torchvision.models as models

    torch.manual_seed(seed)
    model = getattr(models, arch)(weights=None, num_classes=num_classes)
    if world_size > 1:
        model = nn.parallel.DistributedDataParallel(model)
    criterion = nn.CrossEntropyLoss()
    optimizer = torch.optim.SGD(model.parameters(), lr, momentum=0.9, weight_decay=1e-4)
    return model, criterion, optimizer


def synthetic_batches(rank, steps, batch, size, num_classes, seed=1234, fixed=False):
    """Seeded float32 NCHW batches + int64 targets (per-rank generator ``seed + rank``,
    SURVEY.md section 8d)."""
    g = torch.Generator()
    g.manual_seed(seed + rank)
    out = []
    for s in range(steps):
        if fixed and s > 0:
            out.append(out[0])
            continue
        x = torch.randn(batch, 3, size, size, generator=g)
        t = torch.randint(0, num_classes, (batch,), generator=g)
        out.append((x, t))
    return out


def _ddp_worker(rank, world_size, port, arch, num_classes, lr, seed, batches_per_rank, threads, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.set_num_threads(threads)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    try:
        model, criterion, optimizer = build_reference_objects(arch, num_classes, lr, seed, world_size)
        trace = []
        t0 = time.time()
        res = train_epoch_cpu(batches_per_rank[rank], model, criterion, optimizer, world_size, trace)
        wall = time.time() - t0
        if rank == 0:
            # numpy, not tensors: tensors cross the queue as file descriptors owned by this process,
            # which may have exited before the parent reads them
            sd = {k: v.detach().cpu().numpy().copy() for k, v in model.state_dict().items()}
            ret.put({"epoch": res, "trace": trace, "state_dict": sd, "wall": wall})
    finally:
        dist.destroy_process_group()


def run_ddp_oracle(arch, world_size, batches_per_rank, num_classes=1000, lr=0.1, seed=0, port=29611,
                   threads=None):
    """Run ``train_epoch_cpu`` on ``world_size`` gloo ranks (or in-process when 1) over the given
    per-rank batch lists.  Returns rank 0's {"epoch", "trace", "state_dict", "wall"}."""
    if threads is None:
        threads = max(1, len(os.sched_getaffinity(0)) // world_size)
    if world_size == 1:
        torch.set_num_threads(threads)
        model, criterion, optimizer = build_reference_objects(arch, num_classes, lr, seed, 1)
        trace = []
        t0 = time.time()
        res = train_epoch_cpu(batches_per_rank[0], model, criterion, optimizer, 1, trace)
        return {"epoch": res, "trace": trace, "wall": time.time() - t0,
                "state_dict": {k: v.detach().clone() for k, v in model.state_dict().items()}}
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    procs = [ctx.Process(target=_ddp_worker,
                         args=(r, world_size, port, arch, num_classes, lr, seed, batches_per_rank, threads, q))
             for r in range(world_size)]
    for p in procs:
        p.start()
    out = q.get()
    for p in procs:
        p.join()
    out["state_dict"] = {k: torch.from_numpy(np.asarray(v)) for k, v in out["state_dict"].items()}
    return out


def time_reference_step(arch, batch, size, steps, warmup, num_classes=1000, lr=0.1, seed=0):
    """Timed CPU baseline: images/sec of the restated step on all host cores of this process
    (single rank, no process group; the timing is ``time.time()`` around the step as
    imagenet.py:112,148-149)."""
    cores = len(os.sched_getaffinity(0))
    torch.set_num_threads(cores)
    model, criterion, optimizer = build_reference_objects(arch, num_classes, lr, seed, 1)
    data = synthetic_batches(0, 1, batch, size, num_classes)[0]
    train_epoch_cpu([data] * warmup, model, criterion, optimizer, 1)
    t0 = time.time()
    train_epoch_cpu([data] * steps, model, criterion, optimizer, 1)
    dt = time.time() - t0
    return {"images_per_sec": batch * steps / dt, "ms_per_step": 1e3 * dt / steps, "cores": cores,
            "sample": f"{arch} fp32 NCHW batch {batch} at {size}x{size}, {steps} steps after {warmup} warm-up"}


def _timed_ddp_worker(rank, world_size, port, arch, num_classes, lr, seed, batch, size, steps, warmup, threads, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.set_num_threads(threads)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    try:
        model, criterion, optimizer = build_reference_objects(arch, num_classes, lr, seed, world_size)
        data = synthetic_batches(rank, 1, batch, size, num_classes)[0]
        train_epoch_cpu([data] * warmup, model, criterion, optimizer, world_size)
        dist.barrier()
        t0 = time.time()
        res = train_epoch_cpu([data] * steps, model, criterion, optimizer, world_size)
        dist.barrier()
        dt = time.time() - t0
        if rank == 0:
            ret.put({"seconds": dt, "loss_avg": float(res[0])})
    finally:
        dist.destroy_process_group()


def time_reference_ddp(arch, batch_per_rank, size, steps, warmup, world_size=2, num_classes=1000, lr=0.1, seed=0,
                       port=29633):
    """Timed CPU baseline as SURVEY 8d / BASELINE.md 3 define it: the restated step on ``world_size`` gloo ranks
    (the reference's ``--backend gloo`` data-parallel path, imagenet.py:316,440), ``cores // world_size`` threads per
    rank, one fixed synthetic fp32 NCHW batch per rank, ``time.time()`` around the timed iterations with a barrier on
    both sides.  Returns global images/sec with the core count."""
    import torch.multiprocessing as mp

    cores = len(os.sched_getaffinity(0))
    threads = max(1, cores // world_size)
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    procs = [ctx.Process(target=_timed_ddp_worker, args=(r, world_size, port, arch, num_classes, lr, seed, batch_per_rank,
                                                         size, steps, warmup, threads, q))
             for r in range(world_size)]
    for p in procs:
        p.start()
    for p in procs:
        p.join()
    if any(p.exitcode != 0 for p in procs) or q.empty():
        raise RuntimeError(f"gloo reference ranks failed (exit codes {[p.exitcode for p in procs]})")
    out = q.get()
    dt = out["seconds"]
    return {"images_per_sec": batch_per_rank * world_size * steps / dt, "ms_per_step": 1e3 * dt / steps, "cores": cores,
            "threads_per_rank": threads, "world_size": world_size, "loss_avg": out["loss_avg"],
            "sample": f"{arch} fp32 NCHW, DDP over gloo world_size {world_size} x {threads} threads, batch "
                      f"{batch_per_rank}/rank at {size}x{size}, {steps} steps after {warmup} warm-up"}
