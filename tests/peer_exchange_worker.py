"""Worker for tests/test_gpu_peer_exchange.py (torch.distributed.run, one rank per GPU of ONE box).

1. kernel level: b200_peer_exchange_sgd + b200_peer_wait on synthetic symmetric arenas against the closed form
   (sum of the ranks' gradients in rank order, then the update of b200_sgd_flat_f32) - bit-exact;
2. model level: ResNet-18 (float32 activations), FlatDDP(exchange="peer") against FlatDDP(exchange="nccl") on the same
   seeds and per-rank batches for 3 eager steps;
3. the exchange of a ResNet-50-sized arena timed alone.
Rank 0 prints one JSON line.  Every wait inside the kernels times out, so a rank that dies cannot hang the others."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import imagent_distributed_training_pytorch_with_slurm_b200 as b200  # noqa: E402
from imagent_distributed_training_pytorch_with_slurm_b200 import _cabi  # noqa: E402
from oracle import train_cpu  # noqa: E402


def kernel_level(rank, world, dev, n, buckets, steps=3, lr=0.05, wd=1e-4, mu=0.9):
    import torch.distributed._symmetric_memory as symm_mem

    g = torch.Generator(device="cpu").manual_seed(11)
    p0 = torch.randn(n, generator=g)
    grads_all = [[torch.randn(n, generator=torch.Generator().manual_seed(1000 * s + r)) for r in range(world)]
                 for s in range(steps)]
    grad = symm_mem.empty(n, dtype=torch.float32, device=dev).zero_()
    param = symm_mem.empty(n, dtype=torch.float32, device=dev)
    shadow = symm_mem.empty(n, dtype=torch.bfloat16, device=dev).zero_()
    pad = symm_mem.empty(_cabi.peer_pad_words(), dtype=torch.int32, device=dev).zero_()
    param.copy_(p0)
    mom = torch.zeros(n, device=dev)
    hs = [symm_mem.rendezvous(t, dist.group.WORLD) for t in (grad, param, shadow, pad)]
    tabs = [_cabi.peer_table(h.buffer_ptrs) for h in hs]
    torch.cuda.synchronize()
    dist.barrier()
    # closed form on the CPU in float32, summing in rank order like the kernel
    ref_p, ref_m = p0.clone(), torch.zeros(n)
    for s in range(steps):
        grad.copy_(grads_all[s][rank])
        for b, (lo, hi) in enumerate(buckets):
            _cabi.peer_exchange_sgd(tabs[0], tabs[1], tabs[2], tabs[3], mom, rank, world, b, lo, hi, lr, wd, mu, 1.0,
                                    s == 0, 16)
        _cabi.peer_wait(pad, world, len(buckets), rank)
        tot = torch.zeros(n)
        for r in range(world):
            tot = tot + grads_all[s][r]
        gp = torch.addcmul(tot, torch.full_like(ref_p, wd), ref_p)   # fma(wd, p, g) up to contraction
        ref_m = gp.clone() if s == 0 else ref_m * mu + gp
        ref_p = ref_p - lr * ref_m
    torch.cuda.synchronize()
    got = param.cpu()
    # the kernel contracts to FMAs; compare with a one-ulp-of-the-largest-operand bound, and bit-exactly across ranks
    scale = float(torch.maximum(ref_p.abs(), (lr * ref_m).abs()).max())
    err = float((got - ref_p).abs().max())
    lo_t, hi_t = param.clone(), param.clone()
    dist.all_reduce(lo_t, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi_t, op=dist.ReduceOp.MAX)
    return {"max_err": err, "bound": 8 * scale * 2 ** -23 * steps, "replicas_identical": bool(torch.equal(lo_t, hi_t)),
            "shadow_ok": bool(torch.equal(shadow, param.to(torch.bfloat16))), "error_word": int(pad[0].item()),
            "step_counter": int(pad[1].item())}


def model_level(rank, local, world):
    batches = [(x.cuda(), t.cuda()) for x, t in train_cpu.synthetic_batches(rank, 3, 16, 64, 1000)]
    out = {}
    for mode in ("nccl", "peer"):
        torch.manual_seed(100 + rank)
        net = b200.build_model("resnet18", dtype=torch.float32, symmetric=(mode == "peer")).cuda()
        m = b200.FlatDDP(net, device_ids=[local], bucket_cap_mb=8, first_bucket_mb=1, exchange=mode)
        o = b200.FlatSGD(m.parameters(), 0.002, momentum=0.9, weight_decay=1e-4)
        m.attach_optimizer(o)
        r = b200.train(batches, m, b200.CrossEntropyLoss().cuda(), o, 0, world, sync_every=1)
        torch.cuda.synchronize()
        sd = m.state_dict()
        flat = torch.cat([v.flatten().float() for v in sd.values() if v.is_floating_point()])
        lo_t, hi_t = flat.clone(), flat.clone()
        dist.all_reduce(lo_t, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi_t, op=dist.ReduceOp.MAX)
        out[mode] = {"epoch": list(map(float, r[:3])), "flat": flat, "identical": bool(torch.equal(lo_t, hi_t)),
                     "steps": m.module.arena.steps, "error": bool(m.peer_error()), "buckets": len(m.buckets)}
    a, b = out["peer"]["flat"].double(), out["nccl"]["flat"].double()
    rel = float((a - b).norm() / b.norm())
    for mode in out:
        del out[mode]["flat"]
    out["global_rel_weight_diff"] = rel
    return out


def _same_on_all_ranks(t):
    lo_t, hi_t = t.clone().float(), t.clone().float()
    dist.all_reduce(lo_t, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi_t, op=dist.ReduceOp.MAX)
    return bool(torch.equal(lo_t, hi_t))


def graph_level(rank, local, world):
    """8 iterations of ResNet-18 (bf16): NCCL path eager vs peer path as ONE CUDA graph per iteration (3 eager warm-up
    iterations, capture, replays), both with the per-iteration buffer broadcast from rank 0 (NCCL / peer copy)."""
    batches = [(x.cuda(), t.cuda()) for x, t in train_cpu.synthetic_batches(rank, 8, 16, 64, 1000, seed=77)]
    out = {}
    # "peer_auto": the public switch, train(..., graph=True) - it must pick the single-graph runner for the peer path
    # (the chain-of-graphs runner records NCCL all-reduces + a whole-arena SGD and would corrupt the rank-sharded momentum)
    for mode in ("nccl", "peer", "peer_auto"):
        exchange = "nccl" if mode == "nccl" else "peer"
        torch.manual_seed(5)
        net = b200.build_model("resnet18", dtype=torch.bfloat16, symmetric=(exchange == "peer")).cuda()
        m = b200.FlatDDP(net, device_ids=[local], bucket_cap_mb=8, first_bucket_mb=1, exchange=exchange)
        o = b200.FlatSGD(m.parameters(), 0.002, momentum=0.9, weight_decay=1e-4)
        m.attach_optimizer(o)
        runner = b200.GraphedStep(m, b200.CrossEntropyLoss().cuda(), o) if mode == "peer" else (mode == "peer_auto")
        crit = runner.criterion if mode == "peer" else b200.CrossEntropyLoss().cuda()
        r = b200.train(batches, m, crit, o, 0, world, sync_every=1, lag=1, graph=runner)
        torch.cuda.synchronize()
        sd = m.state_dict()
        out[mode] = {"epoch": list(map(float, r[:3])), "steps": m.module.arena.steps,
                     "w": float(sd["module.fc.weight"].double().norm()),
                     "w1": float(sd["module.conv1.weight"].double().norm()),
                     "rv": float(sd["module.bn1.running_var"].double().norm()),
                     "rv_all": _same_on_all_ranks(sd["module.layer4.1.bn2.running_var"]),
                     "captured": bool(mode == "peer" and runner.graph is not None), "error": bool(m.peer_error())}
    try:   # and the chain-of-graphs runner must refuse the peer path outright
        b200.SegmentedGraphStep(m, b200.CrossEntropyLoss().cuda(), o)
        out["segmented_refuses_peer"] = False
    except ValueError:
        out["segmented_refuses_peer"] = True
    return out


def timeout_level(rank, world, dev, n=1 << 16):
    """A peer that never shows up: rank 0 launches the exchange of one bucket, the other ranks launch nothing.  Rank
    0's kernel must give up after its ~20 s wait, latch the error word and leave EVERY arena untouched (its own and,
    through the peer mappings, everybody else's); the iteration counter must not advance; later launches return at
    once without touching anything either."""
    import time

    import torch.distributed._symmetric_memory as symm_mem

    p0 = torch.randn(n, generator=torch.Generator().manual_seed(3))
    grad = symm_mem.empty(n, dtype=torch.float32, device=dev).fill_(1.0)
    param = symm_mem.empty(n, dtype=torch.float32, device=dev)
    shadow = symm_mem.empty(n, dtype=torch.bfloat16, device=dev).zero_()
    pad = symm_mem.empty(_cabi.peer_pad_words(), dtype=torch.int32, device=dev).zero_()
    param.copy_(p0)
    mom = torch.zeros(n, device=dev)
    tabs = [_cabi.peer_table(symm_mem.rendezvous(t, dist.group.WORLD).buffer_ptrs) for t in (grad, param, shadow, pad)]
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.time()
    if rank == 0:
        _cabi.peer_exchange_sgd(tabs[0], tabs[1], tabs[2], tabs[3], mom, rank, world, 0, 0, n, 0.1, 1e-4, 0.9, 1.0, True, 8)
        _cabi.peer_wait(pad, world, 1, rank)
        torch.cuda.synchronize()
        waited = time.time() - t0
        # latched: a second attempt returns immediately
        t1 = time.time()
        _cabi.peer_exchange_sgd(tabs[0], tabs[1], tabs[2], tabs[3], mom, rank, world, 0, 0, n, 0.1, 1e-4, 0.9, 1.0, False, 8)
        _cabi.peer_wait(pad, world, 1, rank)
        torch.cuda.synchronize()
        second = time.time() - t1
    else:
        waited = second = 0.0
    dist.barrier()      # (a host-side barrier over NCCL's own stream: the other ranks just wait here for rank 0)
    torch.cuda.synchronize()
    ok = torch.tensor([float(torch.equal(param.cpu(), p0) and float(mom.abs().sum()) == 0.0
                             and float(shadow.float().abs().sum()) == 0.0)], device=dev)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    return {"error_word_rank0": int(pad[0].item()) if rank == 0 else None, "step_counter": int(pad[1].item()),
            "all_arenas_untouched_on_all_ranks": bool(ok.item() == 1.0), "waited_s": waited, "second_attempt_s": second}


def timing(rank, world, dev, n=25557056):
    import torch.distributed._symmetric_memory as symm_mem

    grad = symm_mem.empty(n, dtype=torch.float32, device=dev).zero_()
    param = symm_mem.empty(n, dtype=torch.float32, device=dev).zero_()
    shadow = symm_mem.empty(n, dtype=torch.bfloat16, device=dev).zero_()
    pad = symm_mem.empty(_cabi.peer_pad_words(), dtype=torch.int32, device=dev).zero_()
    mom = torch.zeros(n, device=dev)
    tabs = [_cabi.peer_table(symm_mem.rendezvous(t, dist.group.WORLD).buffer_ptrs) for t in (grad, param, shadow, pad)]
    torch.cuda.synchronize()
    dist.barrier()
    res = {}
    for ctas in (16, 32, 64, 148):
        def once():
            _cabi.peer_exchange_sgd(tabs[0], tabs[1], tabs[2], tabs[3], mom, rank, world, 0, 0, n, 0.0, 0.0, 0.9, 1.0,
                                    False, ctas)
            _cabi.peer_wait(pad, world, 1, rank)

        for _ in range(3):
            once()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            once()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / 10], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        # NCCL-tests convention for the all-reduce this replaces: bytes = the fp32 gradient arena
        res[str(ctas)] = {"ms": ms, "busbw_GBps": n * 4 / 1e9 / (ms / 1e3) * 2 * (world - 1) / world}
    res["error_word"] = int(pad[0].item())
    return res


def main():
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False
    n = 1 << 20
    buckets = [(n - 4096, n), (n // 2 + 128, n - 4096), (0, n // 2 + 128)]   # uneven, all multiples of 4
    out = {"kernel": kernel_level(rank, world, dev, n, buckets)}
    if os.environ.get("B200_PEER_MODEL", "1") == "1":
        out["model"] = model_level(rank, local, world)
    if os.environ.get("B200_PEER_GRAPH", "1") == "1":
        out["graph"] = graph_level(rank, local, world)
    if os.environ.get("B200_PEER_TIMING", "1") == "1":
        out["timing"] = timing(rank, world, dev)
    if os.environ.get("B200_PEER_TIMEOUT", "0") == "1":     # last: it leaves an error latched in its own pad only
        out["timeout"] = timeout_level(rank, world, dev)
    if rank == 0:
        print(json.dumps(out))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
