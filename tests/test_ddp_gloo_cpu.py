"""world_size-2 data-parallel path on CPU over gloo, with the oracle standing in for the CUDA
kernels: FlatDDP's rank-0 parameter/buffer broadcast, bucketed all-reduce of the flat gradient arena
with the 1/world pre-division, the one-collective metric reduction in train(), and the resulting
trajectory against the stock-DDP golden (tests/golden/golden.json: ddp_resnet18_w2)."""
import json
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, gold, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import imagent_distributed_training_pytorch_with_slurm_b200 as b200
        import imagent_distributed_training_pytorch_with_slurm_b200._cabi as cabi
        import imagent_distributed_training_pytorch_with_slurm_b200.engine as eng
        from oracle import train_cpu
        from tests import cpu_kernel_standins as s

        for name in ("bn_fwd_train", "bn_fwd_eval", "bn_bwd", "ce_topk_fwd", "ce_bwd", "sgd_flat", "input_stage", "maxpool_fwd", "bn_relu_maxpool_fwd",
                 "maxpool_bwd", "pack_grads", "bn_dual_supported", "bn_stats", "bn_fwd_train_dual", "bn_bwd_dual"):
            setattr(cabi, name, getattr(s, name))
        eng._to_device = lambda t: t
        torch.cuda.synchronize = lambda *a, **k: None

        torch.manual_seed(gold["seed"] + 100 * rank)  # ranks start DIFFERENT: rank 0's init must win
        model = b200.build_model("resnet18", device="cpu", dtype=torch.float32)
        model = b200.FlatDDP(model, bucket_cap_mb=8, first_bucket_mb=1)
        crit = b200.CrossEntropyLoss()
        opt = b200.FlatSGD(model.parameters(), gold["lr"], momentum=0.9, weight_decay=1e-4)
        batches = train_cpu.synthetic_batches(rank, gold["steps"], gold["batch_per_rank"], gold["image"], 1000,
                                              seed=gold["data_seed"])
        res = b200.train(batches, model, crit, opt, 0, world, sync_every=1)
        sd = model.state_dict()
        out = {"rank": rank, "epoch": [float(v) for v in res[:3]], "n_buckets": len(model.buckets),
               "bucket_bytes": [(hi - lo) * 4 for lo, hi in model.buckets],
               "l2": {k: float(sd[k].double().norm()) for k in gold["weight_l2"]},
               "nbt": int(sd["module.bn1.num_batches_tracked"]),
               "bn_rv": sd["module.bn1.running_var"].tolist()}
        q.put(out)
    except BaseException as e:  # the parent must hear about it: a silent worker death would leave it waiting
        import traceback

        q.put({"rank": rank, "error": "".join(traceback.format_exception(type(e), e, e.__traceback__))})
        raise
    finally:
        dist.destroy_process_group()


def test_flatddp_two_ranks_match_stock_ddp_golden():
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "golden.json")))["ddp_resnet18_w2"]
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    procs = [ctx.Process(target=_worker, args=(r, 2, 29641, gold, q)) for r in range(2)]
    for p in procs:
        p.start()
    outs = []
    try:
        import time

        deadline = time.time() + 600
        while len(outs) < 2:
            if not q.empty():
                outs.append(q.get())
                if "error" in outs[-1]:
                    break                # the other rank is stuck in a collective: do not wait for it
            elif time.time() > deadline or all(not p.is_alive() for p in procs):
                break
            else:
                time.sleep(0.2)
    finally:
        for p in procs:
            p.join(5)
            if p.is_alive():
                p.terminate()
    errors = [o["error"] for o in outs if "error" in o]
    assert not errors, errors[0]
    assert len(outs) == 2, "a worker exited without reporting"
    outs = sorted(outs, key=lambda o: o["rank"])
    r0, r1 = outs
    # epoch averages (mean over ranks of loss / prec, unweighted over iterations) == stock DDP run
    for a, b in zip(r0["epoch"], gold["epoch_avg"]):
        assert a == pytest.approx(b, rel=2e-4, abs=1e-6)
    assert r0["epoch"] == r1["epoch"]                      # every rank logs the same reduced values
    # updated weights: replicas identical across ranks and equal to the stock trajectory
    for k, v in gold["weight_l2"].items():
        # (the last block's BN bias starts at 0 and is a sum of three tiny, badly conditioned updates)
        assert r0["l2"][k] == pytest.approx(v, rel=5e-3 if k.endswith("bn2.bias") else 2e-4), k
        assert r0["l2"][k] == pytest.approx(r1["l2"][k], rel=1e-6), k
    assert r0["nbt"] == gold["nbt"] == gold["steps"]
    # BatchNorm buffers: rank 0's running statistics were broadcast to rank 1 (DDP _sync_buffers)
    assert r0["bn_rv"] == r1["bn_rv"]
    # buckets: several slices covering the arena, first (last layers) starts communication early
    assert r0["n_buckets"] >= 3 and sum(r0["bucket_bytes"]) >= 46_758_048
