"""Worker for tests/test_gpu_ddp_nccl.py (launched with torch.distributed.run, one rank per GPU):
3 training steps (small learning rate: at random init and tiny batches the first layers' gradients are huge
and chaotic, which would turn rounding differences into visible weight differences) of ResNet-18 (float32 activations) through FlatDDP + FlatSGD against stock
DistributedDataParallel + torch.optim.SGD on the same per-rank batches.  Rank 0 prints one JSON line."""
import json
import os
import sys

import torch
import torch.distributed as dist
import torchvision

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import imagent_distributed_training_pytorch_with_slurm_b200 as b200  # noqa: E402
from oracle import train_cpu  # noqa: E402


def main():
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False
    batches = [(x.cuda(), t.cuda()) for x, t in train_cpu.synthetic_batches(rank, 3, 16, 64, 1000)]

    torch.manual_seed(100 + rank)  # ranks start different: rank 0's weights must win in both implementations
    ours = b200.FlatDDP(b200.build_model("resnet18", dtype=torch.float32).cuda(), device_ids=[local],
                        bucket_cap_mb=8, first_bucket_mb=1)
    o_opt = b200.FlatSGD(ours.parameters(), 0.002, momentum=0.9, weight_decay=1e-4)
    o_res = b200.train(batches, ours, b200.CrossEntropyLoss().cuda(), o_opt, 0, world, sync_every=1)

    torch.manual_seed(100 + rank)
    ref = torch.nn.parallel.DistributedDataParallel(torchvision.models.resnet18(weights=None).cuda(), device_ids=[local])
    r_opt = torch.optim.SGD(ref.parameters(), 0.002, momentum=0.9, weight_decay=1e-4)
    trace = []
    r_res = train_cpu.train_epoch_cpu(batches, ref, torch.nn.CrossEntropyLoss().cuda(), r_opt, world, trace)

    osd, rsd = ours.state_dict(), ref.state_dict()
    worst, worst_key, num, den = 0.0, None, 0.0, 0.0
    for k in rsd:
        if not rsd[k].is_floating_point():
            continue
        a, b = osd[k].double(), rsd[k].double()
        num += float((a - b).pow(2).sum())
        den += float(b.pow(2).sum())
        if b.numel() >= 4096:  # small BatchNorm vectors start at 0/1 and are badly conditioned in relative terms
            rel = ((a - b).norm() / (b.norm() + 1e-12)).item()
            if rel > worst:
                worst, worst_key = rel, k
    # replicas must be identical across ranks after the steps (same averaged gradients everywhere)
    flat = torch.cat([v.flatten().float() for v in osd.values()])
    lo, hi = flat.clone(), flat.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    # the same run again, eager vs the chain of CUDA graphs cut at the collectives (8 iterations: 3 eager
    # warm-up, 1 capture + replay, 4 replays)
    graph_res = {}
    if os.environ.get("B200_TEST_GRAPH", "1") == "1":
        long_batches = [(x.cuda(), t.cuda()) for x, t in train_cpu.synthetic_batches(rank, 8, 16, 64, 1000, seed=77)]
        for use_graph in (False, True):
            torch.manual_seed(5)
            m = b200.FlatDDP(b200.build_model("resnet18", dtype=torch.bfloat16).cuda(), device_ids=[local],
                             bucket_cap_mb=8, first_bucket_mb=1)
            o = b200.FlatSGD(m.parameters(), 0.002, momentum=0.9, weight_decay=1e-4)
            r = b200.train(long_batches, m, b200.CrossEntropyLoss().cuda(), o, 0, world, sync_every=1, lag=1,
                           graph=use_graph)
            sd = m.state_dict()
            graph_res[str(use_graph)] = {"epoch": list(map(float, r[:3])), "steps": m.module.arena.steps,
                                         "w": float(sd["module.fc.weight"].double().norm()),
                                         "rv": float(sd["module.bn1.running_var"].double().norm()),
                                         "nbt": int(sd["module.bn1.num_batches_tracked"])}
    if rank == 0:
        print(json.dumps({"graph": graph_res, "ours": list(map(float, o_res[:3])), "stock": list(map(float, r_res[:3])),
                          "worst_rel_weight_diff": worst, "worst_key": worst_key, "global_rel_weight_diff": (num / den) ** 0.5, "keys_equal": list(osd) == list(rsd),
                          "replicas_identical": bool(torch.equal(lo, hi)), "buckets": len(ours.buckets)}))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
