"""GPU probe (torchrun, >= 2 GPUs): does train(..., graph=True) capture and replay with NCCL collectives inside?"""
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import imagent_distributed_training_pytorch_with_slurm_b200 as b200  # noqa: E402

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
torch.backends.cudnn.benchmark = True
torch.manual_seed(0)
arch = sys.argv[1] if len(sys.argv) > 1 else "resnet50"
model = b200.FlatDDP(b200.build_model(arch).cuda(), device_ids=[local])
crit = b200.CrossEntropyLoss().cuda()
opt = b200.FlatSGD(model.parameters(), 0.1, momentum=0.9, weight_decay=1e-4)
for g in (False, True):
    loader = b200.SyntheticImageNet(256, 24, 256, 224, rank=rank, host=True)
    try:
        t0 = time.time()
        res = b200.train(loader, model, crit, opt, 0, world, sync_every=1, lag=1, graph=g)
        dt = time.time() - t0
        if rank == 0:
            print(f"graph={g}: loss {res[0]:.4f}, {24 * 256 * world / dt:.0f} img/s incl. warm-up/capture", flush=True)
    except Exception as e:  # noqa: BLE001
        print(f"rank {rank} graph={g} failed: {type(e).__name__}: {str(e)[:500]}", flush=True)
        break
dist.destroy_process_group()
