"""GPU soak: a few hundred graph-replayed iterations of ResNet-50 on a learnable synthetic stream; prints the loss
every 25 iterations (it must fall and stay finite) and the peak memory."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import imagent_distributed_training_pytorch_with_slurm_b200 as b200  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
torch.backends.cudnn.benchmark = True
torch.manual_seed(0)
model = b200.build_model("resnet50").cuda()
crit = b200.CrossEntropyLoss().cuda()
opt = b200.FlatSGD(model.parameters(), 0.02, momentum=0.9, weight_decay=1e-4)
g = torch.Generator().manual_seed(7)
protos = torch.randn(100, 3, 64, 64, generator=g)


def batches(n):
    for _ in range(n):
        t = torch.randint(0, 100, (128,), generator=g)
        yield (protos[t] * 0.4 + torch.randn(128, 3, 64, 64, generator=g)).cuda(), t.cuda()


runner = b200.GraphedStep(model, crit, opt)
log = []
for i, (x, t) in enumerate(batches(steps)):
    log.append(runner(x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last), t)[0])
    if (i + 1) % 25 == 0:
        print(i + 1, round(float(torch.stack(log[-25:]).mean()), 4), flush=True)
vals = torch.stack(log).float().cpu()
print("finite:", bool(torch.isfinite(vals).all()), "first/last window:", float(vals[:25].mean()), float(vals[-25:].mean()),
      "peak GB:", round(torch.cuda.max_memory_allocated() / 2**30, 2))
