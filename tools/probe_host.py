"""GPU probe: pure host cost of enqueuing one training step (queue empty), and a cProfile of it."""
import cProfile
import os
import pstats
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import imagent_distributed_training_pytorch_with_slurm_b200 as b200  # noqa: E402
from imagent_distributed_training_pytorch_with_slurm_b200 import _cabi  # noqa: E402

arch = sys.argv[1] if len(sys.argv) > 1 else "resnet50"
torch.backends.cudnn.benchmark = True
dev = torch.device("cuda", 0)
torch.manual_seed(0)
model = b200.FlatDDP(b200.build_model(arch, 1000, dev).cuda())
crit = b200.CrossEntropyLoss()
opt = b200.FlatSGD(model.parameters(), 0.1, momentum=0.9, weight_decay=1e-4)
g = torch.Generator().manual_seed(0)
images = torch.randint(0, 256, (256, 256, 256, 3), dtype=torch.uint8, generator=g).to(dev)
boxes, flips = b200.draw_batch_params(256, 256, 256, g)
boxes, flips = boxes.to(dev), flips.to(dev)
tgt = torch.randint(0, 1000, (256,), generator=g).to(dev)
mean, std = b200.data.IMAGENET_MEAN, b200.data.IMAGENET_STD


def step():
    x = _cabi.input_stage(images, boxes, flips, 224, 224, mean, std, torch.bfloat16, _cabi.LAYOUT_S2D_STEM)
    loss = crit(model(x), tgt)
    opt.zero_grad()
    loss.backward()
    opt.step()


for _ in range(5):
    step()
torch.cuda.synchronize()
ts = []
for _ in range(5):
    t0 = time.perf_counter()
    step()
    ts.append((time.perf_counter() - t0) * 1e3)
    torch.cuda.synchronize()
print(arch, "host ms per step with an empty queue:", [round(t, 2) for t in ts])
pr = cProfile.Profile()
pr.enable()
for _ in range(3):
    step()
    torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
