"""GPU probe: capture one whole training step (input kernel, forward, loss, backward, SGD) in a CUDA graph and
compare replay time with the eager loop."""
import os
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
    return loss


def timeit(fn, n=20):
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


s = torch.cuda.Stream()
s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    for _ in range(5):
        step()
torch.cuda.current_stream().wait_stream(s)
print(arch, "eager ms/step", round(timeit(step), 3), flush=True)
graph = torch.cuda.CUDAGraph()
try:
    with torch.cuda.graph(graph):
        static_loss = step()
    print("captured", flush=True)
    print(arch, "graph ms/step", round(timeit(graph.replay), 3), "loss", float(static_loss), flush=True)
    l0 = float(static_loss)
    for _ in range(30):
        graph.replay()
    print("loss after 30 more replays on the same batch", float(static_loss), "(was", l0, ")")
except Exception as e:  # noqa: BLE001
    print("capture failed:", type(e).__name__, str(e)[:600])
