"""GPU probe: max-pool (K6) forward/backward and the input stage (K1) at the bench shapes, timed alone with CUDA
events (an L2-sized buffer is overwritten between launches).  `--ncu` runs each kernel twice only (for a profiler)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from imagent_distributed_training_pytorch_with_slurm_b200 import _cabi, data  # noqa: E402

quick = "--ncu" in sys.argv
reps = 2 if quick else 20
dev = "cuda"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn):
    ts = []
    for i in range(reps + 2):
        flush.fill_(i & 1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        if i >= 2:
            ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


n = 256
x = torch.randn(n, 64, 112, 112, device=dev).relu().to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
y, arg = _cabi.maxpool_fwd(x)
gy = torch.randn_like(y)
fb = x.numel() * 2 + y.numel() * 3
t = timeit(lambda: _cabi.maxpool_fwd(x))
print(f"maxpool_fwd {t:8.1f} us  {fb / t / 1e3:7.0f} GB/s")
t = timeit(lambda: _cabi.maxpool_bwd(gy, arg, x.shape))
print(f"maxpool_bwd {t:8.1f} us  {fb / t / 1e3:7.0f} GB/s")

g = torch.Generator().manual_seed(1234)
imgs = torch.randint(0, 256, (n, 256, 256, 3), dtype=torch.uint8, generator=g).to(dev)
boxes, flips = data.draw_batch_params(n, 256, 256, g, random_crop=True)
boxes, flips = boxes.to(dev), flips.to(dev)
rd = int((boxes[:, 2].long() * boxes[:, 3].long()).sum()) * 3
for layout, name in ((_cabi.LAYOUT_S2D_STEM, "s2d"), (_cabi.LAYOUT_NHWC, "nhwc")):
    out = _cabi.input_stage(imgs, boxes, flips, 224, 224, (0.485, 0.456, 0.406), (0.229, 0.224, 0.225), torch.bfloat16, layout)
    by = rd + out.numel() * 2
    t = timeit(lambda: _cabi.input_stage(imgs, boxes, flips, 224, 224, (0.485, 0.456, 0.406), (0.229, 0.224, 0.225),
                                         torch.bfloat16, layout))
    print(f"input {name:5s} {t:8.1f} us  {by / t / 1e3:7.0f} GB/s  ({by / 1e6:.1f} MB)")
