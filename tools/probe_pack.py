"""GPU probe: the gradient pack kernel (K5) on ResNet-50's 53 convolution weight gradients (bf16 -> float32 arena),
4 against 8 vectors per thread, L2 flushed between launches.  python tools/probe_pack.py"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import imagent_distributed_training_pytorch_with_slurm_b200 as b200  # noqa: E402
from imagent_distributed_training_pytorch_with_slurm_b200 import _cabi as cabi  # noqa: E402

dev = "cuda"
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
model = b200.build_model("resnet50", dtype=torch.bfloat16).cuda()
a = model.arena
srcs, offs = [], []
for i, shp in enumerate(a.shapes):
    if len(shp) == 4:
        srcs.append(torch.randn(shp, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last))
        offs.append(a.offsets[i])
nbytes = sum(t.numel() * 6 for t in srcs)
for steps in (1, 2, 4, 8, 1, 2, 4):
    cabi.tune_set("pack_steps", steps)
    ts = []
    for _ in range(15):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        cabi.pack_grads(srcs, offs, a.grads)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    t = ts[len(ts) // 2]
    print(json.dumps({"pack_steps": steps, "us": round(t, 1), "MB": round(nbytes / 1e6, 1), "GBps": round(nbytes / t / 1e3)}))
cabi.tune_set("pack_steps", 1)
