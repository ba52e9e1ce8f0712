"""GPU probe: microseconds per BatchNorm site, one cooperative kernel (bn_fused=1) against the three-kernel chain
(bn_fused=0), for the ResNet-50 shapes at batch 256.  Two conditions: 'cold' (L2 flushed before the site) and
'after-producer' (the activation was just written by another kernel, as a convolution does in the step).
    python tools/probe_bn_fused.py > gpurun_out/probe_bn_fused.jsonl"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from imagent_distributed_training_pytorch_with_slurm_b200 import _cabi as cabi  # noqa: E402

dev = "cuda"
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
SHAPES = [(256, 64, 112, 112), (256, 64, 56, 56), (256, 256, 56, 56), (256, 128, 28, 28), (256, 512, 28, 28),
          (256, 256, 14, 14), (256, 1024, 14, 14), (256, 512, 7, 7), (256, 2048, 7, 7)]


def timeit(fn, prep, reps=15):
    ts = []
    for _ in range(reps):
        prep()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    for shape in SHAPES:
        n, c, h, w = shape
        src = torch.randn(n, h, w, c, device=dev).to(torch.bfloat16).permute(0, 3, 1, 2)
        x, z, dy, dy2 = (torch.empty_like(src) for _ in range(4))
        for t in (x, z, dy, dy2):
            t.copy_(src)
        g, b = torch.ones(c, device=dev), torch.zeros(c, device=dev)
        rm, rv = torch.zeros(c, device=dev), torch.ones(c, device=dev)
        mb = x.numel() * 2 / 1e6
        row = {"shape": shape, "MB": round(mb, 1)}
        for cond, prep in (("cold", lambda: flush.zero_()),
                           ("after_producer", lambda: (flush.zero_(), x.copy_(src), dy.copy_(src)))):
            for fused in (0, 1):
                cabi.tune_set("bn_fused", fused)
                tag = f"{cond}_{'fused' if fused else 'chain'}"
                _, saved, _ = cabi.bn_fwd_train(x, None, g, b, rm, rv, None, 1e-5, 0.1, 1)
                row[f"fwd_relu_us_{tag}"] = round(timeit(lambda: cabi.bn_fwd_train(x, None, g, b, rm, rv, None, 1e-5, 0.1, 1), prep), 1)
                _, saved2, mask = cabi.bn_fwd_train(x, z, g, b, rm, rv, None, 1e-5, 0.1, 2)
                row[f"fwd_addrelu_us_{tag}"] = round(timeit(lambda: cabi.bn_fwd_train(x, z, g, b, rm, rv, None, 1e-5, 0.1, 2), prep), 1)
                row[f"bwd_relu_us_{tag}"] = round(timeit(lambda: cabi.bn_bwd(dy, x, None, g, saved, 1), prep), 1)
                row[f"bwd_addrelu2_us_{tag}"] = round(timeit(lambda: cabi.bn_bwd(dy, x, mask, g, saved2, 2, dy2=dy2), prep), 1)
        row["ideal_us_fwd_relu"] = round(3 * mb / 6.5395, 1)
        row["ideal_us_bwd_relu"] = round(5 * mb / 6.5395, 1)
        print(json.dumps(row), flush=True)
    cabi.tune_set("bn_fused", 0)


if __name__ == "__main__":
    main()
