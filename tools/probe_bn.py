"""GPU probe: per-shape GB/s of the BatchNorm kernels under different tuning knobs (L2 flushed
between launches), to pick defaults.  Run on the B200 box: python tools/probe_bn.py"""
import itertools
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from imagent_distributed_training_pytorch_with_slurm_b200 import _cabi as cabi  # noqa: E402

dev = "cuda"
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
SHAPES = [(256, 64, 112, 112), (256, 64, 56, 56), (256, 128, 28, 28), (256, 256, 14, 14), (256, 1024, 14, 14),
          (256, 2048, 7, 7), (256, 512, 7, 7)]


def timeit(fn, reps=9):
    ts = []
    for _ in range(reps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    # (vec, direct-kernel CTAs, reverse, pipeline, pipeline CTAs/SM)
    knobs = [(8, 1184, 1, 1, 2, 0), (8, 1184, 1, 1, 2, 1)]
    if os.environ.get("PROBE_QUICK"):
        knobs = knobs[1:]
    res = []
    for shape in SHAPES:
        n, c, h, w = shape
        x = torch.randn(n, h, w, c, device=dev).to(torch.bfloat16).permute(0, 3, 1, 2)
        z = torch.randn(n, h, w, c, device=dev).to(torch.bfloat16).permute(0, 3, 1, 2)
        dy = torch.randn(n, h, w, c, device=dev).to(torch.bfloat16).permute(0, 3, 1, 2)
        g, b = torch.ones(c, device=dev), torch.zeros(c, device=dev)
        rm, rv = torch.zeros(c, device=dev), torch.ones(c, device=dev)
        e = x.numel() * 2
        for vec, ctas, rev, pipe, pctas, pdl in knobs:
            cabi.tune_set("bn_pdl", pdl)
            cabi.tune_set("bn_vec_bf16", vec)
            cabi.tune_set("bn_target_ctas", ctas)
            cabi.tune_set("bn_stats_reverse", rev)
            cabi.tune_set("bn_pipeline", pipe)
            cabi.tune_set("bn_pipe_ctas_per_sm", pctas)
            row = {"shape": shape, "vec": vec, "ctas": ctas, "rev": rev, "pipe": pipe, "pctas": pctas, "pdl": pdl}
            for act, name in ((1, "relu"), (2, "add_relu")):
                zz = z if act == 2 else None
                _, saved, mask = cabi.bn_fwd_train(x, zz, g, b, rm, rv, None, 1e-5, 0.1, act)
                t = timeit(lambda: cabi.bn_fwd_train(x, zz, g, b, rm, rv, None, 1e-5, 0.1, act))
                row[f"fwd_{name}_GBps"] = round((3 + (act == 2)) * e / t / 1e6)
                t = timeit(lambda: cabi.bn_bwd(dy, x, mask, g, saved, act))
                row[f"bwd_{name}_GBps"] = round((5 + (act == 2)) * e / t / 1e6)
            t = timeit(lambda: cabi.bn_fwd_eval(x, None, g, b, rm, rv, 1e-5, 1))
            row["eval_relu_GBps"] = round(2 * e / t / 1e6)
            res.append(row)
            print(json.dumps(row), flush=True)
        # yardsticks: torch copy and torch read-only reduction on the same tensor
        y = torch.empty_like(x)
        t = timeit(lambda: y.copy_(x))
        t2 = timeit(lambda: x.float().sum()) if x.numel() < (1 << 28) else None
        print(json.dumps({"shape": shape, "torch_copy_GBps": round(2 * e / t / 1e6)}), flush=True)


if __name__ == "__main__":
    main()
