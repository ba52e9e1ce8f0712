"""Loss-curve parity over 200 training steps on a learnable synthetic stream (SURVEY section 8c:
"epoch-level train loss within +-3 % of an fp32 PyTorch-eager run of the same synthetic stream over
>= 200 steps").  Stock objects = what the reference builds (imagenet.py:312-325), run on the GPU in fp32;
ours = B200 model / criterion / optimizer, float32 and bf16 activations, same init, same batches.

Trajectories of a deep net separate chaotically at the level of single steps, so the comparison is on
window means (20 steps), and the yardstick is MEASURED, never asserted on: the stock model is run three
times in float32 (NCHW, channels-last, NCHW with deterministic cuDNN algorithms - same arithmetic, different
summation orders); their per-window mean is the reference curve and their per-window max-min spread says how
far apart two correct implementations end up on this stream.  Our curves must stay within
max(3 % of the reference, 2 x spread) in every window.

The stream is tame on purpose (lr 0.005, batch 64: the loss falls 6.3 -> 4.0 without a cliff).  Three stock
float32 runs of it that differ only in summation order (measured on CPU: 8 threads, 3 threads, channels-last)
stay within 0.8 % of each other over the first 160 steps and 1.7 % at the end, so the 3 % band is the binding
one.  With the previous stream (lr 0.02) stock-vs-stock differed by up to 24 % in the last windows - a stream
like that measures chaos, not the implementation.

File name: statistical test, collected after every deterministic GPU test (and before test_zz_no_alloc.py).
"""
import json
import os

import numpy as np
import pytest
import torch
import torchvision

import imagent_distributed_training_pytorch_with_slurm_b200 as b200

pytestmark = pytest.mark.gpu
STEPS, BATCH, WINDOW, LR = 200, 64, 20, 0.005


def _stream():
    """100 class prototypes buried in noise (signal 0.4), 64x64."""
    g = torch.Generator().manual_seed(7)
    protos = torch.randn(100, 3, 64, 64, generator=g)
    out = []
    for _ in range(STEPS):
        t = torch.randint(0, 100, (BATCH,), generator=g)
        x = protos[t] * 0.4 + torch.randn(BATCH, 3, 64, 64, generator=g)
        out.append((x, t))
    return out


def _curve(model, crit, opt, batches):
    model.train()
    losses = []
    for x, t in batches:
        logits = model(x.cuda())
        loss = crit(logits, t.cuda())
        opt.zero_grad()
        loss.backward()
        opt.step()
        losses.append(loss.detach())
    return torch.stack(losses).float().cpu().numpy()


def _stock(batches, channels_last=False, deterministic=False):
    saved = torch.backends.cudnn.deterministic, torch.backends.cudnn.benchmark
    torch.backends.cudnn.deterministic, torch.backends.cudnn.benchmark = deterministic, False
    try:
        torch.manual_seed(0)
        m = torchvision.models.resnet18(weights=None, num_classes=1000).cuda()
        if channels_last:
            m = m.to(memory_format=torch.channels_last)
            batches = [(x.contiguous(memory_format=torch.channels_last), t) for x, t in batches]
        return _curve(m, torch.nn.CrossEntropyLoss().cuda(),
                      torch.optim.SGD(m.parameters(), LR, momentum=0.9, weight_decay=1e-4), batches)
    finally:
        torch.backends.cudnn.deterministic, torch.backends.cudnn.benchmark = saved


def test_loss_curves_match_fp32_eager_within_tolerance():
    a, b = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False
    try:
        batches = _stream()
        curves = {"eager_fp32": _stock(batches),
                  "eager_fp32_channels_last": _stock(batches, channels_last=True),
                  "eager_fp32_deterministic": _stock(batches, deterministic=True)}
        for name, dtype in (("b200_fp32", torch.float32), ("b200_bf16", torch.bfloat16)):
            torch.manual_seed(0)
            m = b200.build_model("resnet18", num_classes=1000, dtype=dtype).cuda()
            curves[name] = _curve(m, b200.CrossEntropyLoss().cuda(),
                                  b200.FlatSGD(m.parameters(), LR, momentum=0.9, weight_decay=1e-4), batches)
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = a, b
    win = {k: v.reshape(-1, WINDOW).mean(1) for k, v in curves.items()}
    stock = np.stack([win[k] for k in win if k.startswith("eager")])
    ref, spread = stock.mean(0), stock.max(0) - stock.min(0)      # the yardstick: measured, not asserted on
    tol = np.maximum(0.03 * ref, 2.0 * spread)
    d32, d16 = np.abs(win["b200_fp32"] - ref), np.abs(win["b200_bf16"] - ref)
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    if os.path.isdir(out):
        rec = {k: [float(x) for x in v] for k, v in win.items()}
        rec.update(reference=[float(x) for x in ref], stock_spread=[float(x) for x in spread],
                   tolerance=[float(x) for x in tol], dev_b200_fp32=[float(x) for x in d32],
                   dev_b200_bf16=[float(x) for x in d16], lr=LR, batch=BATCH, window=WINDOW)
        json.dump(rec, open(os.path.join(out, "loss_curve_windows.json"), "w"))
    assert ref[-1] < 0.75 * ref[0]                                # the stream is learnable and was being learned
    first = curves["eager_fp32"][0]
    assert abs(curves["b200_fp32"][0] - first) <= 1e-4 * first   # identical start (same init, same batch)
    assert (d32 <= tol).all(), (win["b200_fp32"], ref, tol)
    assert (d16 <= tol).all(), (win["b200_bf16"], ref, tol)
