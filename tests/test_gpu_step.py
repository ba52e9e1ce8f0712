"""Whole training step on the GPU against the oracle (stock torch DDP/ResNet/CE/SGD on CPU) and the
committed golden trajectory.  -m gpu only.

Tolerances: float32 activation mode - loss within 2e-3 relative, per-tensor gradient relative L2
error <= 1e-2 (cuDNN vs CPU convolution order; ResNet-18 at random init amplifies ~1e-6 noise to
~1e-4).  bf16 mode - loss within 1e-2 relative at step 1, gradient errors held to what stock
``torch.autocast(bfloat16)`` reaches on the same batch (see the bf16 test for the measured reason).
"""
import json
import os

import numpy as np
import pytest
import torch

from oracle import train_cpu

import imagent_distributed_training_pytorch_with_slurm_b200 as b200

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(autouse=True)
def _strict_fp32():
    a, b = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = a, b


def _run_ours(arch, batches, dtype, lr, steps):
    torch.manual_seed(0)
    model = b200.build_model(arch, num_classes=1000, device="cuda", dtype=dtype).cuda()
    crit = b200.CrossEntropyLoss().cuda()
    opt = b200.FlatSGD(model.parameters(), lr, momentum=0.9, weight_decay=1e-4)
    model.train()
    trace, grads = [], None
    for s in range(steps):
        x, t = batches[s]
        x, t = x.cuda(), t.cuda()
        logits = model(x)
        loss = crit(logits, t)
        opt.zero_grad()
        loss.backward()
        if s == 0:
            grads = {n: p.grad.detach().float().cpu().clone() for n, p in model.named_parameters()}
        opt.step()
        p1, p5 = b200.accuracy(logits.data, t.data, topk=(1, 5))
        trace.append((loss.item(), p1.item(), p5.item()))
    return model, trace, grads


def _oracle_grads(arch, batch):
    torch.manual_seed(0)
    model, crit, _ = train_cpu.build_reference_objects(arch, 1000, 0.02, 0, 1)
    model.train()
    crit(model(batch[0]), batch[1]).backward()
    return {n: p.grad.clone() for n, p in model.named_parameters()}


def test_resnet18_fp32_step_matches_golden_and_oracle():
    g = json.load(open(os.path.join(GOLD, "golden.json")))["resnet18_w1"]
    batches = train_cpu.synthetic_batches(0, g["steps"], g["batch"], g["image"], 1000)
    model, trace, grads = _run_ours("resnet18", batches, torch.float32, g["lr"], g["steps"])
    for (lo, p1, p5), (glo, gp1, gp5) in zip(trace, g["trace"]):
        assert abs(lo - glo) <= 2e-3 * abs(glo), (trace, g["trace"])
        assert p1 == gp1 and p5 == gp5
    og = _oracle_grads("resnet18", batches[0])
    for n, gr in grads.items():
        rel = ((gr - og[n]).norm() / (og[n].norm() + 1e-12)).item()
        assert rel <= 1e-2, (n, rel)
    # checkpoint layout of the wrapped model == the reference's (imagenet.py:392)
    sd = b200.FlatDDP(model).state_dict()
    assert len(sd) == 122 and next(iter(sd)) == "module.conv1.weight"
    assert all(v.dtype in (torch.float32, torch.int64) for v in sd.values())
    assert int(sd["module.bn1.num_batches_tracked"]) == g["steps"]


def _stock_autocast_grads(arch, batch):
    """Stock torchvision model under torch.autocast(bf16), channels-last, on the GPU: the PyTorch-eager
    mixed-precision incumbent, used as the yardstick for what bf16 activations cost in accuracy."""
    model, crit, _ = train_cpu.build_reference_objects(arch, 1000, 0.02, 0, 1)
    model = model.cuda().to(memory_format=torch.channels_last).train()
    x = batch[0].cuda().contiguous(memory_format=torch.channels_last)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        loss = crit(model(x), batch[1].cuda())
    loss.backward()
    return loss.item(), {n: p.grad.float().cpu() for n, p in model.named_parameters()}


def test_resnet18_bf16_step_within_bf16_tolerance():
    """bf16 activations against the float32 oracle, batch 64.

    What a bf16 pipeline CAN reach was measured first (stock ``torch.autocast(bfloat16)`` against stock float32 on the
    same batch, ResNet-18 at random init, batches 64-128, 64x64 to 224x224, also after 40 training steps): the loss
    agrees to 5e-4, but per-tensor gradient relative L2 errors are 0.02 (fc) ... 0.17-0.27 (layer4) ... 0.39 (conv1),
    i.e. cosine 0.92-0.999 - rounding to 8 bits is amplified by the cancellation in the weight-gradient sums of a
    deep BatchNorm net.  SURVEY 8c's proposed "cosine >= 0.999 / rel-L2 <= 3e-2" is therefore out of reach for ANY
    bf16 implementation, stock autocast included, and the yardstick has to be stock autocast itself:
      * loss within 1e-2 relative of the oracle;
      * every gradient tensor >= 4096 elements: rel-L2 error <= 1.3 x stock autocast's on that tensor + 0.01
        (where stock reaches 3e-2 this demands <= 5e-2; the old form allowed 0.14 there);
      * all gradients concatenated: rel-L2 error <= 1.15 x stock autocast's.
    A backward that is subtly wrong (a mask off by one channel, a missing 1/N, a stale statistic) fails the second and
    third clause by a wide margin; the float32 step test above pins the same code path at 1e-2 per tensor."""
    batches = train_cpu.synthetic_batches(0, 1, 64, 64, 1000)
    _, trace, grads = _run_ours("resnet18", batches, torch.bfloat16, 0.02, 1)
    res = train_cpu.run_ddp_oracle("resnet18", 1, [batches], lr=0.02, seed=0)
    assert abs(trace[0][0] - res["trace"][0][0]) <= 1e-2 * abs(res["trace"][0][0])
    og = _oracle_grads("resnet18", batches[0])
    sloss, sg = _stock_autocast_grads("resnet18", batches[0])

    def rel(a, b):
        return ((a - b).norm() / (b.norm() + 1e-30)).item()

    rows = []
    for n, gr in grads.items():
        if gr.numel() < 4096:
            continue
        ours, stock = rel(gr, og[n]), rel(sg[n], og[n])
        rows.append((n, ours, stock))
    cat = lambda d: torch.cat([d[n].flatten().double() for n in og])  # noqa: E731
    g_ours, g_stock = rel(cat(grads), cat(og)), rel(cat(sg), cat(og))
    out = os.path.join(os.path.dirname(GOLD), "..", "gpurun_out")
    if os.path.isdir(out):
        json.dump({"loss": [trace[0][0], float(res["trace"][0][0]), sloss], "global": [g_ours, g_stock],
                   "per_tensor": rows}, open(os.path.join(out, "bf16_step_gradient_errors.json"), "w"))
    for n, ours, stock in rows:
        assert ours <= 1.3 * stock + 0.01, (n, ours, stock)
    assert g_ours <= 1.15 * g_stock, (g_ours, g_stock)


def test_resnet50_bf16_trains_and_loss_falls():
    """Headline architecture: a few bf16 steps on one fixed batch must drive the loss down, all
    values finite, and the eval-mode forward must agree with the train-mode statistics path."""
    batches = train_cpu.synthetic_batches(0, 1, 32, 64, 1000) * 8
    model, trace, _ = _run_ours("resnet50", batches, torch.bfloat16, 0.01, 8)
    losses = [t[0] for t in trace]
    assert all(np.isfinite(losses))
    assert losses[-1] < losses[0]
    model.eval()
    with torch.no_grad():
        out = model(batches[0][0].cuda())
    assert out.shape == (32, 1000) and torch.isfinite(out.float()).all()


def test_train_and_validate_loops_with_device_input_stage():
    torch.manual_seed(0)
    model = b200.FlatDDP(b200.build_model("resnet18", dtype=torch.bfloat16).cuda())
    crit = b200.CrossEntropyLoss().cuda()
    opt = b200.FlatSGD(model.parameters(), 0.05, momentum=0.9, weight_decay=1e-4)
    loader = b200.SyntheticImageNet(8, 4, src_size=96, out_size=64, host=True)
    lr = b200.adjust_learning_rate(0.05, opt, 31)
    assert opt.param_groups[0]["lr"] == lr == 0.05 * 0.1
    a = b200.train(loader, model, crit, opt, 0, 1)
    assert len(a) == 4 and np.isfinite(a[0]) and a[3] > 0 and 0 <= a[1] <= a[2] <= 100
    vloader = b200.SyntheticImageNet(8, 2, src_size=96, out_size=64, random_crop=False, seed=4321)
    v = b200.validate(vloader, model, crit, 1)
    assert np.isfinite(v[0]) and v[3] > 0
    # per-iteration sync gives the same meters as the deferred drain on the same stream of batches
    torch.manual_seed(0)
    m2 = b200.build_model("resnet18", dtype=torch.bfloat16).cuda()
    o2 = b200.FlatSGD(m2.parameters(), 0.05 * 0.1, momentum=0.9, weight_decay=1e-4)
    l2 = b200.SyntheticImageNet(8, 4, src_size=96, out_size=64, host=True)
    b = b200.train(l2, m2, crit, o2, 0, 1, sync_every=1)
    assert a[0] == pytest.approx(b[0], rel=2e-2) and a[1] == b[1]  # (bit-equality is checked on CPU; cuDNN
    # wgrad is not run-to-run deterministic)


def test_space_to_depth_stem_equals_plain_stem():
    """Same batch through the 7x7/2 stem and through the rearranged 4x4/1 stem (float32): logits,
    loss and the conv1 weight gradient agree to convolution rounding."""
    from imagent_distributed_training_pytorch_with_slurm_b200 import _cabi
    g = torch.Generator().manual_seed(0)
    n, src, out = 8, 80, 64
    images = torch.randint(0, 256, (n, src, src, 3), dtype=torch.uint8, generator=g).cuda()
    boxes, flips = b200.draw_batch_params(n, src, src, g)
    target = torch.randint(0, 1000, (n,), generator=g).cuda()
    mean, std = b200.data.IMAGENET_MEAN, b200.data.IMAGENET_STD
    res = []
    for layout in (_cabi.LAYOUT_NHWC, _cabi.LAYOUT_S2D_STEM):
        torch.manual_seed(0)
        model = b200.build_model("resnet18", dtype=torch.float32).cuda().train()
        x = _cabi.input_stage(images, boxes.cuda(), flips.cuda(), out, out, mean, std, torch.float32, layout)
        logits = model(x)
        loss = b200.CrossEntropyLoss()(logits, target)
        loss.backward()
        res.append((logits.detach().clone(), loss.item(), model.conv1.weight.grad.detach().clone(),
                    model.fc.weight.grad.detach().clone()))
    a, b = res
    assert torch.allclose(a[0], b[0], rtol=1e-3, atol=1e-3)
    assert abs(a[1] - b[1]) <= 1e-4 * abs(a[1])
    for i in (2, 3):
        assert ((a[i] - b[i]).norm() / a[i].norm()).item() < 5e-3


def test_reference_loop_drives_our_objects():
    """Drop-in check: the reference's train() body (restated in oracle/train_cpu.py with only the
    .cuda()/synchronize calls removed: forward, criterion, zero_grad, backward, step, topk-based
    accuracy(), three reduce_tensor, three .item(), meters with n=3) runs unmodified on the B200 model,
    criterion and optimizer, and logs what our own train() logs for the same batches."""
    batches = [(x.cuda(), t.cuda()) for x, t in train_cpu.synthetic_batches(0, 3, 8, 64, 1000)]

    def fresh():
        torch.manual_seed(0)
        m = b200.build_model("resnet18", dtype=torch.float32).cuda()
        return m, b200.CrossEntropyLoss().cuda(), b200.FlatSGD(m.parameters(), 0.02, momentum=0.9, weight_decay=1e-4)

    m1, c1, o1 = fresh()
    trace = []
    ref = train_cpu.train_epoch_cpu(batches, m1, c1, o1, 1, trace)       # the reference's loop
    m2, c2, o2 = fresh()
    ours = b200.train(batches, m2, c2, o2, 0, 1)                          # ours, deferred sync
    for a, b in zip(ref[:3], ours[:3]):
        assert a == pytest.approx(b, rel=1e-4, abs=1e-6)                  # (cuDNN wgrad is not bit-reproducible)
    assert len(trace) == 3 and all(np.isfinite(v[0]) for v in trace)
    # the stock criterion also works on our logits (autograd flows through the single-node network)
    m3, _, o3 = fresh()
    loss = torch.nn.CrossEntropyLoss()(m3(batches[0][0]), batches[0][1])
    loss.backward()
    o3.step()
    assert abs(loss.item() - trace[0][0]) <= 1e-4 * abs(trace[0][0])


def test_graphed_training_matches_eager_training():
    """train(..., graph=True): after the eager warm-up the iteration is one CUDA-graph replay; same
    batches must give the same meters as the eager loop (bf16 path is run-to-run deterministic here up to
    cuDNN's wgrad order, so compare to 1e-3)."""
    def run(graph):
        torch.manual_seed(0)
        m = b200.build_model("resnet18", dtype=torch.bfloat16).cuda()
        opt = b200.FlatSGD(m.parameters(), 0.01, momentum=0.9, weight_decay=1e-4)
        loader = b200.SyntheticImageNet(16, 10, src_size=96, out_size=64, host=True)
        res = b200.train(loader, m, b200.CrossEntropyLoss().cuda(), opt, 0, 1, sync_every=1, lag=1, graph=graph)
        m.eval()
        with torch.no_grad():  # the stem-weight cache must have followed the replays
            x = b200.SyntheticImageNet(4, 1, src_size=96, out_size=64, random_crop=False, seed=9)
            y = m(next(iter(x))[0]).float()
        return res, y, m.arena.steps

    (a, ya, sa), (b, yb, sb) = run(False), run(True)
    assert sa == sb == 10
    assert a[0] == pytest.approx(b[0], rel=1e-3) and a[1] == b[1] and a[2] == b[2]
    assert torch.allclose(ya, yb, rtol=5e-2, atol=5e-2)


def test_checkpoint_round_trip_into_a_bf16_model_evaluates_on_the_loaded_weights():
    """The reference saves ``model.state_dict()`` (imagenet.py:388-392).  Loading that file back into a fresh bf16
    model - bare or wrapped - must make the convolutions use the LOADED weights: they read a bf16 shadow arena and
    a rearranged stem weight, which load_state_dict does not write itself."""
    batches = train_cpu.synthetic_batches(0, 2, 8, 64, 1000)
    trained, _, _ = _run_ours("resnet18", batches, torch.bfloat16, 0.05, 2)
    sd = b200.FlatDDP(trained).state_dict()                     # 'module.'-prefixed float32, as the CLI writes it
    loader = b200.SyntheticImageNet(8, 1, src_size=96, out_size=64, random_crop=False, seed=11)
    x = next(iter(loader))[0]                                    # space-to-depth stem layout: exercises that cache too
    trained.eval()
    with torch.no_grad():
        want = trained(x).float()
    torch.manual_seed(123)                                       # a different random init
    fresh = b200.FlatDDP(b200.build_model("resnet18", dtype=torch.bfloat16).cuda())
    fresh.eval()
    with torch.no_grad():
        before = fresh(x).float()                                # also fills the stem-weight cache with the old weights
    fresh.load_state_dict(sd)
    with torch.no_grad():
        got = fresh(x).float()
    tol = 1e-2 * want.abs().max().item()   # same weights, same kernels; slack only for cuDNN picking another algorithm
    assert torch.allclose(got, want, rtol=0, atol=tol) and not torch.allclose(before, want, rtol=0, atol=10 * tol)
    # bare model, keys without the prefix
    bare = b200.build_model("resnet18", dtype=torch.bfloat16).cuda().eval()
    bare.load_state_dict({k[len("module."):]: v for k, v in sd.items()})
    with torch.no_grad():
        assert torch.allclose(bare(x).float(), want, rtol=0, atol=tol)
    assert torch.equal(bare.arena.shadow, bare.arena.params.to(torch.bfloat16))
