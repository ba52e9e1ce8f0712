"""Host-side logic of the package, run on CPU with the oracle standing in for the CUDA kernels.

What this covers: flat arenas and channels-last weight views, the hand-written backward of the whole
ResNet (tape order, residual fan-out, FC / pooling gradients), FlatSGD plumbing, the deferred-sync
train loop and state_dict compatibility - all against stock torchvision + autograd + torch.optim.SGD.
The CUDA kernels themselves are compared with the oracle in the -m gpu tests.
"""

import pytest
import torch
import torchvision

import imagent_distributed_training_pytorch_with_slurm_b200 as b200


def _stock(arch, seed, num_classes=1000):
    torch.manual_seed(seed)
    return getattr(torchvision.models, arch)(weights=None, num_classes=num_classes)


def test_bottleneck_backward_is_as_accurate_as_stock_float32(cpu_kernels):
    """ResNet-50 at random init and batch 8 is badly conditioned (stock float32 autograd is itself
    ~2e-2 off a float64 run), so the yardstick is float64: our gradients must be no further from it
    than stock float32 autograd is."""
    torch.manual_seed(0)
    model = b200.build_model("resnet50", device="cpu", dtype=torch.float32)
    ref32, ref64 = _stock("resnet50", 0), _stock("resnet50", 0).double()
    g = torch.Generator().manual_seed(3)
    x = torch.randn(8, 3, 64, 64, generator=g)
    t = torch.randint(0, 1000, (8,), generator=g)
    model.train(); ref32.train(); ref64.train()
    b200.CrossEntropyLoss()(model(x), t).backward()
    torch.nn.CrossEntropyLoss()(ref32(x), t).backward()
    torch.nn.CrossEntropyLoss()(ref64(x.double()), t).backward()
    p32, p64 = dict(ref32.named_parameters()), dict(ref64.named_parameters())
    for n, p in model.named_parameters():
        r = p64[n].grad
        ours = ((p.grad.double() - r).norm() / r.norm()).item()
        stock = ((p32[n].grad.double() - r).norm() / r.norm()).item()
        assert ours <= 2.0 * stock + 1e-4, (n, ours, stock)


@pytest.mark.parametrize("arch,size", [("resnet18", 64)])
def test_whole_step_matches_stock_autograd(cpu_kernels, arch, size):
    torch.manual_seed(0)
    model = b200.build_model(arch, num_classes=1000, device="cpu", dtype=torch.float32)
    ref = _stock(arch, 0)
    # same constructor RNG consumption -> identical initial weights, identical key order
    sd, rsd = model.state_dict(), ref.state_dict()
    assert list(sd.keys()) == list(rsd.keys())
    for k in sd:
        assert sd[k].shape == rsd[k].shape and sd[k].dtype == rsd[k].dtype
        assert torch.equal(sd[k].contiguous(), rsd[k]), k

    g = torch.Generator().manual_seed(3)
    x = torch.randn(4, 3, size, size, generator=g)
    t = torch.randint(0, 1000, (4,), generator=g)
    crit, rcrit = b200.CrossEntropyLoss(), torch.nn.CrossEntropyLoss()
    opt = b200.FlatSGD(model.parameters(), 0.05, momentum=0.9, weight_decay=1e-4)
    ropt = torch.optim.SGD(ref.parameters(), 0.05, momentum=0.9, weight_decay=1e-4)
    model.train(); ref.train()
    for step in range(2):
        logits = model(x)
        loss = crit(logits, t)
        opt.zero_grad()
        loss.backward()
        rlogits = ref(x)
        rloss = rcrit(rlogits, t)
        ropt.zero_grad()
        rloss.backward()
        assert torch.allclose(logits, rlogits, rtol=1e-4, atol=1e-4)
        assert abs(loss.item() - rloss.item()) < 1e-5 * max(1.0, abs(rloss.item()))
        rp = dict(ref.named_parameters())
        for n, p in model.named_parameters():
            gr, rg = p.grad, rp[n].grad
            denom = rg.norm().item() + 1e-12
            assert (gr - rg).norm().item() / denom < 5e-4, (step, n, (gr - rg).norm().item() / denom)
        p1, p5 = b200.accuracy(logits.data, t.data, topk=(1, 5))
        _, pred = rlogits.topk(5, 1, True, True)
        hit = pred.t().eq(t.view(1, -1))
        assert p1.item() == pytest.approx(hit[:1].float().sum().item() * 100.0 / 4)
        assert p5.item() == pytest.approx(hit[:5].float().sum().item() * 100.0 / 4)
        opt.step()
        ropt.step()
        for n, p in model.named_parameters():
            assert torch.allclose(p, rp[n], rtol=1e-3, atol=1e-5), (step, n)
    # buffers
    rb = dict(ref.named_buffers())
    for n, b in model.named_buffers():
        assert torch.allclose(b.float(), rb[n].float(), rtol=1e-3, atol=1e-5), n
    # eval forward
    model.eval(); ref.eval()
    with torch.no_grad():
        assert torch.allclose(model(x), ref(x), rtol=5e-3, atol=5e-3)


def test_state_dict_matches_reference_checkpoint_layout(cpu_kernels):
    import json, os
    g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "golden.json")))["ddp_resnet18_w2"]
    torch.manual_seed(0)
    model = b200.FlatDDP(b200.build_model("resnet18", device="cpu", dtype=torch.float32))
    sd = model.state_dict()
    assert len(sd) == g["n_keys"] == 122
    assert next(iter(sd)) == g["first_key"] == "module.conv1.weight"
    assert list(sd)[-1] == g["last_key"] == "module.fc.bias"
    assert sd["module.bn1.num_batches_tracked"].dtype == torch.int64
    assert all(v.is_contiguous() for v in sd.values())
    assert sd["module.conv1.weight"].dtype == torch.float32
    # loads into the stock torchvision module wrapped the way the reference wraps it
    ref = torch.nn.Sequential()
    ref.add_module("module", torchvision.models.resnet18(weights=None))
    ref.load_state_dict(sd)


def test_train_loop_meters_match_per_iteration_sync(cpu_kernels):
    torch.manual_seed(0)
    model = b200.build_model("resnet18", device="cpu", dtype=torch.float32)
    g = torch.Generator().manual_seed(1)
    batches = [(torch.randn(2, 3, 32, 32, generator=g), torch.randint(0, 1000, (2,), generator=g)) for _ in range(3)]
    torch.manual_seed(0)
    m2 = b200.build_model("resnet18", device="cpu", dtype=torch.float32)

    def run(m, sync_every, monkey_cuda):
        crit = b200.CrossEntropyLoss()
        opt = b200.FlatSGD(m.parameters(), 0.01, momentum=0.9, weight_decay=1e-4)
        return b200.train(batches, m, crit, opt, 0, 1, sync_every=sync_every)

    import imagent_distributed_training_pytorch_with_slurm_b200.engine as eng
    orig_sync, orig_to = torch.cuda.synchronize, eng._to_device
    torch.cuda.synchronize = lambda *a, **k: None
    eng._to_device = lambda t: t
    try:
        a = run(model, None, None)
        b = run(m2, 1, None)
    finally:
        torch.cuda.synchronize, eng._to_device = orig_sync, orig_to
    assert a[:3] == b[:3]  # deferred drain == per-iteration sync, bit for bit
    assert a[3] > 0 and b[3] > 0


def test_kernels_refuse_cpu_tensors():
    model = b200.build_model("resnet18", device="cpu", dtype=torch.float32)
    with pytest.raises(b200.B200Error):
        model(torch.randn(1, 3, 32, 32))


def test_second_backward_without_zero_grad_or_step_raises(cpu_kernels):
    """The backward overwrites the gradient arena where autograd would accumulate (the reference clears the gradients
    before every backward, imagenet.py:127, and never accumulates): doing it twice must raise, not silently differ."""
    torch.manual_seed(0)
    model = b200.build_model("resnet18", device="cpu", dtype=torch.float32).train()
    opt = b200.FlatSGD(model.parameters(), 0.01, momentum=0.9, weight_decay=1e-4)
    g = torch.Generator().manual_seed(1)
    x, t = torch.randn(2, 3, 32, 32, generator=g), torch.randint(0, 1000, (2,), generator=g)
    crit = b200.CrossEntropyLoss()
    crit(model(x), t).backward()
    with pytest.raises(RuntimeError, match="does not accumulate"):
        crit(model(x), t).backward()
    opt.zero_grad()                      # re-arms
    crit(model(x), t).backward()
    opt.step()                           # so does step()
    crit(model(x), t).backward()
    model.zero_grad()                    # nn.Module.zero_grad must neither drop the permanent views nor break the guard
    assert all(p.grad is not None for p in model.parameters())
    crit(model(x), t).backward()


def test_load_state_dict_refreshes_the_weights_the_convolutions_read(cpu_kernels):
    """bf16 mode: convolutions read a bf16 shadow of the float32 arena; loading a checkpoint (bare or through the
    data-parallel wrapper, imagenet.py:392's file) must bring it up to date."""
    torch.manual_seed(0)
    src = b200.build_model("resnet18", device="cpu", dtype=torch.bfloat16)
    sd = b200.FlatDDP(src).state_dict()
    torch.manual_seed(1)
    dst = b200.FlatDDP(b200.build_model("resnet18", device="cpu", dtype=torch.bfloat16))
    assert not torch.equal(dst.module.arena.shadow, src.arena.shadow)
    dst.load_state_dict(sd)
    assert torch.equal(dst.module.arena.params, src.arena.params)
    assert torch.equal(dst.module.arena.shadow, src.arena.shadow)
    assert torch.equal(dst.module.conv1.w_compute.float(), src.conv1.weight.detach().to(torch.bfloat16).float())
    torch.manual_seed(2)
    bare = b200.build_model("resnet18", device="cpu", dtype=torch.bfloat16)
    bare.load_state_dict({k[len("module."):]: v for k, v in sd.items()})
    assert torch.equal(bare.arena.shadow, src.arena.shadow)


def test_meter_weight_is_the_reference_channel_count():
    """imagenet.py:142: ``batch_size = input[0].size(0)`` is the CHANNEL count (3) of an NCHW batch; the 16-channel
    stem layout must not change the weight the meters see."""
    import imagent_distributed_training_pytorch_with_slurm_b200.engine as eng

    x = torch.zeros(4, 3, 8, 8)
    assert eng._meter_weight(x) == 3
    s2d = torch.zeros(4, 16, 7, 7)
    s2d._b200_s2d_stem = (8, 8)
    assert eng._meter_weight(s2d) == 3
