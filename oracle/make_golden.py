"""Generate ``tests/golden/*`` (run in the build container, where /root/reference exists):

    python -m oracle.make_golden

Sources of truth, in order of authority:
  * the reference's own functions, imported unmodified through ``oracle/ref_shim.py``
    (accuracy, AverageMeter, adjust_learning_rate);
  * the third-party libraries the reference calls (torch DistributedSampler, torchvision
    transforms on PIL images, stock DDP/SGD/CrossEntropy on CPU/gloo) at the versions installed here
    (recorded in the fixture);
  * the reference's published log ``imagent_sgd.out`` for the LR float values.

The fixtures are small (a few hundred KB) and are what the GPU box, which has no /root/reference,
checks the oracle and the CUDA path against.
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden")


def main():
    sys.path.insert(0, ROOT)
    from oracle.ref_shim import load_reference
    from oracle import train_cpu

    ref = load_reference()
    if ref is None:
        raise SystemExit("reference not on disk; goldens can only be generated in the build container")
    import PIL
    import torchvision
    from PIL import Image
    import torchvision.transforms as T
    import torchvision.transforms.functional as TF
    from torch.utils.data.distributed import DistributedSampler

    os.makedirs(OUT, exist_ok=True)
    g = {"versions": {"torch": torch.__version__, "torchvision": torchvision.__version__,
                      "pillow": PIL.__version__, "numpy": np.__version__}}

    # --- S1: sampler (imagenet.py:346-347,375) -----------------------------------------------
    class _Len:
        def __init__(self, n):
            self.n = n

        def __len__(self):
            return self.n

    samp = []
    for (n, w, r, e) in [(1281167, 1, 0, 0), (1281167, 1, 0, 1), (1281167, 2, 1, 0), (1281167, 2, 0, 0),
                         (1281167, 4, 3, 0), (1281167, 8, 7, 0), (1281167, 8, 7, 1), (1281167, 16, 15, 0),
                         (50000, 16, 0, 0), (10, 4, 3, 0), (3, 8, 5, 2)]:
        s = DistributedSampler(_Len(n), num_replicas=w, rank=r)
        s.set_epoch(e)
        idx = list(iter(s))
        samp.append({"len": n, "world": w, "rank": r, "epoch": e, "count": len(idx),
                     "first8": idx[:8], "last4": idx[-4:],
                     "sum": int(np.sum(np.asarray(idx, dtype=np.int64)))})
    g["sampler"] = samp

    # --- I2: crop boxes + flip bits (transforms.py:928-981,712-722) --------------------------
    crops = []
    for (seed, h, w, count) in [(0, 500, 375, 8), (1, 256, 256, 8), (7, 32, 400, 6), (3, 375, 500, 6)]:
        torch.manual_seed(seed)
        img = Image.fromarray(np.zeros((h, w, 3), dtype=np.uint8))
        seq = []
        for _ in range(count):
            i, j, hh, ww = T.RandomResizedCrop.get_params(img, (0.08, 1.0), (3.0 / 4.0, 4.0 / 3.0))
            fl = bool(torch.rand(1) < 0.5)
            seq.append([i, j, hh, ww, int(fl)])
        crops.append({"seed": seed, "height": h, "width": w, "draws": seq})
    g["crops"] = crops

    # --- E1: LR schedule (imagenet.py:154-162; imagent_sgd.out:274,454,634,814) --------------
    class _Opt:
        param_groups = [{"lr": None}]

    g["lr"] = {str(e): ref.adjust_learning_rate(0.1, _Opt(), e) for e in (0, 1, 29, 30, 59, 60, 89, 90, 99)}
    g["lr_repr_from_log"] = {"30": "0.010000000000000002", "60": "0.0010000000000000002",
                             "90": "0.00010000000000000003"}

    # --- A1 / R2: accuracy() and AverageMeter from the reference itself ----------------------
    gen = torch.Generator().manual_seed(5)
    logits = torch.randn(64, 1000, generator=gen)
    target = torch.randint(0, 1000, (64,), generator=gen)
    for r in range(0, 64, 3):  # plant winners so top-1/top-5 are not all zero
        logits[r, target[r]] += 6.0
    for r in range(1, 64, 5):
        logits[r, target[r]] += 2.5
    p1, p5 = ref.accuracy(logits, target, topk=(1, 5))
    loss = torch.nn.CrossEntropyLoss()(logits, target)
    np.savez_compressed(os.path.join(OUT, "ce_topk_case.npz"), logits=logits.numpy(), target=target.numpy(),
                        loss=np.float32(loss.item()), prec1=np.float32(p1.item()), prec5=np.float32(p5.item()))
    ks = (1, 2, 3, 5, 10, 50, 1000)                     # accuracy() accepts any tuple of k (imagenet.py:63-79)
    g["accuracy_topk"] = {"case": "ce_topk_case.npz", "topk": list(ks),
                          "prec": [float(v.item()) for v in ref.accuracy(logits, target, topk=ks)]}
    m = ref.AverageMeter()
    vals = [5.25, 4.5, 4.0, 3.125]
    for v in vals:
        m.update(v, 3)
    g["meter"] = {"vals": vals, "n": 3, "avg": m.avg, "sum": m.sum, "count": m.count}

    # --- I1/I2: the input stage through torchvision on PIL images ----------------------------
    rng = np.random.default_rng(11)
    imgs = rng.integers(0, 256, (3, 96, 80, 3), dtype=np.uint8)
    # reference mode: Resize((56,56)) stands in for (448,448); same code path, small fixture
    tr = T.Compose([T.Resize(size=(56, 56)), T.ToTensor(), T.Normalize((0.5, 0.5, 0.5), (0.5, 0.5, 0.5))])
    ref_mode = np.stack([tr(Image.fromarray(im)).permute(1, 2, 0).numpy() for im in imgs])
    torch.manual_seed(2)
    boxes, flips, rrc = [], [], []
    mean, std = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)
    for im in imgs:
        pil = Image.fromarray(im)
        i, j, hh, ww = T.RandomResizedCrop.get_params(pil, (0.08, 1.0), (3.0 / 4.0, 4.0 / 3.0))
        fl = bool(torch.rand(1) < 0.5)
        x = TF.resized_crop(pil, i, j, hh, ww, (48, 48))
        if fl:
            x = TF.hflip(x)
        rrc.append(TF.normalize(TF.to_tensor(x), mean, std).permute(1, 2, 0).numpy())
        boxes.append([i, j, hh, ww])
        flips.append(fl)
    np.savez_compressed(os.path.join(OUT, "input_case.npz"), images=imgs, ref_mode_out=ref_mode,
                        boxes=np.asarray(boxes, dtype=np.int32), flips=np.asarray(flips),
                        rrc_out=np.stack(rrc), rrc_mean=np.float32(mean), rrc_std=np.float32(std))

    # --- G3: one SGD trajectory from torch.optim.SGD -----------------------------------------
    gen = torch.Generator().manual_seed(9)
    p = torch.randn(4099, generator=gen).requires_grad_()
    opt = torch.optim.SGD([p], 0.1, momentum=0.9, weight_decay=1e-4, foreach=False)
    p0 = p.detach().clone().numpy()
    grads, ps = [], []
    for s in range(3):
        gr = torch.randn(4099, generator=gen)
        p.grad = gr.clone()
        opt.step()
        grads.append(gr.numpy())
        ps.append(p.detach().clone().numpy())
    np.savez_compressed(os.path.join(OUT, "sgd_case.npz"), p0=p0, grads=np.stack(grads), ps=np.stack(ps))

    # --- B1-B3: BatchNorm(+add)+ReLU from torch on CPU -----------------------------------------
    gen = torch.Generator().manual_seed(13)
    x = (torch.randn(6, 16, 5, 7, generator=gen) * 1.7 + 0.4)
    z = torch.randn(6, 16, 5, 7, generator=gen)
    dy = torch.randn(6, 16, 5, 7, generator=gen)
    bn = torch.nn.BatchNorm2d(16)
    with torch.no_grad():
        bn.weight.copy_(torch.rand(16, generator=gen) + 0.5)
        bn.bias.copy_(torch.randn(16, generator=gen) * 0.3)
    case = {"x": x.numpy(), "z": z.numpy(), "dy": dy.numpy(), "gamma": bn.weight.detach().numpy().copy(),
            "beta": bn.bias.detach().numpy().copy()}
    for name in ("none", "relu", "add_relu"):
        bn.running_mean.zero_()
        bn.running_var.fill_(1.0)
        bn.num_batches_tracked.zero_()
        xx = x.clone().requires_grad_()
        zz = z.clone().requires_grad_()
        bn.zero_grad()
        out = bn(xx)
        if name == "add_relu":
            out = out + zz
        if name != "none":
            out = torch.relu(out)
        out.backward(dy)
        case[f"{name}_y"] = out.detach().numpy()
        case[f"{name}_dx"] = xx.grad.numpy()
        case[f"{name}_dgamma"] = bn.weight.grad.numpy().copy()
        case[f"{name}_dbeta"] = bn.bias.grad.numpy().copy()
        if name == "add_relu":
            case[f"{name}_dz"] = zz.grad.numpy()
    case["running_mean"] = bn.running_mean.numpy().copy()
    case["running_var"] = bn.running_var.numpy().copy()
    bn.eval()
    case["eval_relu_y"] = torch.relu(bn(x)).detach().numpy()
    np.savez_compressed(os.path.join(OUT, "bn_case.npz"), **case)

    # --- G2/M1: stock DDP facts for resnet18 + a 2-rank gloo trajectory ------------------------
    nb = 3
    per_rank = [train_cpu.synthetic_batches(r, nb, 4, 64, 1000) for r in range(2)]
    res = train_cpu.run_ddp_oracle("resnet18", 2, per_rank, lr=0.02, seed=0, port=29621, threads=4)
    sd = res["state_dict"]
    g["ddp_resnet18_w2"] = {
        "batch_per_rank": 4, "image": 64, "steps": nb, "lr": 0.02, "seed": 0, "data_seed": 1234,
        "trace": [list(map(float, t)) for t in res["trace"]],
        "epoch_avg": [float(v) for v in res["epoch"][:3]],
        "n_keys": len(sd), "first_key": next(iter(sd)), "last_key": list(sd)[-1],
        "nbt": int(sd["module.bn1.num_batches_tracked"]),
        "weight_l2": {k: float(sd[k].double().norm()) for k in
                      ("module.conv1.weight", "module.bn1.weight", "module.bn1.running_var",
                       "module.layer1.0.conv1.weight", "module.layer4.1.bn2.bias", "module.fc.weight",
                       "module.fc.bias")},
    }
    single = train_cpu.run_ddp_oracle("resnet18", 1, [per_rank[0]], lr=0.02, seed=0, threads=8)
    g["resnet18_w1"] = {"batch": 4, "image": 64, "steps": nb, "lr": 0.02,
                        "trace": [list(map(float, t)) for t in single["trace"]]}

    with open(os.path.join(OUT, "golden.json"), "w") as f:
        json.dump(g, f, indent=1, sort_keys=True)
    print("wrote", OUT, sorted(os.listdir(OUT)))


if __name__ == "__main__":
    main()
