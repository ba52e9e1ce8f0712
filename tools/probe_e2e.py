"""GPU probe: where does the end-to-end loop lose time against the resident-input measurement?"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import imagent_distributed_training_pytorch_with_slurm_b200 as b200  # noqa: E402

torch.backends.cudnn.benchmark = True
torch.manual_seed(0)
model = b200.FlatDDP(b200.build_model("resnet50").cuda())
crit = b200.CrossEntropyLoss().cuda()
opt = b200.FlatSGD(model.parameters(), 0.1, momentum=0.9, weight_decay=1e-4)
runner = b200.GraphedStep(model, crit, opt)
b200.train(b200.SyntheticImageNet(256, 6, 256, 224, host=True), model, crit, opt, 0, 1, sync_every=1, lag=1, graph=runner)
for name, kw, tkw in [("host + per-step readback (bench e2e)", dict(host=True), dict(sync_every=1, lag=1)),
                      ("host, readback once per epoch", dict(host=True), dict()),
                      ("resident images + per-step readback", dict(host=False), dict(sync_every=1, lag=1)),
                      ("resident images, fixed boxes, epoch readback", dict(host=False, fixed_params=True), dict())]:
    loader = b200.SyntheticImageNet(256, 30, 256, 224, **kw)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    b200.train(loader, model, crit, opt, 0, 1, graph=runner, **tkw)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"{name:48s} {256 * 30 / dt:8.0f} img/s  {1e3 * dt / 30:6.2f} ms/step", flush=True)
