"""GPU probe: a real-data training epoch end to end - generated JPEG folder -> sampler -> decode -> ragged input kernel
-> ResNet-50 bf16 step (CUDA-graph replay) - for each decoder.  python tools/probe_folder_train.py [n_images]"""
import json
import os
import sys
import tempfile
import time

import numpy as np
import torch
from PIL import Image

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import imagent_distributed_training_pytorch_with_slurm_b200 as b200  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 2560
SIZES = [(375, 500), (375, 500), (333, 500), (500, 375), (480, 640), (256, 256)]
root = tempfile.mkdtemp(prefix="b200_folder_")
rng = np.random.default_rng(0)
base = {}
for h, w in set(SIZES):
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    base[(h, w)] = np.stack([127 + 80 * np.sin(0.03 * xx + c) * np.cos(0.02 * yy - c) for c in range(3)], -1)
for i in range(N):
    h, w = SIZES[i % len(SIZES)]
    img = np.clip(base[(h, w)] + rng.normal(0, 12, (h, w, 3)), 0, 255).astype(np.uint8)
    d = os.path.join(root, "train", f"c{i % 8}")
    os.makedirs(d, exist_ok=True)
    Image.fromarray(img).save(os.path.join(d, f"{i:05d}.JPEG"), quality=90)
torch.backends.cudnn.benchmark = True
idx = b200.ImageFolderIndex(root, "train")
for decoder in ("nvjpeg", "pil"):
    torch.manual_seed(0)
    model = b200.FlatDDP(b200.build_model("resnet50", num_classes=len(idx.classes)).cuda())
    crit = b200.CrossEntropyLoss().cuda()
    opt = b200.FlatSGD(model.parameters(), 0.01, momentum=0.9, weight_decay=1e-4)
    sampler = b200.DistributedSampler(idx, num_replicas=1, rank=0)
    loader = b200.ImageFolderLoader(idx, sampler, 256, out_size=224, random_crop=True, decoder=decoder, workers=8, prefetch=3)
    runner = b200.GraphedStep(model, crit, opt)
    for epoch in range(3):
        sampler.set_epoch(epoch)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        r = b200.train(loader, model, crit, opt, epoch, 1, graph=runner)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    print(json.dumps({"decoder": decoder, "images": N, "epoch_s": round(dt, 3), "images_per_s": round(N / dt),
                      "train_loss": round(r[0], 4), "reported_batch_time_sum_s": round(r[3], 3)}), flush=True)
