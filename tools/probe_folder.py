"""GPU probe: images/s of the on-disk front end alone (index -> sampler -> decode -> ragged input kernel) on a generated
dataset of ImageNet-like JPEGs (500x375 and a few other sizes, quality 90, 4:2:0), for both decoders.  The reference does
this stage with PIL in 10 DataLoader worker processes (imagenet.py:280-296,350-359).
    python tools/probe_folder.py [n_images] > gpurun_out/probe_folder.jsonl"""
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

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1536
SIZES = [(375, 500), (375, 500), (333, 500), (500, 375), (480, 640), (256, 256)]
root = tempfile.mkdtemp(prefix="b200_folder_")
rng = np.random.default_rng(0)
t0 = time.time()
base = {}
for h, w in set(SIZES):
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    base[(h, w)] = np.stack([127 + 80 * np.sin(0.03 * xx + c) * np.cos(0.02 * yy - c) for c in range(3)], -1)
for i in range(N):
    h, w = SIZES[i % len(SIZES)]
    img = np.clip(base[(h, w)] + rng.normal(0, 12, (h, w, 3)), 0, 255).astype(np.uint8)
    d = os.path.join(root, "train", f"c{i % 8}")
    os.makedirs(d, exist_ok=True)
    Image.fromarray(img).save(os.path.join(d, f"{i:05d}.JPEG", ), quality=90)
nbytes = sum(os.path.getsize(os.path.join(dp, f)) for dp, _, fs in os.walk(root) for f in fs)
print(json.dumps({"dataset": N, "jpeg_MB": round(nbytes / 1e6, 1), "made_in_s": round(time.time() - t0, 1),
                  "cores": len(os.sched_getaffinity(0))}), flush=True)
idx = b200.ImageFolderIndex(root, "train")
for decoder, workers in (("pil", 8), ("nvjpeg", 8), ("nvjpeg", 16), ("torchvision", 8)):
    for mode, out, crop in (("reference 448 whole image", 448, False), ("random crop 224", 224, True)):
        sampler = b200.DistributedSampler(idx, num_replicas=1, rank=0)
        loader = b200.ImageFolderLoader(idx, sampler, 256, out_size=out, random_crop=crop, decoder=decoder, workers=workers,
                                        prefetch=3)
        for epoch in range(2):                       # epoch 0 warms file cache, allocator and nvJPEG
            sampler.set_epoch(epoch)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            n = 0
            for x, t in loader:
                n += x.shape[0]
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
        print(json.dumps({"decoder": decoder, "workers": workers, "mode": mode, "images_per_s": round(n / dt),
                          "h2d_MB_per_batch": round(loader.h2d_bytes / 1e6, 1),
                          "hardware_engine_images_in_last_batch": getattr(loader, "last_hw_count", None),
                          "hardware_backend_created": getattr(loader, "jpeg_has_hardware", None)}), flush=True)
