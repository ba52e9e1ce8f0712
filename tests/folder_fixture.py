"""A tiny dataset in the layout ``datasets.ImageNet`` reads after extraction (imagenet.py:287-296):
``<root>/{train,val}/<class>/<image>``, written with Pillow.  Images have different sizes and are smooth (low-frequency
colour fields + mild texture) so that JPEG decoders agree closely; a grayscale JPEG, a CMYK JPEG, a PNG, an upper-case
extension, a nested folder and a non-image file exercise the index and the decoder fall-backs."""
import os

import numpy as np
from PIL import Image

SIZES = [(64, 80), (75, 100), (96, 64), (81, 67), (120, 90), (70, 70), (100, 131)]


def _picture(rng, h, w):
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float64)
    img = np.empty((h, w, 3))
    for c in range(3):
        fx, fy, ph = rng.uniform(0.02, 0.12), rng.uniform(0.02, 0.12), rng.uniform(0, 6.28)
        img[:, :, c] = 127 + 90 * np.sin(fx * xx + ph) * np.cos(fy * yy - ph) + rng.normal(0, 4, (h, w))
    return np.clip(img, 0, 255).astype(np.uint8)


def make_dataset(root, classes=("n01", "n02", "n03"), per_class=5, seed=3):
    rng = np.random.default_rng(seed)
    k = 0
    for split, count in (("train", per_class), ("val", 2)):
        for ci, c in enumerate(classes):
            d = os.path.join(root, split, c)
            os.makedirs(d, exist_ok=True)
            for j in range(count):
                h, w = SIZES[k % len(SIZES)]
                k += 1
                im = Image.fromarray(_picture(rng, h, w))
                name, kw = f"img_{j:02d}.JPEG", dict(quality=92)
                if split == "train" and ci == 0 and j == 1:
                    im = im.convert("L")                                  # grayscale JPEG
                elif split == "train" and ci == 1 and j == 2:
                    im = im.convert("CMYK")                               # nvJPEG refuses these: Pillow fall-back
                elif split == "train" and ci == 2 and j == 0:
                    name, kw = f"img_{j:02d}.png", {}                     # not a JPEG at all
                elif split == "train" and ci == 2 and j == 3:
                    kw = dict(quality=95, subsampling=0)                  # 4:4:4 chroma
                elif split == "train" and ci == 0 and j == 4:
                    d2 = os.path.join(d, "more")                          # nested folder (os.walk order)
                    os.makedirs(d2, exist_ok=True)
                    im.save(os.path.join(d2, "deep.jpg"), quality=90)
                    continue
                im.save(os.path.join(d, name), **kw)
    with open(os.path.join(root, "train", classes[0], "notes.txt"), "w") as f:
        f.write("not an image\n")
    return root
