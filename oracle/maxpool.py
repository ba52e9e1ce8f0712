"""Oracle (test infrastructure): 3x3 / stride 2 / pad 1 max-pooling forward and backward on NHWC arrays.

Site: torchvision ``resnet.py:200,270`` (``nn.MaxPool2d(kernel_size=3, stride=2, padding=1)``) and its
autograd gradient.  Semantics of ATen's ``max_pool2d_with_indices``: windows are clipped at the border
(padding never wins), the FIRST maximum in row-major window order is the argmax, and the backward adds
each output gradient to its argmax pixel.
"""
import numpy as np


def maxpool3x3s2_fwd_oracle(x):
    """x: [N,H,W,C].  Returns (y [N,OH,OW,C], argpos uint8 [N,OH,OW,C] with the window position 0..8)."""
    x = np.asarray(x, dtype=np.float32)
    n, h, w, c = x.shape
    oh, ow = (h - 1) // 2 + 1, (w - 1) // 2 + 1
    best = np.full((n, oh, ow, c), -np.inf, dtype=np.float32)
    pos = np.zeros((n, oh, ow, c), dtype=np.uint8)
    seen = np.zeros((n, oh, ow, 1), dtype=bool)
    for kh in range(3):
        for kw in range(3):
            for oy in range(oh):
                iy = oy * 2 - 1 + kh
                if iy < 0 or iy >= h:
                    continue
                ixs = np.arange(ow) * 2 - 1 + kw
                ok = (ixs >= 0) & (ixs < w)
                v = x[:, iy, ixs[ok], :]
                b = best[:, oy, ok, :]
                take = (~seen[:, oy, ok, :]) | (v > b)
                best[:, oy, ok, :] = np.where(take, v, b)
                pos[:, oy, ok, :] = np.where(take, kh * 3 + kw, pos[:, oy, ok, :])
                seen[:, oy, ok, :] = True
    return best, pos


def maxpool3x3s2_bwd_oracle(dy, pos, in_hw):
    dy = np.asarray(dy, dtype=np.float64)
    n, oh, ow, c = dy.shape
    h, w = in_hw
    dx = np.zeros((n, h, w, c), dtype=np.float64)
    for oy in range(oh):
        for ox in range(ow):
            p = pos[:, oy, ox, :].astype(np.int64)
            iy = oy * 2 - 1 + p // 3
            ix = ox * 2 - 1 + p % 3
            nn, cc = np.meshgrid(np.arange(n), np.arange(c), indexing="ij")
            np.add.at(dx, (nn, iy, ix, cc), dy[:, oy, ox, :])
    return dx.astype(np.float32)
