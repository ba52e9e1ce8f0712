"""Oracle (test infrastructure) for the input stage.

Reference mode (imagenet.py:280-283): ``Resize((448,448))`` on a PIL image, ``ToTensor()``,
``Normalize((.5,.5,.5),(.5,.5,.5))``.  North-star mode adds torchvision's RandomResizedCrop box and
RandomHorizontalFlip in front (``torchvision/transforms/transforms.py:928-981,712-722``); on PIL
images ``resized_crop`` is ``crop`` followed by the same PIL resize, and the flip is applied to the
resized image.

The resize is Pillow's ``ImagingResample`` for 8-bit images with the BILINEAR filter (Pillow 12.2.0,
``src/libImaging/Resample.c``; Pillow is a third-party dependency of the reference and is not
vendored there).  Published algorithm, restated:

* per output index ``xx``: ``center = in0 + (xx+0.5)*scale``, ``scale = (in1-in0)/out``,
  ``filterscale = max(scale, 1)``, ``support = 1.0*filterscale``,
  ``xmin = max(0, int(center-support+0.5))``, ``xmax = min(in_size, int(center+support+0.5)) - xmin``,
  weights ``w[x] = tri((x+xmin-center+0.5)/filterscale)`` normalised to sum 1 (all in double);
* coefficients go to fixed point with 22 fractional bits: ``int(0.5 + w*2**22)`` (``-0.5`` for
  negative weights; bilinear has none);
* horizontal pass first, each output ``clip8((2**21 + sum(pix*k)) >> 22)``, producing an 8-bit
  intermediate; the vertical pass does the same on the intermediate.  A pass whose dimension is
  unchanged is skipped (it would be the identity anyway: the only non-zero coefficient is 2**22).
"""
import math

import numpy as np

PRECISION_BITS = 32 - 8 - 2


def _tri(x):
    x = np.abs(x)
    return np.where(x < 1.0, 1.0 - x, 0.0)


def pil_bilinear_coeffs(in_size, out_size, in0=0.0, in1=None):
    """Return (xmin[out], xcount[out], kk[out, ksize] int32) as Pillow's precompute_coeffs +
    normalize_coeffs_8bpc."""
    if in1 is None:
        in1 = float(in_size)
    scale = (in1 - in0) / out_size
    filterscale = scale if scale >= 1.0 else 1.0
    support = 1.0 * filterscale
    ksize = int(math.ceil(support)) * 2 + 1
    xmin = np.zeros(out_size, dtype=np.int32)
    xcnt = np.zeros(out_size, dtype=np.int32)
    kk = np.zeros((out_size, ksize), dtype=np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = in0 + (xx + 0.5) * scale
        lo = int(center - support + 0.5)
        if lo < 0:
            lo = 0
        hi = int(center + support + 0.5)
        if hi > in_size:
            hi = in_size
        n = hi - lo
        w = _tri((np.arange(n, dtype=np.float64) + lo - center + 0.5) * ss)
        ww = 0.0
        for v in w:  # sequential double sum, as the C loop
            ww += float(v)
        if ww != 0.0:
            w = w / ww
        xmin[xx] = lo
        xcnt[xx] = n
        for x in range(n):
            v = float(w[x])
            kk[xx, x] = int(-0.5 + v * (1 << PRECISION_BITS)) if v < 0 else int(0.5 + v * (1 << PRECISION_BITS))
    return xmin, xcnt, kk


def _clip8(acc):
    return np.clip(acc >> PRECISION_BITS, 0, 255).astype(np.uint8)


def pil_resize_bilinear_u8(img, out_h, out_w):
    """img: uint8 [H, W, C].  Returns uint8 [out_h, out_w, C], bit-exact with
    ``PIL.Image.fromarray(img).resize((out_w, out_h), BILINEAR)``."""
    img = np.asarray(img, dtype=np.uint8)
    h, w, c = img.shape
    cur = img
    if out_w != w:
        xmin, xcnt, kk = pil_bilinear_coeffs(w, out_w)
        tmp = np.empty((h, out_w, c), dtype=np.uint8)
        for xx in range(out_w):
            n = xcnt[xx]
            seg = cur[:, xmin[xx]:xmin[xx] + n, :].astype(np.int64)
            acc = (1 << (PRECISION_BITS - 1)) + (seg * kk[xx, :n].astype(np.int64)[None, :, None]).sum(axis=1)
            tmp[:, xx, :] = _clip8(acc)
        cur = tmp
    if out_h != h:
        ymin, ycnt, kk = pil_bilinear_coeffs(h, out_h)
        tmp = np.empty((out_h, cur.shape[1], c), dtype=np.uint8)
        for yy in range(out_h):
            n = ycnt[yy]
            seg = cur[ymin[yy]:ymin[yy] + n, :, :].astype(np.int64)
            acc = (1 << (PRECISION_BITS - 1)) + (seg * kk[yy, :n].astype(np.int64)[:, None, None]).sum(axis=0)
            tmp[yy] = _clip8(acc)
        cur = tmp
    return cur


def normalize_lut(mean, std):
    """float32 table [3, 256]: ``(v/255 - mean)/std`` evaluated in float32 exactly as
    ``ToTensor`` (``img.to(float32).div(255)``, transforms/functional.py:172-176) followed by
    ``Normalize`` (``tensor.sub_(mean).div_(std)``, functional.py:924-928)."""
    v = np.arange(256, dtype=np.float32) / np.float32(255.0)
    out = np.empty((3, 256), dtype=np.float32)
    for ch in range(3):
        out[ch] = (v - np.float32(mean[ch])) / np.float32(std[ch])
    return out


def input_stage_oracle(images_u8, boxes, flips, out_h, out_w, mean, std):
    """images_u8: uint8 [N, Hs, Ws, 3]; boxes: int [N,4] rows (top, left, height, width);
    flips: bool [N].  Returns float32 [N, out_h, out_w, 3] (NHWC)."""
    images_u8 = np.asarray(images_u8)
    n = images_u8.shape[0]
    lut = normalize_lut(mean, std)
    out = np.empty((n, out_h, out_w, 3), dtype=np.float32)
    for s in range(n):
        top, left, bh, bw = (int(v) for v in boxes[s])
        crop = images_u8[s, top:top + bh, left:left + bw, :]
        r = pil_resize_bilinear_u8(crop, out_h, out_w)
        if flips[s]:
            r = r[:, ::-1, :]
        for ch in range(3):
            out[s, :, :, ch] = lut[ch][r[:, :, ch]]
    return out


def rrc_params_oracle(height, width, scale=(0.08, 1.0), ratio=(3.0 / 4.0, 4.0 / 3.0), generator=None):
    """Crop box (top, left, h, w) drawn as torchvision ``RandomResizedCrop.get_params``
    (transforms.py:928-973): up to 10 tries of ``area*U(scale)``, ``exp(U(log ratio))``,
    ``w=round(sqrt(a*r))``, ``h=round(sqrt(a/r))``, then two ``randint`` for the corner;
    centre-crop fallback.  Draw order on the torch CPU generator is the pinned part."""
    import torch

    area = height * width
    log_ratio = torch.log(torch.tensor(ratio))
    for _ in range(10):
        target_area = area * torch.empty(1).uniform_(scale[0], scale[1], generator=generator).item()
        aspect = torch.exp(torch.empty(1).uniform_(log_ratio[0], log_ratio[1], generator=generator)).item()
        w = int(round(math.sqrt(target_area * aspect)))
        h = int(round(math.sqrt(target_area / aspect)))
        if 0 < w <= width and 0 < h <= height:
            i = torch.randint(0, height - h + 1, size=(1,), generator=generator).item()
            j = torch.randint(0, width - w + 1, size=(1,), generator=generator).item()
            return i, j, h, w
    in_ratio = float(width) / float(height)
    if in_ratio < min(ratio):
        w = width
        h = int(round(w / min(ratio)))
    elif in_ratio > max(ratio):
        h = height
        w = int(round(h * max(ratio)))
    else:
        w, h = width, height
    return (height - h) // 2, (width - w) // 2, h, w


def flip_draw_oracle(p=0.5, generator=None):
    """RandomHorizontalFlip (transforms.py:712-722): flip iff ``torch.rand(1) < p``."""
    import torch

    return bool(torch.rand(1, generator=generator) < p)


def space_to_depth_stem_oracle(x_nhwc):
    """[N,H,W,3] -> [N,(H+6)/2,(W+6)/2,16]: zero-pad by 3 (the stem convolution's padding, torchvision
    resnet.py:197), then 2x2 space-to-depth with cell channel (py%2*2 + px%2)*3 + ch; 4 spare channels
    are zero.  Pure re-indexing of ``input_stage_oracle``'s output."""
    x = np.asarray(x_nhwc)
    n, h, w, c = x.shape
    p = np.zeros((n, h + 6, w + 6, c), dtype=x.dtype)
    p[:, 3:3 + h, 3:3 + w, :] = x
    p = p.reshape(n, (h + 6) // 2, 2, (w + 6) // 2, 2, c).transpose(0, 1, 3, 2, 4, 5)  # N, Y, X, dy, dx, ch
    out = np.zeros((n, (h + 6) // 2, (w + 6) // 2, 16), dtype=x.dtype)
    out[..., :12] = p.reshape(n, (h + 6) // 2, (w + 6) // 2, 12)
    return out
