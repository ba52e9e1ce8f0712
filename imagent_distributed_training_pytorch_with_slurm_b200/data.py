"""Input stage on the device, and synthetic ImageNet-shaped loaders.

Replaces the reference's CPU transform pipeline ``Resize((448,448)) -> ToTensor -> Normalize(.5,.5)``
run per image in 10 DataLoader workers (imagenet.py:280-283,350-354) plus the fp32 H2D copy
(imagenet.py:119) by: uint8 NHWC batch -> (pinned) H2D -> ONE fused kernel doing crop box, Pillow-exact
antialiased bilinear resize, flip, /255, (x-mean)/std and the cast to channels-last bf16.

Two modes:
  * reference mode  - box = whole image, no flip, 448x448, mean = std = 0.5 (what imagenet.py does);
  * north-star mode - torchvision RandomResizedCrop(224) + RandomHorizontalFlip boxes/flips, drawn on
    the host from a torch CPU generator in torchvision's order (transforms.py:928-981,712-722), and
    ImageNet mean/std.
"""
import math

import torch

from . import _cabi

IMAGENET_MEAN, IMAGENET_STD = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)
REFERENCE_MEAN, REFERENCE_STD = (0.5, 0.5, 0.5), (0.5, 0.5, 0.5)


_LOG_RATIO_CACHE = {}


def _log_ratio32(ratio):
    key = tuple(ratio)
    if key not in _LOG_RATIO_CACHE:  # float32 bounds, as torchvision's torch.log(torch.tensor(ratio))
        _LOG_RATIO_CACHE[key] = torch.log(torch.tensor(key, dtype=torch.float32)).tolist()
    return _LOG_RATIO_CACHE[key]


class _Draw:
    """Scalar draws from a torch CPU generator with as little Python/dispatcher overhead as possible:
    preallocated 1-element tensors filled in place and read through numpy views.  ``uniform_`` /
    ``random_`` on a 1-element tensor consume the generator exactly like the ``torch.empty(1).uniform_()``,
    ``torch.randint(..., size=(1,))`` and ``torch.rand(1)`` calls torchvision makes (one 32-bit word each)."""

    def __init__(self, generator):
        self.g = generator
        self.f = torch.empty(1, dtype=torch.float32)
        self.i = torch.empty(1, dtype=torch.int64)
        self.fv, self.iv = self.f.numpy(), self.i.numpy()

    def uniform(self, lo, hi):
        self.f.uniform_(lo, hi, generator=self.g)
        return float(self.fv[0])

    def exp_uniform(self, lo, hi):
        self.f.uniform_(lo, hi, generator=self.g).exp_()
        return float(self.fv[0])

    def randint(self, n):
        self.i.random_(0, n, generator=self.g)
        return int(self.iv[0])


def rrc_params(height, width, scale=(0.08, 1.0), ratio=(3.0 / 4.0, 4.0 / 3.0), generator=None, _draw=None):
    """(top, left, h, w) with the draw order and rounding of torchvision's
    ``RandomResizedCrop.get_params`` (transforms.py:928-973)."""
    d = _draw if _draw is not None else _Draw(generator)
    area = height * width
    lo32, hi32 = _log_ratio32(ratio)
    for _ in range(10):
        target_area = area * d.uniform(scale[0], scale[1])
        aspect = d.exp_uniform(lo32, hi32)
        w = int(round(math.sqrt(target_area * aspect)))
        h = int(round(math.sqrt(target_area / aspect)))
        if 0 < w <= width and 0 < h <= height:
            top = d.randint(height - h + 1)
            left = d.randint(width - w + 1)
            return top, left, h, w
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


def draw_batch_params(n, height, width, generator=None, random_crop=True, flip_p=0.5):
    """Per-sample (box, flip) for a batch; per sample the draws are get_params then
    ``torch.rand(1) < p`` - the order a torchvision Compose([RandomResizedCrop, RandomHorizontalFlip])
    consumes the generator in.  Returns (int32 [n,4], uint8 [n]) CPU tensors."""
    if not random_crop:
        return (torch.tensor([[0, 0, height, width]] * n, dtype=torch.int32), torch.zeros(n, dtype=torch.uint8))
    rows, flips = [], []
    d = _Draw(generator)
    for _ in range(n):
        rows.append(rrc_params(height, width, _draw=d))
        flips.append(1 if d.uniform(0.0, 1.0) < flip_p else 0)
    return torch.tensor(rows, dtype=torch.int32), torch.tensor(flips, dtype=torch.uint8)


class DeviceInputStage:
    """uint8 NHWC host (pinned) or device batch -> normalised channels-last activation on the GPU.
    The default output layout is the space-to-depth form B200ResNet's stem consumes directly
    (``_cabi.LAYOUT_S2D_STEM``); pass ``layout=_cabi.LAYOUT_NHWC`` for a plain [N,3,H,W] tensor."""

    def __init__(self, out_size, mean, std, dtype=torch.bfloat16, device=None, layout=_cabi.LAYOUT_S2D_STEM):
        self.out_h, self.out_w = (out_size, out_size) if isinstance(out_size, int) else out_size
        self.mean, self.std, self.dtype, self.layout = tuple(mean), tuple(std), dtype, layout
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)

    def __call__(self, images_u8, boxes, flips):
        dev = self.device
        images_u8 = images_u8.to(dev, non_blocking=True)
        boxes = boxes.to(dev, non_blocking=True)
        flips = None if flips is None else flips.to(dev, non_blocking=True)
        return _cabi.input_stage(images_u8, boxes, flips, self.out_h, self.out_w, self.mean, self.std, self.dtype,
                                 self.layout)


class SyntheticImageNet:
    """Iterable of (input, target) like the reference's DataLoader (imagenet.py:350-354), over
    synthetic uint8 images.

    ``host=True`` keeps the image batch in pinned host memory and copies it to the GPU EVERY step
    (the end-to-end path: what ``pin_memory=True`` + ``.cuda(non_blocking=True)`` do in the
    reference, imagenet.py:119,354), double-buffered on a side stream so the copy of step i+1 overlaps
    the compute of step i.  ``host=False`` keeps the images resident in HBM.  Images are seeded per
    rank (``seed + rank``); one fixed image batch is re-used, the crop boxes / flips / targets are
    drawn per step from the same generator unless ``fixed_params``.

    ``sampler`` (a ``DistributedSampler`` over a virtual dataset of that many images; ``steps`` is then derived):
    the epoch follows the sampler exactly as the reference's loaders follow theirs (imagenet.py:346-359,375) - this
    rank's index stream of the current epoch cut into batches, the last one short (no ``drop_last``), the label of
    virtual image ``i`` being ``i % num_classes``; call ``sampler.set_epoch(e)`` before each epoch.
    """

    def __init__(self, batch_size, steps, src_size=256, out_size=224, num_classes=1000, rank=0, seed=1234,
                 random_crop=True, mean=IMAGENET_MEAN, std=IMAGENET_STD, dtype=torch.bfloat16, host=False,
                 fixed_params=False, device=None, sampler=None):
        self.sampler = sampler
        if sampler is not None:
            steps = -(-len(sampler) // batch_size)
        self.batch_size, self.steps, self.num_classes = batch_size, steps, num_classes
        self.src_h, self.src_w = (src_size, src_size) if isinstance(src_size, int) else src_size
        self.random_crop, self.host, self.fixed_params = random_crop, host, fixed_params
        self.gen = torch.Generator()
        self.gen.manual_seed(seed + rank)
        self.stage = DeviceInputStage(out_size, mean, std, dtype, device)
        img = torch.randint(0, 256, (batch_size, self.src_h, self.src_w, 3), dtype=torch.uint8, generator=self.gen)
        dev = self.stage.device
        if host:
            self.images = img.pin_memory()
            self._stream = torch.cuda.Stream(device=dev)
            self._slots = [{"img": torch.empty_like(img, device=dev), "ready": torch.cuda.Event(),
                            "free": torch.cuda.Event()} for _ in range(2)]
        else:
            self.images = img.to(dev)
        self._fixed = None
        self._queue = None
        self._ids = None             # sampler mode: this epoch's index batches still to be served
        self.h2d_bytes = (img.numel() if host else 0) + batch_size * (16 + 1 + 8)
        if host and sampler is None:
            self._start_producer()   # the first epoch's crop parameters are being drawn while the caller sets up

    def __len__(self):
        return self.steps

    def _epoch_batches(self):
        ids = self.sampler.indices()
        return [ids[i:i + self.batch_size] for i in range(0, len(ids), self.batch_size)]

    def _params(self):
        if self.sampler is not None:
            ids = self._ids.pop(0)
            boxes, flips = draw_batch_params(len(ids), self.src_h, self.src_w, self.gen, self.random_crop)
            return boxes.pin_memory(), flips.pin_memory(), (ids % self.num_classes).pin_memory()
        if self.fixed_params and self._fixed is not None:
            return self._fixed
        boxes, flips = draw_batch_params(self.batch_size, self.src_h, self.src_w, self.gen, self.random_crop)
        target = torch.randint(0, self.num_classes, (self.batch_size,), dtype=torch.int64, generator=self.gen)
        out = (boxes.pin_memory(), flips.pin_memory(), target.pin_memory())
        if self.fixed_params:
            self._fixed = out
        return out

    def _upload(self, slot, first_use, params):
        """Enqueue the H2D copies of one step on the side stream."""
        dev = self.stage.device
        boxes, flips, target = params
        with torch.cuda.stream(self._stream):
            if not first_use:
                self._stream.wait_event(slot["free"])  # the kernel that read this slot has finished
            slot["img"].copy_(self.images, non_blocking=True)
            slot["boxes"] = boxes.to(dev, non_blocking=True)
            slot["flips"] = flips.to(dev, non_blocking=True)
            slot["target"] = target.to(dev, non_blocking=True)
            slot["ready"].record(self._stream)

    def _start_producer(self):
        """Helper thread that draws the per-step (boxes, flips, target) a few steps ahead of the consumer, in the same
        order a single thread would draw them (the generator is only touched by the helper).  Like the workers of a
        DataLoader it starts when the loader is built / an epoch ends, not at the first ``next()``."""
        import queue
        import threading

        if self.sampler is not None:
            self._ids = self._epoch_batches()     # read the sampler's epoch NOW (set_epoch came before iter())
        q = queue.Queue(maxsize=4)

        def work():
            try:
                for _ in range(self.steps):
                    q.put(self._params())
            except BaseException as e:  # surface errors in the consumer
                q.put(e)

        threading.Thread(target=work, daemon=True).start()
        self._queue = q

    def _param_stream(self):
        if self._queue is None:
            self._start_producer()
        q, self._queue = self._queue, None
        for _ in range(self.steps):
            item = q.get()
            if isinstance(item, BaseException):
                raise item
            yield item

    def __iter__(self):
        dev = self.stage.device
        if not self.host:
            if self.sampler is not None:
                self._ids = self._epoch_batches()
            for _ in range(self.steps):
                boxes, flips, target = self._params()
                x = self.stage(self.images[:len(target)], boxes, flips)
                yield x, target.to(dev, non_blocking=True)
            return
        main = torch.cuda.current_stream(dev)
        params = self._param_stream()
        if self.steps > 0:
            self._upload(self._slots[0], True, next(params))
        for i in range(self.steps):
            slot = self._slots[i % 2]
            if i + 1 < self.steps:
                self._upload(self._slots[(i + 1) % 2], i + 1 < 2, next(params))
            main.wait_event(slot["ready"])
            x = _cabi.input_stage(slot["img"][:len(slot["target"])], slot["boxes"], slot["flips"], self.stage.out_h, self.stage.out_w,
                                  self.stage.mean, self.stage.std, self.stage.dtype, self.stage.layout)
            slot["free"].record(main)
            for k in ("boxes", "flips", "target"):
                slot[k].record_stream(main)
            yield x, slot["target"]
