"""The on-disk data front end: class-folder index -> DistributedSampler -> JPEG decode -> ragged uint8 batch -> K1.

Replaces ``datasets.ImageNet(root, split, transform)`` (imagenet.py:287-296), ``DistributedSampler`` +
``set_epoch`` (imagenet.py:346-347,375) and the two ``DataLoader(batch_size, num_workers=10, pin_memory=True)``
objects (imagenet.py:350-359).  What the reference does per image inside a worker process - ``pil_loader`` (open,
``convert("RGB")``), ``Resize((448, 448))``, ``ToTensor``, ``Normalize`` - is split here:

  * index      ``ImageFolderIndex``: ``<root>/<split>/<class>/<image>``, classes = sorted folder names, samples in
               torchvision's ``make_dataset`` order (torchvision/datasets/folder.py), so dataset index i and class index
               mean what they mean in the reference (``datasets.ImageNet`` is an ``ImageFolder`` over the extracted
               split folder);
  * order      this package's ``DistributedSampler`` (bit-exact with torch's), ``drop_last=False`` batching like the
               reference's loaders - the last batch of an epoch is short;
  * decode     ``decoder="pil"``: Pillow in host threads, exactly the reference's pixels;
               ``decoder="nvjpeg"``: nvJPEG's batched API driven directly (``include/b200_jpeg.h``; a LIBRARY, like
               cuDNN for the convolutions): the GPU's JPEG engines where a file allows it, otherwise the GPU-assisted
               decoder with a pool of host threads for the Huffman stage, interleaved RGB written straight into the
               slab, issued from the producer thread on a stream of its own;
               ``decoder="torchvision"``: nvJPEG through ``torchvision.io.decode_jpeg(device="cuda")``, planar output
               re-packed by ``b200_interleave_planar_u8`` (kept as a cross-check; torchvision serialises its decoder).
               Files nvJPEG cannot take (CMYK, PNG-in-disguise, ...) are decoded by Pillow one by one.  JPEG decoders
               differ in IDCT rounding and chroma up-sampling, so these paths are within a few grey levels of Pillow's,
               not bit-exact (measured in tests/test_gpu_folder.py);
  * transform  everything after the decode is the fused input kernel on the RAGGED batch
               (``b200_input_crop_resize_flip_norm_ragged``): images keep their own sizes in one byte slab, the crop box
               of the reference mode is the whole image.

A producer thread keeps ``prefetch`` batches decoded ahead of the consumer (the reference's 10 worker processes).
"""
import os
import queue
import threading
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import torch

from . import _cabi
from .data import IMAGENET_MEAN, IMAGENET_STD, _Draw, rrc_params

IMG_EXTENSIONS = (".jpg", ".jpeg", ".png", ".ppm", ".bmp", ".pgm", ".tif", ".tiff", ".webp")
_ALIGN = 16          # every image starts on a 16-byte boundary of the slab
_SLAB_TAIL = 64      # readable bytes past the last image (the stem-layout kernel fetches whole words)


class ImageFolderIndex:
    """``datasets.ImageNet(root, split)`` / ``ImageFolder`` without the images: ``samples`` = [(path, class index)],
    ``classes`` = sorted class-folder names, ``targets``."""

    def __init__(self, root, split=None):
        self.root = os.path.join(root, split) if split else root
        if not os.path.isdir(self.root):
            raise FileNotFoundError(f"no such dataset folder: {self.root}")
        self.classes = sorted(e.name for e in os.scandir(self.root) if e.is_dir())
        if not self.classes:
            raise FileNotFoundError(f"Couldn't find any class folder in {self.root}.")
        self.class_to_idx = {c: i for i, c in enumerate(self.classes)}
        self.samples = []
        for c in self.classes:
            d = os.path.join(self.root, c)
            for base, _, names in sorted(os.walk(d, followlinks=True)):
                for name in sorted(names):
                    if name.lower().endswith(IMG_EXTENSIONS):
                        self.samples.append((os.path.join(base, name), self.class_to_idx[c]))
        if not self.samples:
            raise FileNotFoundError(f"Found no valid file for the classes in {self.root}.")
        self.targets = [t for _, t in self.samples]

    def __len__(self):
        return len(self.samples)


def _pil_decode(data, limit):
    """bytes -> uint8 [h, w, 3] as torchvision's pil_loader (``Image.open(f).convert("RGB")``).  Images beyond the
    input kernel's 8x down-scaling limit are first shrunk by an integer factor with Pillow's box filter (a handful of
    multi-megapixel files in ImageNet; the reference resizes them in one step, so those few differ slightly)."""
    import io

    from PIL import Image

    img = Image.open(io.BytesIO(data)).convert("RGB")
    f = max(-(-img.height // limit[0]), -(-img.width // limit[1]))
    if f > 1:
        img = img.reduce(f)
    return np.asarray(img)


def _pil_open(data, limit):
    """Header only: (lazy PIL image, integer shrink factor, height, width after the shrink)."""
    import io

    from PIL import Image

    img = Image.open(io.BytesIO(data))
    f = max(-(-img.height // limit[0]), -(-img.width // limit[1]), 1)
    return img, f, -(-img.height // f), -(-img.width // f)


def _pil_decode_into(img, f, dst):
    """Decode (``convert("RGB")``, as pil_loader) straight into ``dst`` = uint8 [h, w, 3] view of the pinned slab."""
    img = img.convert("RGB")
    if f > 1:
        img = img.reduce(f)
    np.copyto(dst, np.asarray(img))


class _PinnedSlabs:
    """A few pinned host buffers reused round-robin for the decoded batches (allocating page-locked memory per batch
    costs more than decoding it).  ``events[i]`` is recorded by the consumer after the host->device copy out of buffer i
    was enqueued; the producer waits on it before writing into that buffer again."""

    def __init__(self, count):
        self.bufs, self.events, self.next = [None] * count, [None] * count, 0

    def take(self, nbytes):
        i = self.next
        self.next = (i + 1) % len(self.bufs)
        if self.events[i] is not None:
            self.events[i].synchronize()
            self.events[i] = None
        if self.bufs[i] is None or self.bufs[i].numel() < nbytes:
            self.bufs[i] = torch.empty(int(nbytes * 1.2) + 4096, dtype=torch.uint8).pin_memory()
        return i, self.bufs[i]


class ImageFolderLoader:
    """Iterable of ``(input, target)`` device tensors over an ``ImageFolderIndex`` in sampler order.

    ``random_crop=False`` is the reference's transform (whole image -> ``out_size``, no flip, pass mean = std = 0.5);
    ``random_crop=True`` draws a RandomResizedCrop box + flip per image from a per-rank CPU generator in torchvision's
    draw order.  ``layout``: the model's space-to-depth stem layout by default, ``_cabi.LAYOUT_NHWC`` for a plain
    channels-last [N,3,H,W] tensor."""

    def __init__(self, index, sampler, batch_size, out_size=448, mean=IMAGENET_MEAN, std=IMAGENET_STD, random_crop=False,
                 dtype=torch.bfloat16, decoder="pil", workers=8, prefetch=2, seed=1234, rank=0, device=None,
                 layout=_cabi.LAYOUT_S2D_STEM):
        if decoder not in ("pil", "nvjpeg", "torchvision"):
            raise ValueError("decoder must be 'pil', 'nvjpeg' or 'torchvision'")
        self.index, self.sampler, self.batch_size = index, sampler, int(batch_size)
        self.out_h, self.out_w = (out_size, out_size) if isinstance(out_size, int) else out_size
        self.mean, self.std, self.random_crop, self.dtype = tuple(mean), tuple(std), random_crop, dtype
        self.decoder, self.workers, self.prefetch, self.layout = decoder, max(1, int(workers)), max(1, int(prefetch)), layout
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.gen = torch.Generator()
        self.gen.manual_seed(seed + rank)
        self._draw = _Draw(self.gen)
        self.limit = (8 * self.out_h, 8 * self.out_w)
        self.h2d_bytes = 0           # bytes copied host -> device by the last batch
        self._slabs = _PinnedSlabs(self.prefetch + 3)
        self._jpeg_tls = threading.local()           # decoder="nvjpeg": one nvJPEG decoder + stream per decode thread
        self.decode_threads = 2                      # batches being decoded on the device at the same time
        self._pool = self._dec = None                # thread pools, created on first use and kept across epochs (the
        #                                              decode threads own nvJPEG decoder objects: creating those per epoch
        #                                              would cost device allocations every time)

    def _pools(self):
        if self._pool is None:
            self._pool = ThreadPoolExecutor(self.workers, thread_name_prefix="b200-io")
            if self.decoder == "nvjpeg":
                self._dec = ThreadPoolExecutor(self.decode_threads, thread_name_prefix="b200-jpeg")
        return self._pool, self._dec

    def close(self):
        """Stop the worker threads (the nvJPEG decoder objects go with their threads)."""
        for ex in (self._pool, self._dec):
            if ex is not None:
                ex.shutdown(wait=False, cancel_futures=True)
        self._pool = self._dec = None

    def __len__(self):
        return -(-len(self.sampler) // self.batch_size)       # drop_last=False

    # ---- host side of one batch (producer thread) -----------------------------------------------------
    def _read(self, path):
        with open(path, "rb") as f:
            return f.read()

    def _boxes(self, sizes):
        rows, flips = [], []
        for h, w in sizes:
            if self.random_crop:
                rows.append(rrc_params(h, w, _draw=self._draw))
                flips.append(1 if self._draw.uniform(0.0, 1.0) < 0.5 else 0)
            else:
                rows.append((0, 0, h, w))
                flips.append(0)
        return torch.tensor(rows, dtype=torch.int32), torch.tensor(flips, dtype=torch.uint8)

    @staticmethod
    def _layout(sizes):
        """byte offsets of the images in the slab (16-byte aligned), total slab bytes"""
        offs, off = [], 0
        for h, w in sizes:
            offs.append(off)
            off += (h * w * 3 + _ALIGN - 1) // _ALIGN * _ALIGN
        return offs, off + _SLAB_TAIL

    def _probe(self, data):
        """(height, width, pixels or None): header only for files nvJPEG will take; anything else (not a JPEG, CMYK,
        beyond the input kernel's down-scaling limit) is decoded here with Pillow."""
        import io

        from PIL import Image

        img = Image.open(io.BytesIO(data))
        if (img.format == "JPEG" and img.mode in ("RGB", "L") and img.height <= self.limit[0]
                and img.width <= self.limit[1]):
            return img.height, img.width, None
        a = _pil_decode(data, self.limit)
        return a.shape[0], a.shape[1], a

    def _host_batch(self, pool, ids):
        paths = [self.index.samples[i][0] for i in ids]
        target = torch.tensor([self.index.samples[i][1] for i in ids], dtype=torch.int64)
        blobs = list(pool.map(self._read, paths))
        if self.decoder == "pil":
            opened = list(pool.map(lambda b: _pil_open(b, self.limit), blobs))      # headers: sizes without decoding
            sizes = [(h, w) for _, _, h, w in opened]
            arrays = None
        else:
            probes = list(pool.map(self._probe, blobs))
            sizes = [(h, w) for h, w, _ in probes]
            arrays = [a for _, _, a in probes]
        offs, total = self._layout(sizes)
        boxes, flips = self._boxes(sizes)            # drawn here: the generator is only ever touched by this thread
        table = torch.tensor([[o, h, w, 3 * w] for o, (h, w) in zip(offs, sizes)], dtype=torch.int64)
        item = {"table": table.pin_memory(), "boxes": boxes.pin_memory(), "flips": flips.pin_memory(),
                "target": target.pin_memory(), "sizes": sizes, "offs": offs, "total": total}
        if self.decoder == "pil":
            # every worker decodes its image straight into its slot of a recycled pinned slab
            slot, slab = self._slabs.take(total)
            sv = slab.numpy()

            def work(k):
                img, f, h, w = opened[k]
                _pil_decode_into(img, f, sv[offs[k]:offs[k] + h * w * 3].reshape(h, w, 3))

            list(pool.map(work, range(len(opened))))
            sv[total - _SLAB_TAIL:total] = 0
            item["slab"], item["slot"] = slab[:total], slot
        elif self.decoder == "nvjpeg":
            item["blobs"], item["arrays"] = blobs, arrays     # decoded by a decode thread (see __iter__)
        else:
            item["blobs"], item["arrays"] = blobs, arrays
        return item

    def _decode_on_device(self, item):
        """Decode thread: nvJPEG batched decode of the whole batch into a device slab, on that thread's own decoder
        object and stream; the consumer only waits for ``item["ready"]``.  Files the header probe kept for Pillow are
        copied in."""
        from . import _jpeg

        dev, tls = self.device, self._jpeg_tls
        with torch.cuda.device(dev):
            if getattr(tls, "jpeg", None) is None:
                # high priority: beside a training step that keeps every SM busy, the decode kernels of the NEXT batches
                # must not queue behind a whole iteration's worth of already-enqueued work
                tls.stream = torch.cuda.Stream(device=dev, priority=-1)
                tls.jpeg = _jpeg.JpegDecoder(cpu_threads=max(1, self.workers // self.decode_threads), prefer_hardware=True)
            jpeg, stream = tls.jpeg, tls.stream
            sizes, offs, total, arrays = item["sizes"], item["offs"], item["total"], item["arrays"]
            with torch.cuda.stream(stream):
                slab = torch.empty(total, dtype=torch.uint8, device=dev)
                slab[total - _SLAB_TAIL:].zero_()
                todo = [i for i, a in enumerate(arrays) if a is None]
                for i, a in enumerate(arrays):
                    if a is not None:
                        slab[offs[i]:offs[i] + a.size].copy_(torch.from_numpy(np.array(a, copy=True)).view(-1))
                if todo:
                    sub = item["table"][todo].contiguous()
                    try:
                        jpeg.decode_batch([item["blobs"][i] for i in todo], slab, sub)
                    except _jpeg.JpegError as e:
                        if e.code not in (_jpeg.NOT_SUPPORTED, _jpeg.BAD_JPEG):
                            raise
                        for i in todo:      # find the file(s) nvJPEG refuses and give exactly those to Pillow
                            try:
                                jpeg.decode_batch([item["blobs"][i]], slab, item["table"][i:i + 1].contiguous())
                            except _jpeg.JpegError:
                                a = _pil_decode(item["blobs"][i], self.limit)
                                slab[offs[i]:offs[i] + a.size].copy_(torch.from_numpy(np.array(a, copy=True)).view(-1))
                ev = torch.cuda.Event()
                ev.record()
            item["slab_dev"], item["ready"] = slab, ev
            item["hw_count"] = jpeg.last_hardware_count
            self.jpeg_has_hardware = jpeg.has_hardware
            stream.synchronize()      # the byte strings and the staging arrays may go once the GPU has them
        return item

    # ---- device side of one batch (consumer thread) ---------------------------------------------------
    def _gpu_decode(self, item):
        """nvJPEG decode (torchvision, a library) of every file the probe cleared, Pillow pixels for the rest; returns
        planar device tensors [3,h,w] uint8 in batch order."""
        from torchvision.io import ImageReadMode, decode_jpeg

        dev = self.device
        todo = [i for i, a in enumerate(item["arrays"]) if a is None]
        raw = [torch.frombuffer(bytearray(item["blobs"][i]), dtype=torch.uint8) for i in todo]
        planes = [None] * len(item["arrays"])
        try:
            done = decode_jpeg(raw, device=dev, mode=ImageReadMode.RGB) if raw else []
        except RuntimeError:
            done = []
            for i, r in zip(todo, raw):      # find the file nvJPEG refuses and hand that one to Pillow
                try:
                    done.append(decode_jpeg(r, device=dev, mode=ImageReadMode.RGB))
                except RuntimeError:
                    done.append(None)
        for i, p in zip(todo, done):
            planes[i] = p
        for i, p in enumerate(planes):
            if p is None:
                a = item["arrays"][i] if item["arrays"][i] is not None else _pil_decode(item["blobs"][i], self.limit)
                planes[i] = torch.from_numpy(np.ascontiguousarray(a.transpose(2, 0, 1))).to(dev)
            if tuple(planes[i].shape[1:]) != tuple(item["sizes"][i]):
                raise _cabi.B200Error(f"decoded size {tuple(planes[i].shape[1:])} differs from the header's "
                                      f"{item['sizes'][i]}")
        return planes

    def _stage(self, item):
        dev = self.device
        sizes, offs, total = item["sizes"], item["offs"], item["total"]
        table = item["table"].to(dev, non_blocking=True)
        if self.decoder == "pil":
            slab = item["slab"].to(dev, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
            self._slabs.events[item["slot"]] = ev          # the producer may refill this buffer once the copy is done
            self.h2d_bytes = total + len(sizes) * (32 + 16 + 1 + 8)
        elif self.decoder == "nvjpeg":
            slab = item["slab_dev"]
            torch.cuda.current_stream(dev).wait_event(item["ready"])
            slab.record_stream(torch.cuda.current_stream(dev))
            self.last_hw_count = item["hw_count"]
            self.h2d_bytes = sum(len(b) for b in item["blobs"]) + len(sizes) * (32 + 16 + 1 + 8)
        else:
            planes = self._gpu_decode(item)
            slab = torch.empty(total, dtype=torch.uint8, device=dev)
            slab[total - _SLAB_TAIL:].zero_()
            src = torch.tensor([[p.data_ptr(), h, w, o] for p, (h, w), o in zip(planes, sizes, offs)], dtype=torch.int64)
            _cabi.interleave_planar(src.to(dev), slab, max(h * w for h, w in sizes))
            for p in planes:
                p.record_stream(torch.cuda.current_stream(dev))
            self.h2d_bytes = sum(len(b) for b in item["blobs"]) + len(sizes) * (32 + 32 + 16 + 1 + 8)
        boxes = item["boxes"].to(dev, non_blocking=True)
        flips = item["flips"].to(dev, non_blocking=True)
        target = item["target"].to(dev, non_blocking=True)
        max_h, max_w = max(h for h, _ in sizes), max(w for _, w in sizes)
        x = _cabi.input_stage_ragged(slab, table, boxes, flips, max_h, max_w, self.out_h, self.out_w, self.mean, self.std,
                                     self.dtype, self.layout)
        self.last_batch = {"slab": slab, "table": table, "sizes": sizes, "boxes": boxes, "flips": flips}   # for tests
        return x, target

    def __iter__(self):
        ids = self.sampler.indices().tolist()
        batches = [ids[i:i + self.batch_size] for i in range(0, len(ids), self.batch_size)]
        q = queue.Queue(maxsize=self.prefetch)
        # decoder="nvjpeg": the producer prepares batches in order (file reads, header probe, crop draws) and hands each
        # to one of `decode_threads` decode threads; the queue carries the futures, so batches still arrive in order
        pool, dec = self._pools()
        stop = threading.Event()

        def put(item):
            while not stop.is_set():       # a consumer that stopped early must not leave this thread blocked forever
                try:
                    q.put(item, timeout=0.2)
                    return True
                except queue.Full:
                    pass
            return False

        def produce():
            try:
                for b in batches:
                    item = self._host_batch(pool, b)
                    if dec is not None:
                        item = dec.submit(self._decode_on_device, item)
                    if not put(item):
                        return
            except BaseException as e:  # noqa: BLE001  (surface it in the consumer)
                put(e)

        threading.Thread(target=produce, daemon=True).start()
        try:
            for _ in batches:
                item = q.get()
                if isinstance(item, BaseException):
                    raise item
                if dec is not None:
                    item = item.result()
                yield self._stage(item)
        finally:
            stop.set()                     # an early exit: the producer stops, decodes not yet started are dropped
            while True:
                try:
                    item = q.get_nowait()
                except queue.Empty:
                    break
                if dec is not None and hasattr(item, "cancel"):
                    item.cancel()
