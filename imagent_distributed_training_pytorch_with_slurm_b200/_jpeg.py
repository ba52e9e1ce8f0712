"""ctypes binding of ``include/b200_jpeg.h`` (libb200_jpeg.so): nvJPEG's batched decode pointed at the ragged input
slab.  A library of its own (it links libnvjpeg); it is loaded only when ``decoder="nvjpeg"`` is asked for, and a missing
library raises - there is no silent fall-back to another decoder."""
import ctypes
import os

import torch

from ._cabi import B200Error

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200_jpeg.so")
_c = ctypes
_vp, _i32, _i64, _sz = _c.c_void_p, _c.c_int, _c.c_int64, _c.c_size_t
NOT_SUPPORTED = 1004          # 1000 + NVJPEG_STATUS_JPEG_NOT_SUPPORTED
BAD_JPEG = 1003

SIGNATURES = {
    "b200_jpeg_create": (_i32, [_i32, _i32, _c.POINTER(_vp)]),
    "b200_jpeg_destroy": (None, [_vp]),
    "b200_jpeg_has_hardware": (_i32, [_vp]),
    "b200_jpeg_last_hardware_count": (_i32, [_vp]),
    "b200_jpeg_info": (_i32, [_vp, _vp, _sz, _c.POINTER(_i32), _c.POINTER(_i32), _c.POINTER(_i32)]),
    "b200_jpeg_decode_batch": (_i32, [_vp, _c.POINTER(_vp), _c.POINTER(_sz), _i32, _vp, _c.POINTER(_i64), _vp]),
    "b200_jpeg_error_string": (_c.c_char_p, [_i32]),
}
_lib = None


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B200Error(f"{LIB_PATH} is missing: build it (python -m "
                            "imagent_distributed_training_pytorch_with_slurm_b200.build); decoder='nvjpeg' has no fall-back")
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


class JpegError(B200Error):
    def __init__(self, code, what):
        self.code = code
        msg = load().b200_jpeg_error_string(code)
        super().__init__(f"{what} failed: {msg.decode() if msg else code} (code {code})")


class JpegDecoder:
    """One nvJPEG batched decoder on the current CUDA device (not thread-safe: one per thread)."""

    def __init__(self, cpu_threads=8, prefer_hardware=True):
        self._h = _vp()
        rc = load().b200_jpeg_create(int(cpu_threads), 1 if prefer_hardware else 0, ctypes.byref(self._h))
        if rc:
            raise JpegError(rc, "b200_jpeg_create")

    def close(self):
        if self._h:
            load().b200_jpeg_destroy(self._h)
            self._h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001  (interpreter shutdown)
            pass

    @property
    def has_hardware(self):
        return bool(load().b200_jpeg_has_hardware(self._h))

    @property
    def last_hardware_count(self):
        return int(load().b200_jpeg_last_hardware_count(self._h))

    def info(self, data):
        """(height, width, components) from the header of a JPEG byte string."""
        h, w, c = _i32(), _i32(), _i32()
        rc = load().b200_jpeg_info(self._h, ctypes.cast(ctypes.c_char_p(data), _vp), len(data), ctypes.byref(h),
                                   ctypes.byref(w), ctypes.byref(c))
        if rc:
            raise JpegError(rc, "b200_jpeg_info")
        return h.value, w.value, c.value

    def decode_batch(self, blobs, slab, table):
        """``blobs``: list of bytes; ``slab``: uint8 CUDA tensor; ``table``: int64 CPU tensor [n,4] = (byte offset, height,
        width, row pitch) per image.  Interleaved RGB rows are written on the current stream."""
        n = len(blobs)
        if not (slab.is_cuda and slab.dtype == torch.uint8 and slab.is_contiguous()):
            raise B200Error("slab must be a contiguous uint8 CUDA tensor")
        if not (table.dtype == torch.int64 and tuple(table.shape) == (n, 4) and table.is_contiguous() and not table.is_cuda):
            raise B200Error("table must be a contiguous int64 CPU tensor [n,4]")
        if n and int((table[:, 0] + table[:, 1] * table[:, 3]).max()) > slab.numel():
            raise B200Error("an image would be written past the end of the slab")
        ptrs = (_vp * n)(*[ctypes.cast(ctypes.c_char_p(b), _vp) for b in blobs])
        lens = (_sz * n)(*[len(b) for b in blobs])
        rc = load().b200_jpeg_decode_batch(self._h, ptrs, lens, n, slab.data_ptr(),
                                           ctypes.cast(table.data_ptr(), _c.POINTER(_i64)),
                                           torch.cuda.current_stream().cuda_stream)
        if rc:
            raise JpegError(rc, "b200_jpeg_decode_batch")
