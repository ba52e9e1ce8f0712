"""ctypes binding of ``include/b200_imagenet.h`` (the C ABI of the sm_100a kernels).

There is no CPU fallback: if the shared library is missing or a call returns a non-zero code this
module raises.  The thin wrappers below take torch CUDA tensors only for their device pointers,
sizes and the current stream; all arithmetic happens in the CUDA kernels.
"""
import ctypes
import os
import threading

import torch

from . import _nvtx

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200_imagenet.so")

F32, BF16 = 0, 1
ACT_NONE, ACT_RELU, ACT_ADD_RELU = 0, 1, 2
LAYOUT_NHWC, LAYOUT_NCHW, LAYOUT_S2D_STEM = 0, 1, 2
ABI_VERSION = 2

_c = ctypes
_vp, _i64, _i32, _f32, _sz = _c.c_void_p, _c.c_int64, _c.c_int, _c.c_float, _c.c_size_t

# name -> (restype, argtypes); mirrors include/b200_imagenet.h one to one
SIGNATURES = {
    "b200_abi_version": (_i32, []),
    "b200_error_string": (_c.c_char_p, [_i32]),
    "b200_tune_set": (_i32, [_c.c_char_p, _i32]),
    "b200_sgd_flat_f32": (_i32, [_vp, _vp, _vp, _vp, _i64, _f32, _f32, _f32, _f32, _i32, _vp]),
    "b200_ce_topk_workspace_bytes": (_sz, [_i64]),
    "b200_ce_topk_fwd": (_i32, [_vp, _vp, _vp, _vp, _vp, _sz, _i64, _i64, _i32, _vp]),
    "b200_ce_topk_fwd_static": (_i32, [_vp, _vp, _vp, _vp, _i64, _i64, _i32, _vp]),
    "b200_target_rank": (_i32, [_vp, _vp, _vp, _i64, _i64, _i32, _vp]),
    "b200_ce_bwd": (_i32, [_vp, _vp, _vp, _vp, _f32, _vp, _i64, _i64, _i32, _vp]),
    "b200_bn_vec": (_i32, [_i32]),
    "b200_bn_workspace_bytes": (_sz, [_i64, _i32, _i32]),
    "b200_bn_fwd_train": (_i32, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _i64, _i32, _f32, _f32,
                                 _i32, _i32, _vp]),
    "b200_bn_fwd_eval": (_i32, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i32, _f32, _i32, _i32, _vp]),
    "b200_bn_bwd": (_i32, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _i64, _i32, _i32, _i32, _vp]),
    "b200_maxpool3x3s2_fwd": (_i32, [_vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _vp]),
    "b200_maxpool3x3s2_bwd": (_i32, [_vp, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _vp]),
    "b200_bn_relu_maxpool3x3s2_fwd": (_i32, [_vp, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _vp]),
    "b200_bn_dual_supported": (_i32, [_i32, _i32]),
    "b200_bn_stats": (_i32, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _i64, _i32, _f32, _f32, _i32, _vp]),
    "b200_bn_fwd_train_dual": (_i32, [_vp] * 12 + [_sz, _i64, _i32, _f32, _f32, _i32, _vp]),
    "b200_bn_bwd_dual": (_i32, [_vp] * 16 + [_sz, _i64, _i32, _i32, _vp]),
    "b200_pack_grads": (_i32, [_c.POINTER(_vp), _c.POINTER(_i64), _c.POINTER(_i64), _i32, _vp, _f32, _i32, _vp]),
    "b200_peer_pad_bytes": (_sz, []),
    "b200_peer_exchange_sgd": (_i32, [_c.POINTER(_vp), _c.POINTER(_vp), _c.POINTER(_vp), _c.POINTER(_vp), _vp, _i32, _i32,
                                      _i32, _i64, _i64, _f32, _f32, _f32, _f32, _i32, _i32, _vp]),
    "b200_peer_bcast": (_i32, [_c.POINTER(_vp), _c.POINTER(_vp), _i64, _i32, _i32, _i32, _vp]),
    "b200_peer_wait": (_i32, [_vp, _i32, _i32, _i32, _i32, _vp]),
    "b200_input_workspace_bytes": (_sz, [_i64, _i32, _i32, _i32, _i32]),
    "b200_input_crop_resize_flip_norm": (_i32, [_vp, _vp, _vp, _vp, _vp, _sz, _i64, _i32, _i32, _i32, _i32,
                                                _c.POINTER(_f32), _c.POINTER(_f32), _i32, _i32, _vp]),
    "b200_input_crop_resize_flip_norm_ragged": (_i32, [_vp, _i64, _vp, _vp, _vp, _vp, _vp, _sz, _i64, _i32, _i32, _i32,
                                                       _i32, _c.POINTER(_f32), _c.POINTER(_f32), _i32, _i32, _vp]),
    "b200_interleave_planar_u8": (_i32, [_vp, _vp, _i64, _i64, _vp]),
}

_lib = None
_lock = threading.Lock()


class B200Error(RuntimeError):
    pass


def load():
    """dlopen the library (once).  Raises B200Error when it has not been built - no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise B200Error(
                f"{LIB_PATH} is missing: build it with `python -m imagent_distributed_training_pytorch_with_slurm_b200.build` "
                "(nvcc, sm_100a). There is no CPU or PyTorch fallback for these kernels.")
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError = header/library mismatch
            fn.restype = res
            fn.argtypes = args
        v = lib.b200_abi_version()
        if v != ABI_VERSION:
            raise B200Error(f"ABI version mismatch: library {v}, binding {ABI_VERSION}")
        _lib = lib
    return _lib


def _check(code, what):
    if code != 0:
        msg = load().b200_error_string(code)
        raise B200Error(f"{what} failed: {msg.decode() if msg else code} (code {code})")


def tune_set(key, value):
    _check(load().b200_tune_set(key.encode(), int(value)), f"b200_tune_set({key})")


def dtype_code(t):
    if t.dtype == torch.float32:
        return F32
    if t.dtype == torch.bfloat16:
        return BF16
    raise B200Error(f"unsupported activation dtype {t.dtype}")


def _ptr(t):
    return None if t is None else t.data_ptr()


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def _stream():
    """cudaStream_t of torch's current stream on the current device (the fast C accessor when this torch
    build has it: the Python ``current_stream()`` object costs ~15 us per call, 200+ calls per step)."""
    if _raw_stream is not None:
        return _raw_stream(torch.cuda.current_device())
    return torch.cuda.current_stream().cuda_stream


def _need_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise B200Error("B200 kernels need CUDA tensors; there is no CPU path")


# ---- launch accounting / per-kernel timing -----------------------------------------------------
class _Stats:
    """Counts our kernel launches and, when enabled, brackets chosen entry points with CUDA events
    on the launching stream so bench.py can report achieved GB/s of one kernel family."""

    def __init__(self):
        self.launches = 0
        self.timed = None  # set of family names or None
        self.records = []  # (family, bytes, start_event, end_event, bytes by SURVEY 8d's plain model)

    def reset(self):
        self.launches = 0
        self.records = []

    def begin(self, family):
        if self.timed is not None and family in self.timed:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            return e
        return None

    def end(self, family, n_launch, start, nbytes, nbytes_8d=None):
        """``nbytes``: what this call has to move given the fusions it implements; ``nbytes_8d``: the figure SURVEY
        8(d)'s un-fused per-site model assigns to the same work (3/4 E*s forward, 5/6 E*s backward per BatchNorm site,
        no mask, no fan-in), reported beside it."""
        self.launches += n_launch
        if start is not None:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            self.records.append((family, nbytes, start, e, nbytes if nbytes_8d is None else nbytes_8d))

    def summary(self):
        out = {}
        for fam, nbytes, s, e, n8d in self.records:
            d = out.setdefault(fam, {"calls": 0, "bytes": 0, "ms": 0.0, "bytes_8d": 0})
            d["calls"] += 1
            d["bytes"] += nbytes
            d["bytes_8d"] += n8d
            d["ms"] += s.elapsed_time(e)
        return out


stats = _Stats()

_workspaces = {}
_bn_ws_bytes = {}


def _bn_workspace(x, m, c, dt):
    """The shared BatchNorm scratch buffer, large enough for an [m, c] activation (size query cached)."""
    key = (m, c, dt)
    n = _bn_ws_bytes.get(key)
    if n is None:
        n = _bn_ws_bytes[key] = load().b200_bn_workspace_bytes(m, c, dt)
    return workspace(x.device, n, "bn")



_retired = []  # outgrown scratch buffers are kept alive: a captured CUDA graph may still point at them


def workspace(dev, nbytes, tag, zero=False):
    """Cached per-(device, tag) scratch buffer, grown on demand (PyTorch's allocator owns it)."""
    key = (dev.index, tag)
    buf = _workspaces.get(key)
    if buf is None or buf.numel() < nbytes:
        if buf is not None:
            _retired.append(buf)
        n = max(int(nbytes), 256)
        buf = torch.zeros(n, dtype=torch.uint8, device=dev) if zero else torch.empty(n, dtype=torch.uint8, device=dev)
        _workspaces[key] = buf
    return buf


# ---- K4b ------------------------------------------------------------------------------------------
def sgd_flat(p, g, buf, shadow, lr, weight_decay, momentum, grad_scale, first_step):
    _need_cuda(p, g, buf, shadow)
    n = p.numel()
    assert p.dtype == g.dtype == buf.dtype == torch.float32 and g.numel() == n and buf.numel() == n
    assert p.is_contiguous() and g.is_contiguous() and buf.is_contiguous()
    if shadow is not None:
        assert shadow.dtype == torch.bfloat16 and shadow.numel() == n and shadow.is_contiguous()
    s = stats.begin("sgd")
    _check(load().b200_sgd_flat_f32(_ptr(p), _ptr(g), _ptr(buf), _ptr(shadow), n, lr, weight_decay, momentum,
                                    grad_scale, 1 if first_step else 0, _stream()), "b200_sgd_flat_f32")
    stats.end("sgd", 1, s, n * ((16 if first_step else 20) + (2 if shadow is not None else 0)))


# ---- K3 -------------------------------------------------------------------------------------------
def ce_topk_fwd(logits, target, out3=None):
    """Returns (out3 float32[3] = loss, top1 %, top5 %; row_lse float32[N])."""
    _need_cuda(logits, target)
    assert logits.dim() == 2 and logits.is_contiguous() and target.dtype == torch.int64 and target.is_contiguous()
    n, c = logits.shape
    assert target.numel() == n
    lib = load()
    if out3 is None:
        out3 = torch.empty(3, dtype=torch.float32, device=logits.device)
    assert out3.dtype == torch.float32 and out3.numel() == 3 and out3.is_contiguous()
    row_lse = torch.empty(n, dtype=torch.float32, device=logits.device)
    nbytes = lib.b200_ce_topk_workspace_bytes(n)
    ws = workspace(logits.device, nbytes, "ce", zero=True)
    s = stats.begin("ce_topk_fwd")
    _check(lib.b200_ce_topk_fwd(_ptr(logits), _ptr(target), _ptr(out3), _ptr(row_lse), _ptr(ws), ws.numel(), n, c,
                                dtype_code(logits), _stream()), "b200_ce_topk_fwd")
    stats.end("ce_topk_fwd", 1, s, n * c * logits.element_size() + 8 * n + 12)
    return out3, row_lse


CE_STATIC_MAX_ROWS = 16384


def ce_topk_fwd_static(logits, target, out3, row_lse):
    """Loss + top-1/top-5 like ``ce_topk_fwd`` for code paths that must not allocate: nothing is
    allocated here either (no torch tensor is created; ``out3`` float32[3] and ``row_lse`` float32[N] are
    the caller's), and the kernel's scratch is a fixed array inside the library.  Meant for hooks that
    run when the process may no longer ask the driver or PyTorch's allocator for device memory.  At most
    ``CE_STATIC_MAX_ROWS`` rows; calls that could overlap on the GPU must not share a device.  The first
    call on a device should happen outside stream capture."""
    _need_cuda(logits, target, out3, row_lse)
    n, c = logits.shape
    if n > CE_STATIC_MAX_ROWS:
        raise B200Error(f"b200_ce_topk_fwd_static handles at most {CE_STATIC_MAX_ROWS} rows (its built-in scratch is "
                        f"fixed-size); got {n}. Use ce_topk_fwd, which takes a workspace.")
    if not (logits.is_contiguous() and target.is_contiguous() and target.dtype == torch.int64 and target.numel() == n):
        raise B200Error("logits [N,C] contiguous and int64 target [N] expected")
    if not (out3.dtype == torch.float32 and out3.numel() == 3 and out3.is_contiguous()
            and row_lse.dtype == torch.float32 and row_lse.numel() == n and row_lse.is_contiguous()):
        raise B200Error("out3 must be float32[3] and row_lse float32[N], both preallocated by the caller")
    _check(load().b200_ce_topk_fwd_static(_ptr(logits), _ptr(target), _ptr(out3), _ptr(row_lse), n, c,
                                          dtype_code(logits), _stream()), "b200_ce_topk_fwd_static")
    stats.launches += 1


def target_rank(logits, target):
    """int32[N]: how many classes rank ahead of each row's target (ties: the lower class index is ahead)."""
    _need_cuda(logits, target)
    assert logits.dim() == 2 and logits.is_contiguous() and target.dtype == torch.int64 and target.is_contiguous()
    n, c = logits.shape
    assert target.numel() == n
    rank = torch.empty(n, dtype=torch.int32, device=logits.device)
    s = stats.begin("target_rank")
    _check(load().b200_target_rank(_ptr(logits), _ptr(target), _ptr(rank), n, c, dtype_code(logits), _stream()),
           "b200_target_rank")
    stats.end("target_rank", 1, s, n * c * logits.element_size() + 12 * n)
    return rank


def ce_bwd(logits, target, row_lse, grad_out=None, scale=1.0):
    _need_cuda(logits, target, row_lse, grad_out)
    n, c = logits.shape
    dlogits = torch.empty_like(logits)
    if grad_out is not None:
        assert grad_out.dtype == torch.float32 and grad_out.numel() == 1
    s = stats.begin("ce_bwd")
    _check(load().b200_ce_bwd(_ptr(logits), _ptr(target), _ptr(row_lse), _ptr(grad_out), scale, _ptr(dlogits), n, c,
                              dtype_code(logits), _stream()), "b200_ce_bwd")
    stats.end("ce_bwd", 1, s, 2 * n * c * logits.element_size() + 12 * n)
    return dlogits


# ---- K2 -------------------------------------------------------------------------------------------
def _rows_channels(x):
    """(M, C) of a channels-last 4-d tensor or a contiguous 2-d [M, C] tensor."""
    if x.dim() == 4:
        if not x.is_contiguous(memory_format=torch.channels_last):
            raise B200Error("BatchNorm kernels need channels-last (NHWC) activations")
        n, c, h, w = x.shape
        return n * h * w, c
    if x.dim() == 2 and x.is_contiguous():
        return x.shape[0], x.shape[1]
    raise B200Error("expected a channels-last 4-d or contiguous 2-d activation")


def bn_vec(dtype):
    return load().b200_bn_vec(F32 if dtype == torch.float32 else BF16)


def bn_fwd_train(x, z, gamma, beta, running_mean, running_var, num_batches_tracked, eps, momentum, act):
    """Returns (y, saved float32[4,C], mask uint8 or None)."""
    _need_cuda(x, z, gamma, beta)
    m, c = _rows_channels(x)
    lib = load()
    dt = dtype_code(x)
    y = torch.empty_like(x)
    saved = torch.empty(4, c, dtype=torch.float32, device=x.device)
    mask = None
    if act == ACT_ADD_RELU:
        assert z is not None and z.shape == x.shape and z.dtype == x.dtype
        _rows_channels(z)
        # +16: the bulk-async backward fetches the mask in 16-byte units and may read past the last byte
        mask = torch.empty(m * (c // lib.b200_bn_vec(dt)) + 16, dtype=torch.uint8, device=x.device)
    ws = _bn_workspace(x, m, c, dt)
    s = stats.begin("bn_fwd")
    _check(lib.b200_bn_fwd_train(_ptr(x), _ptr(z), _ptr(y), _ptr(mask), _ptr(gamma), _ptr(beta), _ptr(running_mean),
                                 _ptr(running_var), _ptr(num_batches_tracked), _ptr(saved), _ptr(ws), ws.numel(), m, c,
                                 eps, momentum, act, dt, _stream()), "b200_bn_fwd_train")
    e = m * c * x.element_size()
    stats.end("bn_fwd", 3, s, (4 if act == ACT_ADD_RELU else 3) * e + (mask.numel() - 16 if mask is not None else 0),
              (4 if act == ACT_ADD_RELU else 3) * e)
    return y, saved, mask


def bn_fwd_eval(x, z, gamma, beta, running_mean, running_var, eps, act):
    _need_cuda(x, z, gamma, beta, running_mean, running_var)
    m, c = _rows_channels(x)
    y = torch.empty_like(x)
    s = stats.begin("bn_eval")
    _check(load().b200_bn_fwd_eval(_ptr(x), _ptr(z), _ptr(y), _ptr(gamma), _ptr(beta), _ptr(running_mean),
                                   _ptr(running_var), m, c, eps, act, dtype_code(x), _stream()), "b200_bn_fwd_eval")
    stats.end("bn_eval", 1, s, (3 if act == ACT_ADD_RELU else 2) * m * c * x.element_size())
    return y


_NOT_SUPPORTED = 801  # cudaErrorNotSupported


def bn_bwd(dy, x, mask, gamma, saved, act, dgamma=None, dbeta=None, dy2=None):
    """Returns (dx, dz or None, dgamma, dbeta); dgamma/dbeta may be preallocated float32 views
    (e.g. slices of the flat gradient buffer) and are overwritten.  ``dy2``: optional second gradient
    branch, summed with ``dy`` inside the kernels (residual fan-in)."""
    _need_cuda(dy, x, gamma, saved, dy2)
    if dy2 is not None and (dy2.shape != dy.shape or dy2.dtype != dy.dtype or dy2.stride() != dy.stride()):
        raise B200Error("dy2 must match dy")
    m, c = _rows_channels(x)
    if dy.shape != x.shape or dy.dtype != x.dtype:
        raise B200Error("dy must match x")
    _rows_channels(dy)
    lib = load()
    dt = dtype_code(x)
    dx = torch.empty_like(x)
    dz = torch.empty_like(x) if act == ACT_ADD_RELU else None
    if dgamma is None:
        dgamma = torch.empty(c, dtype=torch.float32, device=x.device)
    if dbeta is None:
        dbeta = torch.empty(c, dtype=torch.float32, device=x.device)
    ws = _bn_workspace(x, m, c, dt)
    s = stats.begin("bn_bwd")
    args = (_ptr(x), _ptr(mask), _ptr(gamma), _ptr(saved), _ptr(dx), _ptr(dz), _ptr(dgamma), _ptr(dbeta), _ptr(ws),
            ws.numel(), m, c, act, dt, _stream())
    rc = lib.b200_bn_bwd(_ptr(dy), _ptr(dy2), *args)
    if rc == _NOT_SUPPORTED and dy2 is not None:  # shape outside the bulk-async kernels: add the branches first
        dy = dy + dy2
        rc = lib.b200_bn_bwd(_ptr(dy), None, *args)
    _check(rc, "b200_bn_bwd")
    e = m * c * x.element_size()
    # algorithmic bytes: the two-pass model of SURVEY 8d; a second gradient branch is read in both passes
    # (+2E: it replaces the separate add kernel's 3E)
    stats.end("bn_bwd", 3, s, ((6 if act == ACT_ADD_RELU else 5) + (2 if dy2 is not None else 0)) * e
              + (2 * (mask.numel() - 16) if mask is not None else 0), (6 if act == ACT_ADD_RELU else 5) * e)
    return dx, dz, dgamma, dbeta


def bn_dual_supported(x):
    return bool(load().b200_bn_dual_supported(x.shape[1], dtype_code(x)))


def bn_stats(x, gamma, beta, running_mean, running_var, num_batches_tracked, eps, momentum):
    """Statistics + finalize only (no output tensor).  Returns saved float32[4,C]."""
    _need_cuda(x, gamma, beta)
    m, c = _rows_channels(x)
    lib = load()
    dt = dtype_code(x)
    saved = torch.empty(4, c, dtype=torch.float32, device=x.device)
    ws = _bn_workspace(x, m, c, dt)
    s = stats.begin("bn_fwd")
    _check(lib.b200_bn_stats(_ptr(x), _ptr(gamma), _ptr(beta), _ptr(running_mean), _ptr(running_var),
                             _ptr(num_batches_tracked), _ptr(saved), _ptr(ws), ws.numel(), m, c, eps, momentum, dt,
                             _stream()), "b200_bn_stats")
    stats.end("bn_fwd", 2, s, m * c * x.element_size(), 3 * m * c * x.element_size())   # 8d: a whole 'none' site
    return saved


def bn_fwd_train_dual(x, x2, saved2, gamma, beta, running_mean, running_var, num_batches_tracked, eps, momentum):
    """y = relu(bn(x) + x2*saved2.scale + saved2.shift).  Returns (y, saved, mask).  Check
    ``bn_dual_supported`` first; unsupported shapes raise."""
    _need_cuda(x, x2, saved2, gamma, beta)
    m, c = _rows_channels(x)
    if x2.shape != x.shape or x2.dtype != x.dtype:
        raise B200Error("x2 must match x")
    _rows_channels(x2)
    lib = load()
    dt = dtype_code(x)
    y = torch.empty_like(x)
    saved = torch.empty(4, c, dtype=torch.float32, device=x.device)
    mask = torch.empty(m * (c // lib.b200_bn_vec(dt)) + 16, dtype=torch.uint8, device=x.device)
    ws = _bn_workspace(x, m, c, dt)
    s = stats.begin("bn_fwd")
    rc = lib.b200_bn_fwd_train_dual(_ptr(x), _ptr(x2), _ptr(y), _ptr(mask), _ptr(gamma), _ptr(beta), _ptr(running_mean),
                                    _ptr(running_var), _ptr(num_batches_tracked), _ptr(saved), _ptr(saved2), _ptr(ws),
                                    ws.numel(), m, c, eps, momentum, dt, _stream())
    _check(rc, "b200_bn_fwd_train_dual")
    stats.end("bn_fwd", 3, s, 4 * m * c * x.element_size() + mask.numel() - 16, 4 * m * c * x.element_size())
    return y, saved, mask


def bn_bwd_dual(dy, dy2, x, x2, mask, gamma, gamma2, saved, saved2, dgamma, dbeta, dgamma2, dbeta2):
    """Backward of ``bn_fwd_train_dual``.  Returns (dx, dx2)."""
    _need_cuda(dy, dy2, x, x2, mask, gamma, gamma2, saved, saved2)
    m, c = _rows_channels(x)
    _rows_channels(dy)
    lib = load()
    dt = dtype_code(x)
    dx, dx2 = torch.empty_like(x), torch.empty_like(x2)
    ws = _bn_workspace(x, m, c, dt)
    s = stats.begin("bn_bwd")
    rc = lib.b200_bn_bwd_dual(_ptr(dy), _ptr(dy2), _ptr(x), _ptr(x2), _ptr(mask), _ptr(gamma), _ptr(gamma2), _ptr(saved),
                              _ptr(saved2), _ptr(dx), _ptr(dx2), _ptr(dgamma), _ptr(dbeta), _ptr(dgamma2), _ptr(dbeta2),
                              _ptr(ws), ws.numel(), m, c, dt, _stream())
    _check(rc, "b200_bn_bwd_dual")
    e = m * c * x.element_size()
    stats.end("bn_bwd", 3, s, (8 + (2 if dy2 is not None else 0)) * e + 2 * (mask.numel() - 16), 11 * e)  # 8d: 5E + 6E
    return dx, dx2


# ---- K5 -------------------------------------------------------------------------------------------
def pack_grads(tensors, offsets, dst, scale=1.0):
    """Cast-copy each tensor (all of one dtype, each dense in memory) to ``dst[offset : offset+numel]``
    (float32 flat buffer) in one launch per 64 tensors."""
    if not tensors:
        return
    _need_cuda(dst, *tensors)
    n = len(tensors)
    dt = dtype_code(tensors[0])
    srcs = (_vp * n)(*[t.data_ptr() for t in tensors])
    offs = (_i64 * n)(*[int(o) for o in offsets])
    cnts = (_i64 * n)(*[t.numel() for t in tensors])
    s = stats.begin("pack")
    _check(load().b200_pack_grads(srcs, offs, cnts, n, _ptr(dst), scale, dt, _stream()), "b200_pack_grads")
    stats.end("pack", (n + 63) // 64, s, sum(t.numel() * (t.element_size() + 4) for t in tensors))


# ---- K6 -------------------------------------------------------------------------------------------
def maxpool_fwd(x):
    """3x3/2/1 max-pool of a channels-last activation.  Returns (y, argmax uint8 [N,OH,OW,C])."""
    _need_cuda(x)
    if x.dim() != 4 or not x.is_contiguous(memory_format=torch.channels_last):
        raise B200Error("max-pool kernel needs a channels-last 4-d activation")
    n, c, h, w = x.shape
    oh, ow = (h - 1) // 2 + 1, (w - 1) // 2 + 1
    y = torch.empty((n, c, oh, ow), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
    arg = torch.empty((n, oh, ow, c), dtype=torch.uint8, device=x.device)
    s = stats.begin("maxpool_fwd")
    _check(load().b200_maxpool3x3s2_fwd(_ptr(x), _ptr(y), _ptr(arg), n, c, h, w, dtype_code(x), _stream()),
           "b200_maxpool3x3s2_fwd")
    stats.end("maxpool_fwd", 1, s, x.numel() * x.element_size() + y.numel() * (x.element_size() + 1))
    return y, arg


def bn_relu_maxpool_fwd(x, saved):
    """max-pool(relu(x * scale + shift)) with ``saved`` = float32 [4,C] of ``bn_stats`` / ``bn_fwd_train`` (rows 2, 3 =
    scale, shift): the stem's BatchNorm apply + ReLU + pool in one pass, bit-identical to the two kernels run one after
    the other, without the normalised activation ever touching memory.  Returns (y, argmax)."""
    _need_cuda(x, saved)
    if x.dim() != 4 or not x.is_contiguous(memory_format=torch.channels_last):
        raise B200Error("max-pool kernel needs a channels-last 4-d activation")
    n, c, h, w = x.shape
    if saved.dtype != torch.float32 or tuple(saved.shape) != (4, c) or not saved.is_contiguous():
        raise B200Error("saved must be the float32 [4,C] statistics tensor of the BatchNorm forward")
    oh, ow = (h - 1) // 2 + 1, (w - 1) // 2 + 1
    y = torch.empty((n, c, oh, ow), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
    arg = torch.empty((n, oh, ow, c), dtype=torch.uint8, device=x.device)
    s = stats.begin("bn_relu_maxpool_fwd")
    _check(load().b200_bn_relu_maxpool3x3s2_fwd(_ptr(x), saved.data_ptr() + 2 * c * 4, _ptr(y), _ptr(arg), n, c, h, w,
                                                dtype_code(x), _stream()), "b200_bn_relu_maxpool3x3s2_fwd")
    nbytes = x.numel() * x.element_size() + y.numel() * (x.element_size() + 1)
    # SURVEY 8d counts this work as the apply pass of a bn+relu site (2 E*s) plus the pool (x + y + argmax)
    stats.end("bn_relu_maxpool_fwd", 1, s, nbytes, nbytes + 2 * x.numel() * x.element_size())
    return y, arg


def maxpool_bwd(dy, arg, in_shape, dy2=None):
    """dx of the 3x3/2/1 max-pool; ``dy2`` (same shape as dy) is a second gradient branch summed on the fly."""
    _need_cuda(dy, arg, dy2)
    n, c, h, w = in_shape
    if not dy.is_contiguous(memory_format=torch.channels_last):
        raise B200Error("max-pool kernel needs a channels-last gradient")
    if dy2 is not None and (dy2.shape != dy.shape or dy2.dtype != dy.dtype
                            or not dy2.is_contiguous(memory_format=torch.channels_last)):
        raise B200Error("max-pool kernel: the second gradient must match the first (shape, dtype, channels-last)")
    dx = torch.empty((n, c, h, w), dtype=dy.dtype, device=dy.device, memory_format=torch.channels_last)
    s = stats.begin("maxpool_bwd")
    _check(load().b200_maxpool3x3s2_bwd(_ptr(dy), _ptr(dy2), _ptr(arg), _ptr(dx), n, c, h, w, dtype_code(dy), _stream()),
           "b200_maxpool3x3s2_bwd")
    stats.end("maxpool_bwd", 1, s, dx.numel() * dx.element_size() + dy.numel() * (dy.element_size() + 1)
              + (dy2.numel() * dy2.element_size() if dy2 is not None else 0))
    return dx


# ---- K4c ------------------------------------------------------------------------------------------
def peer_pad_words():
    return int(load().b200_peer_pad_bytes()) // 4


def peer_table(ptrs):
    """Host array of device pointers (one per rank) in the form the peer entry points take."""
    return (_vp * len(ptrs))(*[int(p) for p in ptrs])


def peer_exchange_sgd(grad_tab, param_tab, shadow_tab, pad_tab, momentum, rank, world, bucket, lo, hi, lr,
                      weight_decay, momentum_coef, grad_scale, first_step, max_ctas=32):
    _need_cuda(momentum)
    s = stats.begin("peer_exchange")
    _check(load().b200_peer_exchange_sgd(grad_tab, param_tab, shadow_tab, pad_tab, _ptr(momentum), rank, world, bucket,
                                         lo, hi, lr, weight_decay, momentum_coef, grad_scale, int(bool(first_step)),
                                         max_ctas, _stream()), "b200_peer_exchange_sgd")
    # per rank: reads world gradient shards + p + buf, writes buf + world x (p + shadow)
    n = (hi - lo) // world
    stats.end("peer_exchange", 1, s, n * (4 * world + 8 + 4 + world * (4 + (2 if shadow_tab is not None else 0))))


def peer_bcast(buf_tab, pad_tab, nbytes, rank, world, root=0):
    s = stats.begin("peer_bcast")
    _check(load().b200_peer_bcast(buf_tab, pad_tab, nbytes, rank, world, root, _stream()), "b200_peer_bcast")
    stats.end("peer_bcast", 1, s, 0 if rank == root else 2 * nbytes)


def peer_wait(pad, world, n_buckets, rank, bcast_readers_of_me=False):
    _need_cuda(pad)
    s = stats.begin("peer_wait")
    _check(load().b200_peer_wait(_ptr(pad), world, n_buckets, rank, int(bool(bcast_readers_of_me)), _stream()),
           "b200_peer_wait")
    stats.end("peer_wait", 1, s, 0)


# ---- K1 -------------------------------------------------------------------------------------------
def input_stage(images, boxes, flips, out_h, out_w, mean, std, out_dtype=torch.bfloat16, layout=LAYOUT_NHWC):
    """images uint8 [N,Hs,Ws,3] (device), boxes int32 [N,4] (device), flips uint8 [N] (device) or
    None.  Returns [N,3,out_h,out_w] logical NCHW: channels-last strides for LAYOUT_NHWC,
    contiguous for LAYOUT_NCHW; for LAYOUT_S2D_STEM a channels-last [N,16,(out_h+6)/2,(out_w+6)/2]
    tensor tagged ``_b200_s2d_stem`` (see include/b200_imagenet.h)."""
    _need_cuda(images, boxes, flips)
    assert images.dtype == torch.uint8 and images.dim() == 4 and images.shape[3] == 3 and images.is_contiguous()
    assert boxes.dtype == torch.int32 and boxes.is_contiguous() and boxes.shape == (images.shape[0], 4)
    if flips is not None:
        assert flips.dtype == torch.uint8 and flips.is_contiguous() and flips.numel() == images.shape[0]
    n, sh, sw, _ = images.shape
    lib = load()
    if layout == LAYOUT_NHWC:
        out = torch.empty((n, 3, out_h, out_w), dtype=out_dtype, device=images.device,
                          memory_format=torch.channels_last)
    elif layout == LAYOUT_S2D_STEM:
        if out_h % 2 or out_w % 2:
            raise B200Error("the space-to-depth stem layout needs even output sizes")
        out = torch.empty((n, 16, (out_h + 6) // 2, (out_w + 6) // 2), dtype=out_dtype, device=images.device,
                          memory_format=torch.channels_last)
        out._b200_s2d_stem = (out_h, out_w)  # tells B200ResNet to run the rearranged stem convolution
    else:
        out = torch.empty((n, 3, out_h, out_w), dtype=out_dtype, device=images.device)
    nbytes = lib.b200_input_workspace_bytes(n, sh, sw, out_h, out_w)
    if nbytes == 0:
        raise B200Error("invalid input-stage geometry")
    ws = workspace(images.device, nbytes, "input")
    mean3 = (_f32 * 3)(*[float(v) for v in mean])
    std3 = (_f32 * 3)(*[float(v) for v in std])
    s = stats.begin("input")
    _nvtx.push("b200.input")
    rc = (lib.b200_input_crop_resize_flip_norm(_ptr(images), _ptr(boxes), _ptr(flips), _ptr(out), _ptr(ws), ws.numel(),
                                                n, sh, sw, out_h, out_w, mean3, std3,
                                                F32 if out_dtype == torch.float32 else BF16, layout, _stream()))
    _nvtx.pop()
    _check(rc, "b200_input_crop_resize_flip_norm")
    stats.end("input", 2, s, n * 3 * out_h * out_w * out.element_size())  # + box bytes, added by the caller
    return out


def _input_out(n, out_h, out_w, out_dtype, layout, device):
    if layout == LAYOUT_NHWC:
        return torch.empty((n, 3, out_h, out_w), dtype=out_dtype, device=device, memory_format=torch.channels_last)
    if layout == LAYOUT_S2D_STEM:
        if out_h % 2 or out_w % 2:
            raise B200Error("the space-to-depth stem layout needs even output sizes")
        out = torch.empty((n, 16, (out_h + 6) // 2, (out_w + 6) // 2), dtype=out_dtype, device=device,
                          memory_format=torch.channels_last)
        out._b200_s2d_stem = (out_h, out_w)
        return out
    return torch.empty((n, 3, out_h, out_w), dtype=out_dtype, device=device)


def input_stage_ragged(slab, table, boxes, flips, max_h, max_w, out_h, out_w, mean, std, out_dtype=torch.bfloat16,
                       layout=LAYOUT_NHWC):
    """The input stage on a batch of images of different sizes: ``slab`` uint8 [bytes] (device) holds the interleaved
    RGB images, ``table`` int64 [N,4] (device) = (byte offset, height, width, row pitch) per image, ``boxes`` int32
    [N,4] each inside its image.  ``max_h`` / ``max_w`` bound the image sizes of the batch.  Output as ``input_stage``."""
    _need_cuda(slab, table, boxes, flips)
    n = table.shape[0]
    assert slab.dtype == torch.uint8 and slab.dim() == 1 and slab.is_contiguous()
    assert table.dtype == torch.int64 and table.shape == (n, 4) and table.is_contiguous()
    assert boxes.dtype == torch.int32 and boxes.is_contiguous() and boxes.shape == (n, 4)
    if flips is not None:
        assert flips.dtype == torch.uint8 and flips.is_contiguous() and flips.numel() == n
    lib = load()
    out = _input_out(n, out_h, out_w, out_dtype, layout, slab.device)
    nbytes = lib.b200_input_workspace_bytes(n, max_h, max_w, out_h, out_w)
    if nbytes == 0:
        raise B200Error("invalid input-stage geometry")
    ws = workspace(slab.device, nbytes, "input")
    mean3 = (_f32 * 3)(*[float(v) for v in mean])
    std3 = (_f32 * 3)(*[float(v) for v in std])
    s = stats.begin("input")
    _check(lib.b200_input_crop_resize_flip_norm_ragged(_ptr(slab), slab.numel(), _ptr(table), _ptr(boxes), _ptr(flips),
                                                       _ptr(out), _ptr(ws), ws.numel(), n, max_h, max_w, out_h, out_w,
                                                       mean3, std3, F32 if out_dtype == torch.float32 else BF16, layout,
                                                       _stream()), "b200_input_crop_resize_flip_norm_ragged")
    stats.end("input", 2, s, n * 3 * out_h * out_w * out.element_size())
    return out


def interleave_planar(table, slab, max_pixels):
    """Re-pack a batch of planar [3,h,w] uint8 images (what nvJPEG returns) into the interleaved ragged ``slab``;
    ``table`` int64 [N,4] (device) = (address of the planar image, height, width, byte offset in the slab)."""
    _need_cuda(table, slab)
    assert table.dtype == torch.int64 and table.dim() == 2 and table.shape[1] == 4 and table.is_contiguous()
    assert slab.dtype == torch.uint8 and slab.is_contiguous()
    s = stats.begin("interleave")
    _check(load().b200_interleave_planar_u8(_ptr(table), _ptr(slab), table.shape[0], int(max_pixels), _stream()),
           "b200_interleave_planar_u8")
    stats.end("interleave", 1, s, 0)
