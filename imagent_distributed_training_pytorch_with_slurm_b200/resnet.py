"""torchvision-compatible ResNets executed as one hand-scheduled forward/backward.

Replaces ``models.resnet18(pretrained=False, num_classes=...)`` (imagenet.py:312) and what
``loss.backward()`` (imagenet.py:128) does to it.  Same constructor RNG consumption, same
``state_dict`` keys / shapes / dtypes as the torchvision module (resnet.py:166-285), but:

* parameters, their gradients and the momentum buffers live in ONE flat float32 arena each
  (``ParamArena``); the ``nn.Parameter`` objects are views into it, ``p.grad`` is a permanent view of
  the gradient arena.  Convolution weights are stored KRSC (channels-last) so cuDNN needs no
  per-call transposes; a bf16 shadow arena with the same layout feeds the bf16 convolutions and is
  refreshed by the fused SGD kernel;
* the network is a single autograd node: the forward records a tape, the backward walks it in
  reverse calling cuDNN (dgrad/wgrad), the B200 BatchNorm(+add)+ReLU kernels and cuBLAS directly,
  writing every parameter gradient straight into the gradient arena and telling the data-parallel
  wrapper how far it has got so bucket all-reduces overlap the rest of the backward;
* activations are channels-last, bf16 (default) or float32;
* convolution weight gradients (needed only at the next gradient flush) are computed on a second CUDA
  stream, concurrently with the data-gradient -> BatchNorm-backward chain that the rest of the backward
  is waiting for (``use_wgrad_stream``; +4 % on ResNet-50).

BatchNorm sites: ``bn+relu`` resnet.py:93-94,147-152,269-270; ``bn+add+relu`` :97-103,155-161;
``bn`` (downsample) :240-243.  Convolutions and the final Linear stay on cuDNN / cuBLAS.
"""
import math

import torch
import torch.nn as nn

from . import _cabi, _nvtx

_ALIGN = 32  # floats: every parameter starts on a 128-byte boundary of the arena
# resolved overloads: calling the packet (torch.ops.aten.convolution) re-resolves the overload on every call
_conv = torch.ops.aten.convolution.default
_conv_bwd = torch.ops.aten.convolution_backward.default


class ParamArena:
    """Flat storage for parameters / gradients / momentum (float32) and the bf16 shadow."""

    def __init__(self, shapes, device, shadow, symmetric=False):
        self.offsets = []
        off = 0
        for shp in shapes:
            self.offsets.append(off)
            n = int(math.prod(shp))
            off += (n + _ALIGN - 1) // _ALIGN * _ALIGN
        self.numel = off
        self.shapes = list(shapes)
        self.device = device
        # symmetric=True: parameters, gradients and shadow come from torch's symmetric-memory allocator, so that the
        # ranks of one box can map each other's arenas (FlatDDP(exchange="peer")); momentum stays private.
        self.symmetric = bool(symmetric)
        if symmetric:
            import torch.distributed._symmetric_memory as symm_mem

            def zeros(n, dtype):
                return symm_mem.empty(n, dtype=dtype, device=device).zero_()
        else:
            def zeros(n, dtype):
                return torch.zeros(n, dtype=dtype, device=device)
        self.params = zeros(off, torch.float32)
        self.grads = zeros(off, torch.float32)
        self.momentum = torch.zeros(off, dtype=torch.float32, device=device)
        self.shadow = zeros(off, torch.bfloat16) if shadow else None
        self.steps = 0
        self.applied_step = -1   # FlatDDP(exchange="peer") sets this to `steps` when the update ran inside the backward
        # True between a backward and the next zero_grad()/step(): the backward OVERWRITES the gradient arena, so a
        # second backward in that window would silently drop the first one's gradients where autograd accumulates
        self.grads_pending = False

    def view(self, flat, i):
        shp = self.shapes[i]
        seg = flat[self.offsets[i]: self.offsets[i] + int(math.prod(shp))]
        if len(shp) == 4:  # [K,C,R,S] stored as K,R,S,C
            k, c, r, s = shp
            return seg.view(k, r, s, c).permute(0, 3, 1, 2)
        return seg.view(*shp)

    def refresh_shadow(self):
        if self.shadow is not None:
            self.shadow.copy_(self.params)


class _Conv(nn.Module):
    def __init__(self, tv):
        super().__init__()
        self.stride, self.padding, self.dilation, self.groups = tv.stride, tv.padding, tv.dilation, tv.groups
        self.weight = None  # set by _adopt
        self.w_compute = None
        self.grad_view = None
        self.side = None       # [cuda stream or None] shared with the network: where weight gradients are computed
        self.grad_offset = 0   # element offset of this weight's slot in the gradient arena
        self.pending = None    # list shared with the network: (wgrad tensor, arena offset) awaiting one pack launch

    def fwd(self, x):
        return _conv(x, self.w_compute, None, self.stride, self.padding, self.dilation, False, (0, 0), self.groups)

    def _cb(self, gy, x, mask):
        return _conv_bwd(gy, x, self.w_compute, None, self.stride, self.padding, self.dilation, False, (0, 0),
                         self.groups, mask)

    def bwd(self, gy, x, need_gx=True):
        side = self.side[0] if self.side is not None else None
        if side is not None and need_gx and gy.is_cuda:
            # weight gradient on the side stream (nothing downstream needs it before the bucket flush), data
            # gradient on the main stream where the rest of the backward is waiting for it
            main = torch.cuda.current_stream()
            side.wait_stream(main)
            with torch.cuda.stream(side):
                _, gw, _ = self._cb(gy, x, (False, True, False))
            gy.record_stream(side)
            x.record_stream(side)
            self.side[1] = True  # work pending on the side stream: the next flush has to join it
            gx = self._cb(gy, x, (True, False, False))[0]
        else:
            gx, gw, _ = self._cb(gy, x, (need_gx, True, False))
        dense = gw.is_contiguous(memory_format=torch.channels_last) if gw.dim() == 4 else gw.is_contiguous()
        if self.pending is not None and dense:
            self.pending.append((gw, self.grad_offset))  # K,R,S,C memory order == the arena slot's order
        else:
            self.grad_view.copy_(gw)
        return gx


class _StemS2D:
    """The stem convolution (7x7, stride 2, padding 3, 3->64; torchvision resnet.py:197) on an input
    that the input kernel already delivered zero-padded and 2x2 space-to-depth'ed ([N,16,H/2+3,W/2+3]):
    with the weights rearranged the same way it is a 4x4, stride-1, unpadded convolution over 16
    channels - identical sums, ~2.3x faster in cuDNN than the 3-channel strided form (measured
    1.60+0.91 ms -> 0.52+0.59 ms fwd+wgrad at N=256).  The rearranged weights are rebuilt from the
    (shadow) parameter once per optimizer step; the weight gradient is mapped back to [64,3,7,7]."""

    def __init__(self, conv, arena):
        if (conv.weight.shape[1:] != (3, 7, 7) or tuple(conv.stride) != (2, 2) or tuple(conv.padding) != (3, 3)
                or tuple(conv.dilation) != (1, 1) or conv.groups != 1):
            raise ValueError("space-to-depth stem needs the torchvision 7x7/2/3 RGB stem")
        self.conv, self.arena, self._w, self._at = conv, arena, None, -1
        self._flat, self._idx = None, None

    def _gather_index(self, k, device):
        """int64 [K,16,4,4] (channels-last): for every element of the rearranged weight, its position in the flat
        [K,3,7,7] weight, or K*147 (a slot that holds zero) for the padded taps and the 4 spare channels."""
        zero = k * 147
        src = torch.arange(zero, dtype=torch.int64).reshape(k, 3, 7, 7)
        w8 = torch.nn.functional.pad(src, (0, 1, 0, 1), value=zero)                 # [K,3,8,8]
        w = w8.reshape(k, 3, 4, 2, 4, 2).permute(0, 3, 5, 1, 2, 4)                 # K, dy, dx, ch, R, S
        w = torch.nn.functional.pad(w.reshape(k, 12, 4, 4), (0, 0, 0, 0, 0, 4), value=zero)
        return w.contiguous(memory_format=torch.channels_last).to(device)

    def weight(self):
        if self._at != self.arena.steps or self._w is None:
            w7 = self.conv.w_compute
            k = w7.shape[0]
            if self._idx is None or self._idx.device != w7.device or self._flat.dtype != w7.dtype:
                self._idx = self._gather_index(k, w7.device)
                self._flat = torch.zeros(k * 147 + 1, dtype=w7.dtype, device=w7.device)   # last slot stays zero
            self._flat[:-1].view(k, 3, 7, 7).copy_(w7)                                   # 1 kernel
            self._w, self._at = self._flat[self._idx], self.arena.steps                   # 1 kernel (gather)
        return self._w

    def fwd(self, x):
        return _conv(x, self.weight(), None, (1, 1), (0, 0), (1, 1), False, (0, 0), 1)

    def bwd(self, gy, x):
        _, gw, _ = _conv_bwd(gy, x, self.weight(), None, (1, 1), (0, 0), (1, 1), False, (0, 0), 1,
                             (False, True, False))
        k = gw.shape[0]
        g = gw[:, :12].reshape(k, 2, 2, 3, 4, 4).permute(0, 3, 4, 1, 5, 2).reshape(k, 3, 8, 8)
        self.conv.grad_view.copy_(g[:, :, :7, :7])


class _BN(nn.Module):
    def __init__(self, tv):
        super().__init__()
        self.eps, self.momentum, self.num_features = tv.eps, tv.momentum, tv.num_features
        self.weight = None
        self.bias = None
        self.g_weight = None
        self.g_bias = None

    def fwd_train(self, x, act, z=None):
        return _cabi.bn_fwd_train(x, z, self.weight, self.bias, self.running_mean, self.running_var,
                                  self.num_batches_tracked, self.eps, self.momentum, act)

    def fwd_eval(self, x, act, z=None):
        return _cabi.bn_fwd_eval(x, z, self.weight, self.bias, self.running_mean, self.running_var, self.eps, act)

    def bwd(self, dy, x, saved, mask, act, dy2=None):
        dx, dz, _, _ = _cabi.bn_bwd(dy, x, mask, self.weight, saved, act, self.g_weight, self.g_bias, dy2)
        return dx, dz


class _Linear(nn.Module):
    def __init__(self, tv):
        super().__init__()
        self.in_features, self.out_features = tv.in_features, tv.out_features
        self.weight = None
        self.bias = None


class _Block(nn.Module):
    """BasicBlock (resnet.py:59-105) or Bottleneck (resnet.py:108-163)."""

    def __init__(self, tv):
        super().__init__()
        self.n_conv = 3 if hasattr(tv, "conv3") else 2
        for i in range(1, self.n_conv + 1):
            setattr(self, f"conv{i}", _Conv(getattr(tv, f"conv{i}")))
            setattr(self, f"bn{i}", _BN(getattr(tv, f"bn{i}")))
        if tv.downsample is not None:
            self.downsample = nn.Sequential(_Conv(tv.downsample[0]), _BN(tv.downsample[1]))
        else:
            self.downsample = None

    def convs(self):
        return [(getattr(self, f"conv{i}"), getattr(self, f"bn{i}")) for i in range(1, self.n_conv + 1)]

    def fwd_train(self, x):
        tape = {"in": x, "x": [], "a": [], "saved": []}
        cur = x
        pairs = self.convs()
        for conv, bn in pairs[:-1]:
            t = conv.fwd(cur)
            cur, saved, _ = bn.fwd_train(t, _cabi.ACT_RELU)
            tape["x"].append(t)
            tape["a"].append(cur)
            tape["saved"].append(saved)
        conv, bn = pairs[-1]
        t = conv.fwd(cur)
        fused = None
        if self.downsample is not None:
            dbn = self.downsample[1]
            # (Running this branch on the side stream beside the main branch was tried: 19.13-19.15 ms per ResNet-50
            # step with it against 19.15 without, same box - both branches are HBM-bound most of the time.)
            td = self.downsample[0].fwd(x)
            tape["td"] = td
            if _cabi.bn_dual_supported(t):
                # "dual" form: the downsample BatchNorm is only reduced to {scale, shift}; its normalised output
                # is never written - the block's last kernel applies both BatchNorms, the add and the ReLU
                saved_d = _cabi.bn_stats(td, dbn.weight, dbn.bias, dbn.running_mean, dbn.running_var,
                                         dbn.num_batches_tracked, dbn.eps, dbn.momentum)
                fused = _cabi.bn_fwd_train_dual(t, td, saved_d, bn.weight, bn.bias, bn.running_mean, bn.running_var,
                                                bn.num_batches_tracked, bn.eps, bn.momentum)
            else:
                zd, saved_d, _ = dbn.fwd_train(td, _cabi.ACT_NONE)
            tape["saved_d"], tape["dual"] = saved_d, fused is not None
        else:
            zd = x
        out, saved, mask = fused if fused is not None else bn.fwd_train(t, _cabi.ACT_ADD_RELU, zd)
        tape["x"].append(t)
        tape["saved"].append(saved)
        tape["mask"] = mask
        return out, tape

    def fwd_eval(self, x):
        cur = x
        pairs = self.convs()
        for conv, bn in pairs[:-1]:
            cur = bn.fwd_eval(conv.fwd(cur), _cabi.ACT_RELU)
        conv, bn = pairs[-1]
        t = conv.fwd(cur)
        zd = x if self.downsample is None else self.downsample[1].fwd_eval(self.downsample[0].fwd(x), _cabi.ACT_NONE)
        return bn.fwd_eval(t, _cabi.ACT_ADD_RELU, zd)

    def bwd(self, g_out, tape, need_gx=True):
        """``g_out``: gradient of the block output, either one tensor or the pair (main branch,
        identity branch) left un-added by the block after this one - the BatchNorm backward kernels sum
        the pair on the fly.  Returns such a pair for the block before."""
        pairs = self.convs()
        conv, bn = pairs[-1]
        g_out, g_out2 = g_out if isinstance(g_out, tuple) else (g_out, None)
        g_td = None
        if tape.get("dual"):
            dbn = self.downsample[1]
            g, g_td = _cabi.bn_bwd_dual(g_out, g_out2, tape["x"].pop(), tape["td"], tape.pop("mask"), bn.weight, dbn.weight,
                                        tape["saved"].pop(), tape["saved_d"], bn.g_weight, bn.g_bias, dbn.g_weight,
                                        dbn.g_bias)
            g_z = None
        else:
            g, g_z = bn.bwd(g_out, tape["x"].pop(), tape["saved"].pop(), tape.pop("mask"), _cabi.ACT_ADD_RELU, g_out2)
        del g_out, g_out2
        for conv_prev, bn_prev in reversed(pairs[:-1]):
            g = conv.bwd(g, tape["a"].pop())
            g, _ = bn_prev.bwd(g, tape["x"].pop(), tape["saved"].pop(), None, _cabi.ACT_RELU)
            conv = conv_prev
        x_in = tape.pop("in")
        g = conv.bwd(g, x_in, need_gx)
        if self.downsample is not None:
            if g_td is None:
                g_td, _ = self.downsample[1].bwd(g_z, tape["td"], tape["saved_d"], None, _cabi.ACT_NONE)
            del g_z
            tape.pop("td"), tape.pop("saved_d")
            g2 = self.downsample[0].bwd(g_td, x_in, need_gx)
        else:
            g2 = g_z
        return (g, g2) if need_gx else None


class _WholeNet(torch.autograd.Function):
    """The network as one autograd node (so ``loss.backward()`` drives our own backward pass)."""

    @staticmethod
    def forward(ctx, x, anchor, net):
        logits, tape = net._forward_train(x)
        ctx.net, ctx.tape = net, tape
        return logits

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g_logits):
        net, tape = ctx.net, ctx.tape
        ctx.tape = None
        net._backward(tape, g_logits)
        return None, None, None


class B200ResNet(nn.Module):
    """Drop-in for a torchvision ResNet (``models.<arch>(weights=None, num_classes=...)``)."""

    def __init__(self, arch="resnet18", num_classes=1000, device=None, dtype=torch.bfloat16, wgrad_stream=True,
                 symmetric=False):
        super().__init__()
        import torchvision.models as tvm

        if dtype not in (torch.bfloat16, torch.float32):
            raise ValueError("activation dtype must be torch.bfloat16 or torch.float32")
        ctor = getattr(tvm, arch, None)
        if ctor is None:
            raise ValueError(f"unknown torchvision arch {arch!r}")
        tv = ctor(weights=None, num_classes=num_classes)  # same RNG draws as the reference's constructor
        if not isinstance(tv, tvm.ResNet):
            raise ValueError(f"{arch!r} is not a torchvision ResNet; only that family is implemented")
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device())
        device = torch.device(device)  # kernels raise on non-CUDA tensors: there is no CPU path
        if device.type == "cuda" and device.index is None:
            device = torch.device("cuda", torch.cuda.current_device())
        self.arch, self.act_dtype, self.device = arch, dtype, device

        self.conv1, self.bn1 = _Conv(tv.conv1), _BN(tv.bn1)
        if (tv.maxpool.kernel_size, tv.maxpool.stride, tv.maxpool.padding) != (3, 2, 1):
            raise ValueError("only the torchvision stem max-pool (3x3, stride 2, padding 1) is implemented")
        for li in range(1, 5):
            setattr(self, f"layer{li}", nn.Sequential(*[_Block(b) for b in getattr(tv, f"layer{li}")]))
        self.fc = _Linear(tv.fc)
        self._symmetric = bool(symmetric)   # read by _adopt when it allocates the arenas
        self._adopt(tv)
        if wgrad_stream and device.type == "cuda":
            self.use_wgrad_stream(True)
        self.__dict__["_stem_s2d"] = None
        self._anchor = torch.zeros((), device=device, requires_grad=True)
        # load_state_dict (directly or through the data-parallel wrapper) writes the float32 arena only: bring the
        # bf16 shadow and the rearranged stem weight, which is what the convolutions actually read, up to date
        self.register_load_state_dict_post_hook(lambda module, incompatible: module.weights_changed())
        self.__dict__["_sink"] = None  # data-parallel wrapper (grad_ready / backward_done); not a submodule
        self.grad_predivide = 1.0  # 1/world_size, set by the wrapper (DDP's pre-division)
        self.fuse_stem_pool = True  # training forward: bn1 apply + ReLU + max-pool as one kernel (same bits)

    # ---- storage --------------------------------------------------------------------------------
    def _adopt(self, tv):
        """Move the torchvision module's parameters/buffers into flat arenas, keeping key order."""
        self.__dict__["_pending"] = []
        self.__dict__["_side"] = [None, False]  # [side stream or None, un-joined work pending on it]
        tv_params = dict(tv.named_parameters())
        tv_mods = dict(tv.named_modules())
        names = [n for n, _ in tv.named_parameters()]
        self.arena = ParamArena([tuple(tv_params[n].shape) for n in names], self.device,
                                shadow=self.act_dtype == torch.bfloat16, symmetric=self._symmetric)
        a = self.arena
        self.param_names = names
        self.param_index = {n: i for i, n in enumerate(names)}
        bn_mods = [n for n, m in tv.named_modules() if isinstance(m, nn.BatchNorm2d)]
        nfeat = sum(tv_mods[n].num_features for n in bn_mods)
        # buffer arena: [running_mean/var float32 ... | num_batches_tracked int64 ...] in one byte buffer
        fbytes = 4 * 2 * nfeat
        fbytes = (fbytes + 15) // 16 * 16
        if self._symmetric:
            import torch.distributed._symmetric_memory as symm_mem

            self.buffer_arena = symm_mem.empty(fbytes + 8 * len(bn_mods), dtype=torch.uint8, device=self.device).zero_()
        else:
            self.buffer_arena = torch.zeros(fbytes + 8 * len(bn_mods), dtype=torch.uint8, device=self.device)
        fview = self.buffer_arena[:4 * 2 * nfeat].view(torch.float32)
        iview = self.buffer_arena[fbytes:].view(torch.int64)
        own = dict(self.named_modules())
        boff = 0
        for bi, n in enumerate(bn_mods):
            m, src = own[n], tv_mods[n]
            c = src.num_features
            m.register_buffer("running_mean", fview[boff:boff + c])
            m.register_buffer("running_var", fview[boff + c:boff + 2 * c])
            m.register_buffer("num_batches_tracked", iview[bi])
            m.running_mean.copy_(src.running_mean)
            m.running_var.copy_(src.running_var)
            boff += 2 * c
        with torch.no_grad():
            for i, n in enumerate(names):
                modname, leaf = n.rsplit(".", 1)
                m = own[modname]
                pv = a.view(a.params, i)
                pv.copy_(tv_params[n])
                p = nn.Parameter(pv)
                p._b200_arena, p._b200_index = a, i
                p.grad = a.view(a.grads, i)
                setattr(m, leaf, p)
                if isinstance(m, _Conv):
                    m.grad_view = a.view(a.grads, i)
                    m.grad_offset, m.pending, m.side = a.offsets[i], self._pending, self._side
                    m.w_compute = a.view(a.shadow, i) if a.shadow is not None else p.detach()
                elif isinstance(m, _BN):
                    if leaf == "weight":
                        m.g_weight = a.view(a.grads, i)
                    else:
                        m.g_bias = a.view(a.grads, i)
                elif isinstance(m, _Linear):
                    src = a.shadow if a.shadow is not None else a.params
                    setattr(m, "w_compute" if leaf == "weight" else "b_compute", a.view(src, i))
                    setattr(m, "g_" + leaf, a.view(a.grads, i))
        a.refresh_shadow()
        # arena offsets at which the backward reports progress (start of each block, then the stem)
        self._marks = {}
        for li in range(1, 5):
            for bi, _ in enumerate(getattr(self, f"layer{li}")):
                self._marks[(li, bi)] = a.offsets[self.param_index[f"layer{li}.{bi}.conv1.weight"]]
        self._marks["fc"] = a.offsets[self.param_index["fc.weight"]]

    def weights_changed(self):
        """Call after writing parameters other than through ``load_state_dict`` or the optimizer (e.g. an in-place
        edit of ``p.data``): refreshes the bf16 shadow arena and drops the cached space-to-depth stem weight."""
        self.arena.refresh_shadow()
        if self._stem_s2d is not None:
            self._stem_s2d._at = -1

    def mark_offsets(self):
        """Arena offsets (descending) at which gradients become final during the backward."""
        return sorted(set(self._marks.values()) | {0}, reverse=True)

    def use_wgrad_stream(self, enable=True):
        """Compute convolution weight gradients on a second CUDA stream, concurrently with the
        data-gradient / BatchNorm-backward chain (they are only needed at the next gradient flush)."""
        self._side[0] = torch.cuda.Stream(device=self.device) if enable else None

    def _flush_grads(self):
        """One K5 launch writes every weight gradient produced since the last flush into the arena."""
        if self._side[0] is not None and self._side[1]:
            torch.cuda.current_stream().wait_stream(self._side[0])
            self._side[1] = False
        if self._pending:
            _cabi.pack_grads([t for t, _ in self._pending], [o for _, o in self._pending], self.arena.grads)
            self._pending.clear()

    def _progress(self, offset):
        """Backward has finished everything at arena offsets >= ``offset``."""
        sink = self._sink
        if sink is not None and sink.wants(offset):
            self._flush_grads()
            sink.grad_ready(offset)
        elif offset == 0:
            self._flush_grads()

    def attach_sink(self, sink):
        self.__dict__["_sink"] = sink  # bypass nn.Module registration (the wrapper owns us, not vice versa)

    # ---- execution ------------------------------------------------------------------------------
    def _stem(self, x):
        """(convolution object, input) for the stem: the space-to-depth form when the input kernel
        delivered that layout, the plain 7x7/2 convolution otherwise."""
        if getattr(x, "_b200_s2d_stem", None) is not None:
            if self._stem_s2d is None:
                self.__dict__["_stem_s2d"] = _StemS2D(self.conv1, self.arena)
            return self._stem_s2d
        return self.conv1

    def _prep_input(self, x):
        if getattr(x, "_b200_s2d_stem", None) is not None and x.dtype == self.act_dtype and x.device == self.device:
            return x
        if x.device != self.device:
            raise _cabi.B200Error("input must already be on the model's CUDA device")
        if x.dtype != self.act_dtype or not x.is_contiguous(memory_format=torch.channels_last):
            x = x.to(dtype=self.act_dtype, memory_format=torch.channels_last)
        return x

    def forward(self, x):
        x = self._prep_input(x)
        if self.training and torch.is_grad_enabled():
            return _WholeNet.apply(x, self._anchor, self)
        if self.training:
            return self._forward_train(x)[0]
        return self._forward_eval(x)

    def _forward_train(self, x):
        with _nvtx.range("b200.forward"):
            return self._forward_train_impl(x)

    def _forward_train_impl(self, x):
        stem = self._stem(x)
        tape = {"x0": x, "stem": stem}
        t = stem.fwd(x)
        bn = self.bn1
        if self.fuse_stem_pool:
            # statistics only; the apply + ReLU happens inside the pool kernel and relu(bn1(x)) is never materialised
            # (the backward recomputes the ReLU mask from t and routes the pool's gradient through the argmax)
            saved = _cabi.bn_stats(t, bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.num_batches_tracked,
                                   bn.eps, bn.momentum)
            cur, idx = _cabi.bn_relu_maxpool_fwd(t, saved)
        else:
            a, saved, _ = bn.fwd_train(t, _cabi.ACT_RELU)
            cur, idx = _cabi.maxpool_fwd(a)
            del a
        tape["t1"], tape["saved1"] = t, saved
        tape["pool_idx"], tape["pool_in_shape"] = idx, t.shape
        tape["blocks"] = []
        for li in range(1, 5):
            for blk in getattr(self, f"layer{li}"):
                cur, bt = blk.fwd_train(cur)
                tape["blocks"].append(bt)
        tape["last_shape"] = cur.shape
        feat = cur.mean((2, 3))
        tape["feat"] = feat
        logits = torch.addmm(self.fc.b_compute, feat, self.fc.w_compute.t())
        return logits, tape

    def _forward_eval(self, x):
        a = self.bn1.fwd_eval(self._stem(x).fwd(x), _cabi.ACT_RELU)
        cur, _ = _cabi.maxpool_fwd(a)
        for li in range(1, 5):
            for blk in getattr(self, f"layer{li}"):
                cur = blk.fwd_eval(cur)
        feat = cur.mean((2, 3))
        return torch.addmm(self.fc.b_compute, feat, self.fc.w_compute.t())

    def zero_grad(self, set_to_none=True):
        """``p.grad`` is a permanent view of the gradient arena and every backward overwrites all of it, so there is
        nothing to clear (and the views must not be dropped); this only re-arms the accumulation guard."""
        self.arena.grads_pending = False

    def _backward(self, tape, g_logits):
        with _nvtx.range("b200.backward"):
            return self._backward_impl(tape, g_logits)

    def _backward_impl(self, tape, g_logits):
        if self.arena.grads_pending:
            raise RuntimeError(
                "backward() called twice without optimizer.zero_grad() or optimizer.step() in between: this model "
                "writes gradients in place of the previous ones (the reference clears them before every backward, "
                "imagenet.py:127) and does not accumulate")
        self.arena.grads_pending = True
        sink = self._sink
        fc = self.fc
        g_logits = g_logits.contiguous()
        pre = self.grad_predivide
        feat = tape.pop("feat")
        # Linear backward on cuBLAS; DDP's 1/world pre-division rides on alpha
        g_feat = torch.addmm(feat, g_logits, fc.w_compute, beta=0.0, alpha=pre)
        if g_logits.dtype == torch.float32:
            torch.addmm(fc.g_weight, g_logits.t(), feat, beta=0.0, alpha=pre, out=fc.g_weight)
            torch.sum(g_logits, 0, out=fc.g_bias)
            if pre != 1.0:
                fc.g_bias.mul_(pre)
        else:
            fc.g_weight.copy_(torch.addmm(fc.w_compute, g_logits.t(), feat, beta=0.0, alpha=pre))
            fc.g_bias.copy_(g_logits.sum(0, dtype=torch.float32) * pre)
        self._progress(self._marks["fc"])
        n, c, h, w = tape.pop("last_shape")
        # gradient of the global average pool: one row of C values repeated over the H*W positions (channels-last, so
        # the copy moves whole rows)
        g = (g_feat * (1.0 / (h * w))).view(n, 1, 1, c).expand(n, h, w, c).contiguous().permute(0, 3, 1, 2)
        blocks = tape.pop("blocks")
        for li in range(4, 0, -1):
            layer = getattr(self, f"layer{li}")
            for bi in range(len(layer) - 1, -1, -1):
                g = layer[bi].bwd(g, blocks.pop())
                self._progress(self._marks[(li, bi)])
        # the two gradient branches of the first block meet at the pooled tensor: the pool's backward sums them
        g = _cabi.maxpool_bwd(g[0], tape.pop("pool_idx"), tape.pop("pool_in_shape"), g[1])
        g, _ = self.bn1.bwd(g, tape.pop("t1"), tape.pop("saved1"), None, _cabi.ACT_RELU)
        stem = tape.pop("stem")
        if stem is self.conv1:
            stem.bwd(g, tape.pop("x0"), need_gx=False)
        else:
            stem.bwd(g, tape.pop("x0"))
        self._progress(0)
        self._flush_grads()
        if sink is not None:
            sink.backward_done()


def build_model(arch="resnet18", num_classes=1000, device=None, dtype=torch.bfloat16, symmetric=False):
    """``models.resnet18(pretrained=False, num_classes=num_classes).cuda()`` (imagenet.py:312) with
    ``arch`` generalised; call after ``torch.manual_seed`` like the reference does.  ``symmetric=True`` places the
    arenas in peer-mappable memory (needed by ``FlatDDP(exchange="peer")``)."""
    return B200ResNet(arch, num_classes, device, dtype, symmetric=symmetric)
