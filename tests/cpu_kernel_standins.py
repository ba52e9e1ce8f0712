"""CPU stand-ins for the C-ABI kernel wrappers, built from the oracle (test infrastructure only).

Used by the ``cpu_kernels`` fixture and by multi-process gloo tests to run the host-side logic of
the package without a GPU.  Signatures match ``_cabi``'s wrappers.
"""
import numpy as np
import torch

from oracle import batchnorm as obn
from oracle import cross_entropy as oce
from oracle import input_pipeline as oin
from oracle import maxpool as omp
from oracle import sgd as osgd


def _rows(x):
    if x.dim() == 4:
        return x.permute(0, 2, 3, 1).reshape(-1, x.shape[1])
    return x


def _like(rows, ref):
    t = torch.from_numpy(np.ascontiguousarray(rows)).to(ref.dtype)
    if ref.dim() == 4:
        n, c, h, w = ref.shape
        return t.view(n, h, w, c).permute(0, 3, 1, 2)
    return t


def bn_fwd_train(x, z, gamma, beta, running_mean, running_var, num_batches_tracked, eps, momentum, act):
    xr = _rows(x).float().numpy()
    zr = None if z is None else _rows(z).float().numpy()
    y, mean, invstd, rm, rv, pos = obn.bn_train_fwd_oracle(xr, gamma.detach().numpy(), beta.detach().numpy(),
                                                           running_mean.numpy(), running_var.numpy(), eps, momentum,
                                                           act, zr)
    running_mean.copy_(torch.from_numpy(rm))
    running_var.copy_(torch.from_numpy(rv))
    if num_batches_tracked is not None:
        num_batches_tracked += 1
    scale = gamma.detach().numpy() * invstd
    saved = torch.from_numpy(np.stack([mean, invstd, scale, beta.detach().numpy() - mean * scale]).astype(np.float32))
    mask = torch.from_numpy(pos) if act == obn.ACT_ADD_RELU else None
    return _like(y, x), saved, mask


def bn_fwd_eval(x, z, gamma, beta, running_mean, running_var, eps, act):
    zr = None if z is None else _rows(z).float().numpy()
    y = obn.bn_eval_fwd_oracle(_rows(x).float().numpy(), gamma.detach().numpy(), beta.detach().numpy(),
                               running_mean.numpy(), running_var.numpy(), eps, act, zr)
    return _like(y, x)


def bn_bwd(dy, x, mask, gamma, saved, act, dgamma=None, dbeta=None, dy2=None):
    if dy2 is not None:
        dy = dy + dy2
    xr = _rows(x).float().numpy()
    mean, invstd, scale, shift = (saved[i].numpy() for i in range(4))
    if act == obn.ACT_RELU:
        pos = (xr.astype(np.float64) * scale + shift) > 0
    elif act == obn.ACT_ADD_RELU:
        pos = mask.numpy()
    else:
        pos = None
    dx, dz, dg, db = obn.bn_train_bwd_oracle(_rows(dy).float().numpy(), xr, gamma.detach().numpy(), mean, invstd, act,
                                             pos)
    dg, db = torch.from_numpy(dg), torch.from_numpy(db)
    if dgamma is not None:
        dgamma.copy_(dg)
        dg = dgamma
    if dbeta is not None:
        dbeta.copy_(db)
        db = dbeta
    return _like(dx, x), (None if dz is None else _like(dz, x)), dg, db


def ce_topk_fwd(logits, target, out3=None):
    loss, t1, t5, _, lse = oce.ce_topk_oracle(logits.detach().float().numpy(), target.numpy())
    res = torch.tensor([loss, t1, t5], dtype=torch.float32)
    if out3 is not None:
        out3.copy_(res)
        res = out3
    return res, torch.from_numpy(lse)


def ce_bwd(logits, target, row_lse, grad_out=None, scale=1.0):
    g = float(scale) * (1.0 if grad_out is None else float(grad_out))
    _, _, _, d, _ = oce.ce_topk_oracle(logits.detach().float().numpy(), target.numpy(), grad_scale=g)
    return torch.from_numpy(d).to(logits.dtype)


def sgd_flat(p, g, buf, shadow, lr, weight_decay, momentum, grad_scale, first_step):
    pn, bn = osgd.sgd_step_oracle(p.numpy(), g.numpy(), buf.numpy(), lr, momentum, weight_decay, first_step,
                                  grad_scale)
    p.copy_(torch.from_numpy(pn))
    buf.copy_(torch.from_numpy(bn))
    if shadow is not None:
        shadow.copy_(p)


def input_stage(images, boxes, flips, out_h, out_w, mean, std, out_dtype=torch.bfloat16, layout=0):
    fl = np.zeros(images.shape[0], dtype=bool) if flips is None else flips.numpy().astype(bool)
    out = oin.input_stage_oracle(images.numpy(), boxes.numpy(), fl, out_h, out_w, mean, std)
    t = torch.from_numpy(out).to(out_dtype).permute(0, 3, 1, 2)  # NHWC memory, NCHW shape
    if layout == 2:
        s2d = torch.from_numpy(oin.space_to_depth_stem_oracle(out)).to(out_dtype).permute(0, 3, 1, 2)
        s2d._b200_s2d_stem = (out_h, out_w)
        return s2d
    return t if layout == 0 else t.contiguous()


def maxpool_fwd(x):
    y, pos = omp.maxpool3x3s2_fwd_oracle(x.permute(0, 2, 3, 1).float().numpy())
    return torch.from_numpy(y).to(x.dtype).permute(0, 3, 1, 2), torch.from_numpy(pos)


def bn_relu_maxpool_fwd(x, saved):
    c = x.shape[1]
    a = torch.relu(torch.addcmul(saved[3].view(1, c, 1, 1), x.float(), saved[2].view(1, c, 1, 1))).to(x.dtype)
    return maxpool_fwd(a)


def maxpool_bwd(dy, arg, in_shape, dy2=None):
    g = dy.permute(0, 2, 3, 1).float().numpy()
    if dy2 is not None:
        g = g + dy2.permute(0, 2, 3, 1).float().numpy()
    dx = omp.maxpool3x3s2_bwd_oracle(g, arg.numpy(), (in_shape[2], in_shape[3]))
    return torch.from_numpy(dx).to(dy.dtype).permute(0, 3, 1, 2)


def pack_grads(tensors, offsets, dst, scale=1.0):
    for t, o in zip(tensors, offsets):
        n = t.numel()
        flat = t.permute(0, 2, 3, 1).reshape(-1) if t.dim() == 4 else t.reshape(-1)  # memory (K,R,S,C) order
        dst[o:o + n].copy_(flat.float() * scale)


def bn_dual_supported(x):
    return True


def bn_stats(x, gamma, beta, running_mean, running_var, num_batches_tracked, eps, momentum):
    _, saved, _ = bn_fwd_train(x, None, gamma, beta, running_mean, running_var, num_batches_tracked, eps, momentum,
                               obn.ACT_NONE)
    return saved


def bn_fwd_train_dual(x, x2, saved2, gamma, beta, running_mean, running_var, num_batches_tracked, eps, momentum):
    z = (x2.double() * saved2[2].double().view(1, -1, 1, 1) + saved2[3].double().view(1, -1, 1, 1)).to(x2.dtype)
    return bn_fwd_train(x, z, gamma, beta, running_mean, running_var, num_batches_tracked, eps, momentum,
                        obn.ACT_ADD_RELU)


def bn_bwd_dual(dy, dy2, x, x2, mask, gamma, gamma2, saved, saved2, dgamma, dbeta, dgamma2, dbeta2):
    dx, dz, _, _ = bn_bwd(dy, x, mask, gamma, saved, obn.ACT_ADD_RELU, dgamma, dbeta, dy2)
    dx2, _, _, _ = bn_bwd(dz, x2, None, gamma2, saved2, obn.ACT_NONE, dgamma2, dbeta2)
    return dx, dx2
