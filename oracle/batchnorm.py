"""Oracle (test infrastructure): BatchNorm2d(+residual add)+ReLU forward/backward on [M, C] rows.

Sites: torchvision ``resnet.py:93-94,147-152,269-270`` (bn+relu), ``:97-103,155-161``
(bn+add+relu), ``:240-243`` (bn only); arithmetic ``torch/nn/modules/batchnorm.py:160-206`` with
the defaults eps=1e-5, momentum=0.1: normalise with the biased batch variance, update
``running_var`` with the unbiased one, ``num_batches_tracked += 1``.

Rows are the flattened N*H*W positions of a channels-last tensor, so ``x`` is ``[M, C]``.
Statistics are accumulated in float64 (the exact answer the float32 kernels are compared with under
a stated tolerance); outputs are returned as float32.
"""
import numpy as np

ACT_NONE, ACT_RELU, ACT_ADD_RELU = 0, 1, 2


def bn_train_fwd_oracle(x, gamma, beta, running_mean, running_var, eps=1e-5, momentum=0.1,
                        act=ACT_NONE, z=None):
    x64 = np.asarray(x, dtype=np.float64)
    m = x64.shape[0]
    mean = x64.mean(axis=0)
    var = x64.var(axis=0)  # biased
    invstd = 1.0 / np.sqrt(var + eps)
    v = (x64 - mean) * invstd * np.asarray(gamma, np.float64) + np.asarray(beta, np.float64)
    if act == ACT_ADD_RELU:
        v = v + np.asarray(z, dtype=np.float64)
    y = np.maximum(v, 0.0) if act != ACT_NONE else v
    unbiased = var * (m / max(m - 1, 1))
    new_rm = (1 - momentum) * np.asarray(running_mean, np.float64) + momentum * mean
    new_rv = (1 - momentum) * np.asarray(running_var, np.float64) + momentum * unbiased
    f = np.float32
    return (y.astype(f), mean.astype(f), invstd.astype(f), new_rm.astype(f), new_rv.astype(f), v > 0)


def bn_eval_fwd_oracle(x, gamma, beta, running_mean, running_var, eps=1e-5, act=ACT_NONE, z=None):
    x64 = np.asarray(x, dtype=np.float64)
    invstd = 1.0 / np.sqrt(np.asarray(running_var, np.float64) + eps)
    v = (x64 - np.asarray(running_mean, np.float64)) * invstd * np.asarray(gamma, np.float64) \
        + np.asarray(beta, np.float64)
    if act == ACT_ADD_RELU:
        v = v + np.asarray(z, dtype=np.float64)
    y = np.maximum(v, 0.0) if act != ACT_NONE else v
    return y.astype(np.float32)


def bn_train_bwd_oracle(dy, x, gamma, mean, invstd, act=ACT_NONE, relu_mask=None):
    """Returns (dx, dz_or_None, dgamma, dbeta).  ``relu_mask`` is the boolean ``pre-activation > 0``
    from the forward (ignored for ACT_NONE).  dz (the residual branch's gradient) equals the masked
    dy."""
    dy64 = np.asarray(dy, dtype=np.float64)
    x64 = np.asarray(x, dtype=np.float64)
    m = x64.shape[0]
    if act != ACT_NONE:
        dy64 = dy64 * relu_mask
    xhat = (x64 - np.asarray(mean, np.float64)) * np.asarray(invstd, np.float64)
    dbeta = dy64.sum(axis=0)
    dgamma = (dy64 * xhat).sum(axis=0)
    dx = np.asarray(gamma, np.float64) * np.asarray(invstd, np.float64) * (dy64 - dbeta / m - xhat * dgamma / m)
    f = np.float32
    dz = dy64.astype(f) if act == ACT_ADD_RELU else None
    return dx.astype(f), dz, dgamma.astype(f), dbeta.astype(f)


def bn_fwd_bwd_chunked_oracle(x, gamma, beta, dy, act=ACT_NONE, z=None, dy2=None, eps=1e-5, chunk_rows=65536):
    """The same arithmetic as ``bn_train_fwd_oracle`` + ``bn_train_bwd_oracle`` for activations too large to hold in
    float64 at once (a BASELINE-size site is 802,816 x 256): every pass walks ``chunk_rows`` rows at a time, all sums
    are float64.  ``dy2``: second gradient branch (the residual fan-in autograd would add first).  Returns a dict with
    mean, invstd (float64), y, dx, dz (float32, dz None unless add+relu), dgamma, dbeta (float64)."""
    m, c = x.shape
    g64, b64 = np.asarray(gamma, np.float64), np.asarray(beta, np.float64)
    chunks = [(r, min(r + chunk_rows, m)) for r in range(0, m, chunk_rows)]
    s1 = np.zeros(c)
    for r0, r1 in chunks:
        s1 += np.asarray(x[r0:r1], np.float64).sum(axis=0)
    mean = s1 / m
    s2 = np.zeros(c)
    for r0, r1 in chunks:
        d = np.asarray(x[r0:r1], np.float64) - mean
        s2 += (d * d).sum(axis=0)
    var = s2 / m
    invstd = 1.0 / np.sqrt(var + eps)
    y = np.empty((m, c), np.float32)
    dbeta, dgamma = np.zeros(c), np.zeros(c)
    for r0, r1 in chunks:
        xhat = (np.asarray(x[r0:r1], np.float64) - mean) * invstd
        v = xhat * g64 + b64
        if act == ACT_ADD_RELU:
            v = v + np.asarray(z[r0:r1], np.float64)
        y[r0:r1] = np.maximum(v, 0.0) if act != ACT_NONE else v
        g = np.asarray(dy[r0:r1], np.float64)
        if dy2 is not None:
            g = g + np.asarray(dy2[r0:r1], np.float64)
        if act != ACT_NONE:
            g = g * (v > 0)
        dbeta += g.sum(axis=0)
        dgamma += (g * xhat).sum(axis=0)
    dx = np.empty((m, c), np.float32)
    dz = np.empty((m, c), np.float32) if act == ACT_ADD_RELU else None
    for r0, r1 in chunks:
        xhat = (np.asarray(x[r0:r1], np.float64) - mean) * invstd
        g = np.asarray(dy[r0:r1], np.float64)
        if dy2 is not None:
            g = g + np.asarray(dy2[r0:r1], np.float64)
        if act != ACT_NONE:
            v = xhat * g64 + b64
            if act == ACT_ADD_RELU:
                v = v + np.asarray(z[r0:r1], np.float64)
            g = g * (v > 0)
        dx[r0:r1] = g64 * invstd * (g - dbeta / m - xhat * dgamma / m)
        if dz is not None:
            dz[r0:r1] = g
    return {"mean": mean, "invstd": invstd, "var_unbiased": var * (m / max(m - 1, 1)), "y": y, "dx": dx, "dz": dz,
            "dgamma": dgamma, "dbeta": dbeta}
