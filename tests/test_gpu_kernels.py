"""Parity of every CUDA kernel with the oracle, called through the C ABI (ctypes).  -m gpu only.

Tolerances (stated per test): integer/byte results bit-exact; float32-I/O kernels within a few ulp /
1e-5 relative of the float64-accumulating oracle; bf16-I/O kernels within bf16 rounding of it.
"""
import os

import numpy as np
import pytest
import torch

from oracle import batchnorm as obn
from oracle import cross_entropy as oce
from oracle import input_pipeline as oin
from oracle import maxpool as omp
from oracle import sgd as osgd

import imagent_distributed_training_pytorch_with_slurm_b200._cabi as cabi

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
DEV = "cuda"


def _bf16_round(a):
    return torch.from_numpy(np.asarray(a, dtype=np.float32)).to(torch.bfloat16).float().numpy()


# ------------------------------------------------------------------------------------------------ K4b
@pytest.mark.parametrize("n", [1, 3, 4099, 1 << 20, 25_557_032])
@pytest.mark.parametrize("shadow", [False, True])
def test_sgd_flat_matches_oracle(n, shadow):
    rng = np.random.default_rng(n)
    p0 = rng.standard_normal(n).astype(np.float32)
    p = torch.from_numpy(p0.copy()).to(DEV)
    buf = torch.zeros(n, device=DEV)
    sh = torch.zeros(n, dtype=torch.bfloat16, device=DEV) if shadow else None
    po, bo = p0, None
    for step in range(3):
        g0 = rng.standard_normal(n).astype(np.float32)
        g = torch.from_numpy(g0).to(DEV)
        cabi.sgd_flat(p, g, buf, sh, 0.1, 1e-4, 0.9, 1.0, step == 0)
        prev, prev_b = po, (bo if bo is not None else np.zeros_like(po))
        po, bo = osgd.sgd_step_oracle(po, g0, bo, 0.1, 0.9, 1e-4, first_step=(step == 0))
        big = np.maximum.reduce([np.abs(prev), np.abs(po), np.abs(bo), np.abs(prev_b), np.abs(g0)]).astype(np.float32)
        tol = 3 * np.spacing(big)  # <= 3 ulp of the largest operand (fused vs separately rounded multiply-adds)
        assert (np.abs(p.cpu().numpy() - po) <= tol).all(), step
        assert (np.abs(buf.cpu().numpy() - bo) <= tol).all(), step
        # keep the two trajectories from drifting apart by rounding: continue from the GPU state
        po, bo = p.cpu().numpy(), buf.cpu().numpy()
        if shadow:
            assert torch.equal(sh, p.to(torch.bfloat16))


def test_sgd_golden_trajectory_and_grad_scale():
    c = np.load(os.path.join(GOLD, "sgd_case.npz"))
    p = torch.from_numpy(c["p0"].copy()).to(DEV)
    buf = torch.zeros_like(p)
    for s in range(len(c["grads"])):
        g = torch.from_numpy(c["grads"][s] * 4.0).to(DEV)  # un-averaged sum of 4 ranks ...
        cabi.sgd_flat(p, g, buf, None, 0.1, 1e-4, 0.9, 0.25, s == 0)  # ... scaled back by 1/world
        assert np.allclose(p.cpu().numpy(), c["ps"][s], rtol=0, atol=4e-7 * (s + 1))


# ------------------------------------------------------------------------------------------------ K3
def _ce_case(n, c, seed, dtype):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, c, generator=g) * 3.0
    t = torch.randint(0, c, (n,), generator=g)
    for r in range(0, n, 3):
        x[r, t[r]] += 7.0
    for r in range(1, n, 4):
        x[r, t[r]] += 2.0
    return x.to(dtype), t


@pytest.mark.parametrize("n,c", [(1, 8), (7, 13), (64, 1000), (256, 1000), (33, 1001), (5, 4099)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_ce_topk_matches_oracle(n, c, dtype):
    x, t = _ce_case(n, c, n * 31 + c, dtype)
    xo = x.float().numpy()  # the oracle sees exactly the values the kernel sees
    loss, t1, t5, dl, lse = oce.ce_topk_oracle(xo, t.numpy())
    out3, row_lse = cabi.ce_topk_fwd(x.to(DEV), t.to(DEV))
    out3 = out3.cpu().numpy()
    assert abs(out3[0] - loss) <= 2e-6 * abs(loss)         # float32 accumulation vs float64
    assert out3[1] == t1 and out3[2] == t5                 # counts: bit-exact
    assert np.allclose(row_lse.cpu().numpy(), lse, rtol=2e-6, atol=2e-6)
    gout = torch.tensor(0.5, device=DEV)
    d = cabi.ce_bwd(x.to(DEV), t.to(DEV), row_lse, gout, 2.0).float().cpu().numpy()  # 0.5 * 2.0 = 1
    if dtype == torch.float32:
        assert np.abs(d - dl).max() <= 2e-7
    else:
        assert np.abs(d - dl).max() <= np.abs(dl).max() * 2 ** -8 + 1e-9  # one bf16 rounding


def test_ce_topk_golden_reference_accuracy():
    c = np.load(os.path.join(GOLD, "ce_topk_case.npz"))
    out3, _ = cabi.ce_topk_fwd(torch.from_numpy(c["logits"]).to(DEV), torch.from_numpy(c["target"]).to(DEV))
    out3 = out3.cpu().numpy()
    assert out3[1] == c["prec1"] and out3[2] == c["prec5"]  # the reference's accuracy() output
    assert abs(out3[0] - c["loss"]) <= 2e-6 * c["loss"]


def test_accuracy_for_any_topk_matches_reference_golden_and_oracle():
    """accuracy() takes any tuple of k (imagenet.py:63-79); checked against the reference's own output on the golden
    case (bit-exact) and against the oracle on bf16 logits, through the C ABI (b200_target_rank) and the wrapper."""
    import json

    from oracle import step_helpers as osh
    import imagent_distributed_training_pytorch_with_slurm_b200 as b200

    a = json.load(open(os.path.join(GOLD, "golden.json")))["accuracy_topk"]
    c = np.load(os.path.join(GOLD, a["case"]))
    x, t = torch.from_numpy(c["logits"]).to(DEV), torch.from_numpy(c["target"]).to(DEV)
    got = b200.accuracy(x, t, topk=tuple(a["topk"]))
    assert all(v.shape == (1,) and v.dtype == torch.float32 for v in got)
    assert [v.item() for v in got] == a["prec"]
    rank = cabi.target_rank(x, t).cpu().numpy()
    assert np.array_equal(rank, osh.rank_of_target(c["logits"], c["target"]))
    for n, cc, dtype in ((33, 1001, torch.bfloat16), (5, 4099, torch.float32), (256, 1000, torch.bfloat16)):
        xb, tb = _ce_case(n, cc, 3 * n + cc, dtype)
        ks = (1, 2, 7, 5, cc)
        want = osh.accuracy_oracle(xb.float().numpy(), tb.numpy(), ks)
        assert [v.item() for v in b200.accuracy(xb.to(DEV), tb.to(DEV), topk=ks)] == [float(w) for w in want]
    with pytest.raises(RuntimeError):
        b200.accuracy(x, t, topk=(1, 1001))                 # the reference's topk(maxk) raises here too
    # the (1, 5) result cached by the criterion is dropped when the logits are modified in place afterwards
    crit = b200.CrossEntropyLoss()
    xb, tb = _ce_case(64, 1000, 9, torch.float32)
    xb, tb = xb.to(DEV), tb.to(DEV)
    crit(xb, tb)
    before = [v.item() for v in b200.accuracy(xb, tb, topk=(1, 5))]
    xb.neg_()
    after = [v.item() for v in b200.accuracy(xb, tb, topk=(1, 5))]
    want = osh.accuracy_oracle(xb.cpu().numpy(), tb.cpu().numpy(), (1, 5))
    assert after == [float(w) for w in want] and after != before


def test_ce_topk_ties_and_determinism():
    # all-equal rows: the target ranks behind the classes with a lower index
    x = torch.zeros(6, 16, device=DEV)
    t = torch.tensor([0, 1, 4, 5, 15, 3], device=DEV)
    out3, _ = cabi.ce_topk_fwd(x, t)
    o = out3.cpu().numpy()
    assert o[1] == np.float32(100.0 / 6) * 1 and o[2] == np.float32(np.float32(4) * np.float32(100.0 / 6))
    assert abs(o[0] - np.log(16.0)) < 1e-6
    x, t = _ce_case(256, 1000, 5, torch.bfloat16)
    a = cabi.ce_topk_fwd(x.to(DEV), t.to(DEV))[0].clone()
    for _ in range(5):
        b = cabi.ce_topk_fwd(x.to(DEV), t.to(DEV))[0]
        assert torch.equal(a, b)  # fixed-order reduction: bit-reproducible


# ------------------------------------------------------------------------------------------------ K2
def _rows_t(a, dtype):  # [M, C] numpy -> device tensor
    return torch.from_numpy(np.ascontiguousarray(a)).to(dtype).to(DEV)


def _bn_check(m, c, act, dtype, vec, seed, eps=1e-5, momentum=0.1):
    if dtype == torch.bfloat16:
        cabi.tune_set("bn_vec_bf16", vec)
    rng = np.random.default_rng(seed)
    x = (rng.standard_normal((m, c)) * rng.uniform(0.5, 2.0, c) + rng.uniform(-3, 3, c)).astype(np.float32)
    z = rng.standard_normal((m, c)).astype(np.float32)
    dy = rng.standard_normal((m, c)).astype(np.float32)
    gamma = rng.uniform(0.5, 1.5, c).astype(np.float32)
    beta = (rng.standard_normal(c) * 0.3).astype(np.float32)
    rm0, rv0 = rng.standard_normal(c).astype(np.float32), rng.uniform(0.5, 2, c).astype(np.float32)
    xd, zd, dyd = _rows_t(x, dtype), _rows_t(z, dtype), _rows_t(dy, dtype)
    xs, zs, dys = xd.float().cpu().numpy(), zd.float().cpu().numpy(), dyd.float().cpu().numpy()
    gd, bd = torch.from_numpy(gamma).to(DEV), torch.from_numpy(beta).to(DEV)
    rm, rv = torch.from_numpy(rm0.copy()).to(DEV), torch.from_numpy(rv0.copy()).to(DEV)
    nbt = torch.tensor(5, dtype=torch.int64, device=DEV)
    y, saved, mask = cabi.bn_fwd_train(xd, zd if act == cabi.ACT_ADD_RELU else None, gd, bd, rm, rv, nbt, eps,
                                       momentum, act)
    yo, mean, invstd, rmo, rvo, pos = obn.bn_train_fwd_oracle(xs, gamma, beta, rm0, rv0, eps, momentum, act, zs)
    bf = dtype == torch.bfloat16
    sv = saved.cpu().numpy()
    assert np.allclose(sv[0], mean, rtol=1e-5, atol=1e-5)
    assert np.allclose(sv[1], invstd, rtol=2e-5)
    assert np.allclose(rm.cpu().numpy(), rmo, rtol=1e-5, atol=1e-5)
    assert np.allclose(rv.cpu().numpy(), rvo, rtol=2e-5, atol=1e-6)
    assert int(nbt) == 6
    yk = y.float().cpu().numpy()
    if bf:
        assert np.abs(yk - yo).max() <= np.abs(yo).max() * 2 ** -7
    else:
        assert np.allclose(yk, yo, rtol=2e-5, atol=2e-5)
    # backward, fed with the kernel's own saved statistics.  Elements whose pre-activation is within
    # rounding of zero may take either side of the ReLU (float32 fma in the kernel vs float64 in the
    # oracle); they are excluded from the element-wise checks and bounded in number.
    dx, dz, dg, db = cabi.bn_bwd(dyd, xd, mask, gd, saved, act)
    dxo, dzo, dgo, dbo = obn.bn_train_bwd_oracle(dys, xs, gamma, mean, invstd, act, pos)
    pre = (xs.astype(np.float64) - mean) * invstd * gamma + beta + (zs if act == cabi.ACT_ADD_RELU else 0.0)
    safe = (np.abs(pre) > 1e-4) if act != cabi.ACT_NONE else np.ones_like(pre, dtype=bool)
    assert (~safe).mean() < 1e-3
    scale = np.abs(dgo).max() + np.abs(dbo).max()
    # an excluded element that took the other side of the ReLU moves its channel's sums by its own gradient
    xhat = (xs.astype(np.float64) - mean) * invstd
    slack_b = (np.abs(dys) * ~safe).sum(0)
    slack_g = (np.abs(dys * xhat) * ~safe).sum(0)
    assert np.all(np.abs(db.cpu().numpy() - dbo) <= 1e-4 * np.abs(dbo) + 1e-4 * scale + 1e-3 + slack_b)
    assert np.all(np.abs(dg.cpu().numpy() - dgo) <= 1e-4 * np.abs(dgo) + 1e-4 * scale + 1e-3 + slack_g)
    dxk = dx.float().cpu().numpy()
    if bf:
        assert (np.abs(dxk - dxo)[safe]).max() <= np.abs(dxo).max() * 2 ** -7
    else:
        assert np.allclose(dxk[safe], dxo[safe], rtol=1e-4, atol=1e-5 * (np.abs(dxo).max() + 1))
    if act == cabi.ACT_ADD_RELU:
        assert np.array_equal(dz.float().cpu().numpy()[safe], (dys * pos)[safe])  # masked copy: exact
    # two-branch incoming gradient (residual fan-in): same result as adding the branches first
    half = (dyd.float() * 0.5).to(dtype)
    rest = (dyd.float() - half.float()).to(dtype)
    if torch.equal((half.float() + rest.float()).to(dtype), dyd):
        dx2, dz2, dg2, db2 = cabi.bn_bwd(half, xd, mask, gd, saved, act, dy2=rest)
        if bf:
            assert (dx2.float() - dx.float()).abs().max().item() <= np.abs(dxo).max() * 2 ** -7
        else:
            assert torch.allclose(dx2, dx, rtol=1e-5, atol=1e-6)
        assert torch.allclose(dg2, dg, rtol=1e-4, atol=1e-4 * scale + 1e-3)
    # eval mode with the updated running statistics
    ye = cabi.bn_fwd_eval(xd, zd if act == cabi.ACT_ADD_RELU else None, gd, bd, rm, rv, eps, act)
    yeo = obn.bn_eval_fwd_oracle(xs, gamma, beta, rm.cpu().numpy(), rv.cpu().numpy(), eps, act, zs)
    if bf:
        assert np.abs(ye.float().cpu().numpy() - yeo).max() <= np.abs(yeo).max() * 2 ** -7
    else:
        assert np.allclose(ye.cpu().numpy(), yeo, rtol=2e-5, atol=2e-5)


SHAPES = [(2, 8), (210, 16), (1000, 64), (4099, 24), (777, 256), (300, 2048), (50176, 64), (12544, 512), (3, 40)]


@pytest.mark.parametrize("m,c", SHAPES)
@pytest.mark.parametrize("act", [cabi.ACT_NONE, cabi.ACT_RELU, cabi.ACT_ADD_RELU])
def test_bn_f32_matches_oracle(m, c, act):
    _bn_check(m, c, act, torch.float32, 4, seed=m + c + act)


@pytest.mark.parametrize("m,c", SHAPES)
@pytest.mark.parametrize("act", [cabi.ACT_NONE, cabi.ACT_RELU, cabi.ACT_ADD_RELU])
@pytest.mark.parametrize("vec", [4, 8])
def test_bn_bf16_matches_oracle(m, c, act, vec):
    if c % vec:
        pytest.skip("channel count not a multiple of the vector width")
    try:
        _bn_check(m, c, act, torch.bfloat16, vec, seed=m + c + act)
    finally:
        cabi.tune_set("bn_vec_bf16", 8)


@pytest.mark.parametrize("m,c", [(4096, 64), (12544, 256), (3001, 128), (1568, 2048), (50176, 64)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_bn_single_kernel_form_is_bit_identical_to_the_three_kernel_chain(m, c, dtype):
    """The cooperative one-kernel-per-site form (statistics / fold / apply with grid-wide barriers inside) uses the
    chain's chunking, partial layout and summation order, so every output must have the same bits: forward (none,
    relu, add+relu, statistics only), running statistics, backward with and without a second gradient branch."""
    cabi.tune_set("bn_vec_bf16", 8)
    g = torch.Generator().manual_seed(m + c)
    mk = lambda: (torch.randn(m, c, generator=g) * 1.5 + 0.3).to(dtype).to(DEV).view(1, m, 1, c).permute(0, 3, 1, 2)  # noqa: E731
    x, z, dy, dy2 = mk(), mk(), mk(), mk()
    gamma, beta = torch.rand(c, generator=g).add(0.5).to(DEV), torch.randn(c, generator=g).to(DEV)
    res = []
    for fused in (0, 1):
        cabi.tune_set("bn_fused", fused)
        out = []
        for act, zz in ((cabi.ACT_NONE, None), (cabi.ACT_RELU, None), (cabi.ACT_ADD_RELU, z)):
            rm, rv = torch.zeros(c, device=DEV), torch.ones(c, device=DEV)
            nbt = torch.zeros((), dtype=torch.int64, device=DEV)
            y, saved, mask = cabi.bn_fwd_train(x, zz, gamma, beta, rm, rv, nbt, 1e-5, 0.1, act)
            out += [y, saved, rm, rv, nbt] + ([mask[:-16]] if mask is not None else [])
            for second in (None, dy2):
                dx, dz, dg, db = cabi.bn_bwd(dy, x, mask, gamma, saved, act, dy2=second)
                out += [dx, dg, db] + ([dz] if dz is not None else [])
        rm, rv = torch.zeros(c, device=DEV), torch.ones(c, device=DEV)
        out += [cabi.bn_stats(x, gamma, beta, rm, rv, None, 1e-5, 0.1), rm, rv]
        torch.cuda.synchronize()
        res.append([t.clone() for t in out])
    cabi.tune_set("bn_fused", 0)      # the default (the single-kernel form measured slower, see csrc/bn_fused.cuh)
    assert len(res[0]) == len(res[1])
    for i, (a, b) in enumerate(zip(*res)):
        assert torch.equal(a, b), i


def test_bn_baseline_size_add_relu_with_fan_in_matches_the_chunked_float64_oracle():
    """One BASELINE-size bn+add+relu site (ResNet-50 layer1, N=256: [256,256,56,56] = 802,816 x 256, bf16) - the path
    that carries most of the step's bytes - against the oracle itself, not only its properties: forward output,
    statistics, and the backward with the mask and a second gradient branch (residual fan-in).  Tolerances: statistics
    1e-5 relative (float32 accumulation over 802,816 rows vs float64), outputs one bf16 rounding of the float64 value
    (the gradient sum of two branches is rounded from float32 once) on all but < 1e-5 of the elements (those whose
    pre-activation lies within float32 rounding of the ReLU boundary may fall on either side), d(gamma)/d(beta) 2e-3 of
    the vector's scale."""
    from oracle import batchnorm as obn
    m, c = 256 * 56 * 56, 256
    cabi.tune_set("bn_vec_bf16", 8)
    g = torch.Generator(device=DEV).manual_seed(3)
    mk = lambda s, o: (torch.randn(m, c, generator=g, device=DEV) * s + o).to(torch.bfloat16)  # noqa: E731
    x, z, dy, dy2 = mk(1.7, 0.4), mk(1.0, 0.0), mk(1.0, 0.0), mk(0.5, 0.0)
    gamma = (torch.rand(c, generator=g, device=DEV) + 0.5)
    beta = torch.randn(c, generator=g, device=DEV) * 0.3
    v4 = lambda t: t.view(1, m, 1, c).permute(0, 3, 1, 2)  # noqa: E731
    rm, rv = torch.zeros(c, device=DEV), torch.ones(c, device=DEV)
    y, saved, mask = cabi.bn_fwd_train(v4(x), v4(z), gamma, beta, rm, rv, None, 1e-5, 0.1, cabi.ACT_ADD_RELU)
    dx, dz, dg, db = cabi.bn_bwd(v4(dy), v4(x), mask, gamma, saved, cabi.ACT_ADD_RELU, dy2=v4(dy2))
    torch.cuda.synchronize()
    f32 = lambda t: t.float().cpu().numpy().reshape(m, c)  # noqa: E731
    o = obn.bn_fwd_bwd_chunked_oracle(f32(x), gamma.cpu().numpy(), beta.cpu().numpy(), f32(dy), obn.ACT_ADD_RELU, f32(z),
                                      f32(dy2))
    sv = saved.cpu().numpy().astype(np.float64)
    assert np.allclose(sv[0], o["mean"], rtol=1e-5, atol=1e-6) and np.allclose(sv[1], o["invstd"], rtol=1e-5)
    assert np.allclose(rv.cpu().numpy(), 0.9 + 0.1 * o["var_unbiased"], rtol=1e-5)

    def close_bf16(got, want, what):
        got = got.permute(0, 2, 3, 1).float().cpu().numpy().reshape(m, c)
        err = np.abs(got - want)
        tol = np.abs(want) * 2.0 ** -8 + np.abs(want).max() * 2.0 ** -16 + 1e-6
        bad = err > tol
        # elements whose pre-activation sits within float32 rounding of zero may legitimately fall on the other side of
        # the ReLU; they are counted, not tolerated silently
        assert bad.mean() < 1e-5, (what, float(bad.mean()), float(err.max()))

    close_bf16(y, o["y"], "y")
    close_bf16(dz, o["dz"], "dz")
    close_bf16(dx, o["dx"], "dx")
    scale_g, scale_b = np.abs(o["dgamma"]).max(), np.abs(o["dbeta"]).max()
    assert np.abs(dg.cpu().numpy() - o["dgamma"]).max() <= 2e-3 * scale_g     # (a boundary element moves a sum by ~1)
    assert np.abs(db.cpu().numpy() - o["dbeta"]).max() <= 2e-3 * scale_b


def test_bn_golden_case_from_torch():
    c = np.load(os.path.join(GOLD, "bn_case.npz"))

    def cl(a):
        return torch.from_numpy(a).to(DEV).contiguous(memory_format=torch.channels_last)

    x, z, dy = cl(c["x"]), cl(c["z"]), cl(c["dy"])
    g, b = torch.from_numpy(c["gamma"]).to(DEV), torch.from_numpy(c["beta"]).to(DEV)
    for name, act in (("none", 0), ("relu", 1), ("add_relu", 2)):
        rm, rv = torch.zeros(16, device=DEV), torch.ones(16, device=DEV)
        nbt = torch.zeros((), dtype=torch.int64, device=DEV)
        y, saved, mask = cabi.bn_fwd_train(x, z if act == 2 else None, g, b, rm, rv, nbt, 1e-5, 0.1, act)
        assert y.shape == x.shape and y.is_contiguous(memory_format=torch.channels_last)
        assert np.allclose(y.cpu().numpy(), c[f"{name}_y"], rtol=1e-5, atol=1e-5)
        dx, dz, dg, db = cabi.bn_bwd(dy, x, mask, g, saved, act)
        assert np.allclose(dx.cpu().numpy(), c[f"{name}_dx"], rtol=1e-4, atol=1e-5)
        assert np.allclose(dg.cpu().numpy(), c[f"{name}_dgamma"], rtol=1e-4, atol=1e-4)
        assert np.allclose(db.cpu().numpy(), c[f"{name}_dbeta"], rtol=1e-4, atol=1e-4)
        if act == 2:
            assert np.allclose(dz.cpu().numpy(), c["add_relu_dz"], rtol=1e-6, atol=1e-6)
    assert np.allclose(rm.cpu().numpy(), c["running_mean"], rtol=1e-5, atol=1e-6)
    assert np.allclose(rv.cpu().numpy(), c["running_var"], rtol=1e-5, atol=1e-6)
    assert int(nbt) == 1


def test_bn_large_mean_does_not_cancel():
    """|mean| >> std: the shifted-sum statistics must not lose the variance."""
    m, c = 8192, 64
    rng = np.random.default_rng(0)
    x = (rng.standard_normal((m, c)) * 0.01 + 100.0).astype(np.float32)
    xd = torch.from_numpy(x).to(DEV)
    g, b = torch.ones(c, device=DEV), torch.zeros(c, device=DEV)
    _, saved, _ = cabi.bn_fwd_train(xd, None, g, b, torch.zeros(c, device=DEV), torch.ones(c, device=DEV), None, 1e-5,
                                    0.1, 0)
    var = x.astype(np.float64).var(axis=0)
    invstd = 1.0 / np.sqrt(var + 1e-5)
    assert np.allclose(saved[1].cpu().numpy(), invstd, rtol=1e-3)


@pytest.mark.parametrize("shape", [(256, 64, 112, 112), (256, 256, 56, 56), (256, 2048, 7, 7)])
def test_bn_full_size_properties(shape):
    """BASELINE sizes (ResNet-50, N=256, bf16): checked through properties that do not need the
    oracle to chew 0.8 G elements - normalised output has zero mean / unit variance per channel,
    dbeta = sum(dy), dx is orthogonal to 1 and to x-hat."""
    n, c, h, w = shape
    g = torch.Generator(device=DEV).manual_seed(1)
    x = (torch.randn(n, h, w, c, device=DEV, generator=g) * 1.5 + 0.7).to(torch.bfloat16).permute(0, 3, 1, 2)
    dy = torch.randn(n, h, w, c, device=DEV, generator=g).to(torch.bfloat16).permute(0, 3, 1, 2)
    gamma, beta = torch.ones(c, device=DEV), torch.zeros(c, device=DEV)
    rm, rv = torch.zeros(c, device=DEV), torch.ones(c, device=DEV)
    y, saved, _ = cabi.bn_fwd_train(x, None, gamma, beta, rm, rv, None, 1e-5, 0.1, cabi.ACT_NONE)
    yf = y.float()
    assert yf.mean((0, 2, 3)).abs().max().item() < 3e-3
    assert (yf.pow(2).mean((0, 2, 3)) - 1).abs().max().item() < 1e-2
    xf = x.float()
    assert torch.allclose(saved[0], xf.mean((0, 2, 3)), rtol=1e-4, atol=1e-4)
    dx, _, dg, db = cabi.bn_bwd(dy, x, None, gamma, saved, cabi.ACT_NONE)
    m = n * h * w
    assert torch.allclose(db, dy.float().sum((0, 2, 3)), rtol=1e-3, atol=1e-3 * m ** 0.5)
    dxf = dx.float()
    assert (dxf.sum((0, 2, 3)).abs() / m).max().item() < 1e-3           # sum(dx) = 0
    xhat = (xf - saved[0].view(1, c, 1, 1)) * saved[1].view(1, c, 1, 1)
    assert ((dxf * xhat).sum((0, 2, 3)).abs() / m).max().item() < 1e-2  # sum(dx * xhat) = 0


# ------------------------------------------------------------------------------------------------ K1
def _run_input(images, boxes, flips, oh, ow, mean, std, dtype, layout):
    out = cabi.input_stage(torch.from_numpy(images).to(DEV), torch.from_numpy(np.asarray(boxes, dtype=np.int32)).to(DEV),
                           torch.from_numpy(np.asarray(flips, dtype=np.uint8)).to(DEV), oh, ow, mean, std, dtype, layout)
    return out.float().cpu().permute(0, 2, 3, 1).numpy()  # logical NCHW -> NHWC values


def test_input_stage_golden_bit_exact():
    c = np.load(os.path.join(GOLD, "input_case.npz"))
    n = len(c["images"])
    full = np.tile(np.array([[0, 0, 96, 80]], dtype=np.int32), (n, 1))
    for layout in (cabi.LAYOUT_NHWC, cabi.LAYOUT_NCHW):
        out = _run_input(c["images"], full, np.zeros(n), 56, 56, (.5, .5, .5), (.5, .5, .5), torch.float32, layout)
        assert np.array_equal(out, c["ref_mode_out"])      # reference mode == torchvision on PIL, bit-exact
        out = _run_input(c["images"], c["boxes"], c["flips"], 48, 48, c["rrc_mean"].tolist(), c["rrc_std"].tolist(),
                         torch.float32, layout)
        assert np.array_equal(out, c["rrc_out"])           # crop + resize + flip + normalise, bit-exact
    out = _run_input(c["images"], c["boxes"], c["flips"], 48, 48, c["rrc_mean"].tolist(), c["rrc_std"].tolist(),
                     torch.bfloat16, cabi.LAYOUT_NHWC)
    assert np.array_equal(out, _bf16_round(c["rrc_out"]))  # bf16 = one rounding of the float32 result


@pytest.mark.parametrize("sh,sw,oh,ow", [(256, 256, 224, 224), (500, 375, 448, 448), (100, 120, 224, 224),
                                         (375, 500, 224, 224), (33, 57, 7, 9), (224, 224, 224, 224),
                                         (640, 480, 96, 96)])
def test_input_stage_matches_oracle_bit_exact(sh, sw, oh, ow):
    rng = np.random.default_rng(sh * 7 + sw)
    n = 5
    images = rng.integers(0, 256, (n, sh, sw, 3), dtype=np.uint8)
    g = torch.Generator().manual_seed(sh + sw)
    from imagent_distributed_training_pytorch_with_slurm_b200.data import draw_batch_params
    boxes, flips = draw_batch_params(n, sh, sw, g)
    boxes, flips = boxes.numpy().copy(), flips.numpy().copy()
    boxes[0] = (0, 0, sh, sw)            # whole image
    boxes[1] = (sh - 1, sw - 1, 1, 1)    # single-pixel box in the corner (pure up-scaling)
    boxes[2] = (0, sw // 2, sh, 1)       # one column
    flips[0], flips[1] = 0, 1
    mean, std = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)
    ref = oin.input_stage_oracle(images, boxes, flips.astype(bool), oh, ow, mean, std)
    out = _run_input(images, boxes, flips, oh, ow, mean, std, torch.float32, cabi.LAYOUT_NHWC)
    assert np.array_equal(out, ref)


def test_input_stage_full_batch_properties():
    """BASELINE size (256 x 256x256 -> 224x224 bf16): flip is an exact mirror, and a constant image
    stays constant through resize and normalisation."""
    n = 256
    g = torch.Generator().manual_seed(0)
    images = torch.randint(0, 256, (n, 256, 256, 3), dtype=torch.uint8, generator=g)
    images[0] = 77
    from imagent_distributed_training_pytorch_with_slurm_b200.data import draw_batch_params
    boxes, _ = draw_batch_params(n, 256, 256, g)
    dev = images.to(DEV)
    mean, std = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)
    a = cabi.input_stage(dev, boxes.to(DEV), torch.zeros(n, dtype=torch.uint8, device=DEV), 224, 224, mean, std)
    b = cabi.input_stage(dev, boxes.to(DEV), torch.ones(n, dtype=torch.uint8, device=DEV), 224, 224, mean, std)
    assert a.is_contiguous(memory_format=torch.channels_last) and a.dtype == torch.bfloat16
    assert torch.equal(a, b.flip(3))
    lut = oin.normalize_lut(mean, std)
    for ch in range(3):
        assert torch.all(a[0, ch].float() == float(_bf16_round(lut[ch][77])))
    # three samples against the oracle
    ref = oin.input_stage_oracle(images[5:8].numpy(), boxes[5:8].numpy(), np.zeros(3, bool), 224, 224, mean, std)
    assert np.array_equal(a[5:8].float().cpu().permute(0, 2, 3, 1).numpy(), _bf16_round(ref))


# ------------------------------------------------------------------------------------------------ K6
@pytest.mark.parametrize("shape", [(2, 8, 8, 8), (3, 16, 7, 9), (2, 64, 12, 5), (1, 8, 1, 1), (2, 24, 2, 3)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_maxpool_matches_oracle(shape, dtype):
    n, c, h, w = shape
    g = torch.Generator().manual_seed(h * 100 + w)
    x = torch.randn(n, h, w, c, generator=g).relu().to(dtype)       # post-ReLU: many exact ties at 0
    x[0, 0, 0, :] = float("-inf") if h * w > 1 else x[0, 0, 0, :]
    xd = x.to(DEV).permute(0, 3, 1, 2)
    y, arg = cabi.maxpool_fwd(xd)
    yo, pos = omp.maxpool3x3s2_fwd_oracle(x.float().numpy())
    assert np.array_equal(y.float().cpu().permute(0, 2, 3, 1).numpy(), yo)   # max of inputs: exact
    assert np.array_equal(arg.cpu().numpy(), pos)                            # argmax incl. tie rule: exact
    dy = torch.randn(yo.shape, generator=g).to(dtype)
    dx = cabi.maxpool_bwd(dy.to(DEV).permute(0, 3, 1, 2), arg, xd.shape)
    dxo = omp.maxpool3x3s2_bwd_oracle(dy.float().numpy(), pos, (h, w))
    got = dx.float().cpu().permute(0, 2, 3, 1).numpy()
    if dtype == torch.float32:
        assert np.allclose(got, dxo, rtol=1e-6, atol=1e-6)
    else:
        assert np.abs(got - dxo).max() <= np.abs(dxo).max() * 2 ** -8 + 1e-9
    # fan-in: a second gradient branch is summed on the fly (float32 sum, no intermediate rounding)
    dy2 = torch.randn(yo.shape, generator=g).to(dtype)
    dx2 = cabi.maxpool_bwd(dy.to(DEV).permute(0, 3, 1, 2), arg, xd.shape, dy2.to(DEV).permute(0, 3, 1, 2))
    dxo2 = omp.maxpool3x3s2_bwd_oracle(dy.float().numpy() + dy2.float().numpy(), pos, (h, w))
    got2 = dx2.float().cpu().permute(0, 2, 3, 1).numpy()
    if dtype == torch.float32:
        assert np.allclose(got2, dxo2, rtol=1e-6, atol=1e-6)
    else:
        assert np.abs(got2 - dxo2).max() <= np.abs(dxo2).max() * 2 ** -8 + 1e-9


def test_maxpool_bf16_packed_path_follows_the_scalar_rule_with_nan_inf_and_signed_zero():
    """The bf16 kernel compares packed pairs; windows holding a NaN are redone by the scalar rule.  Checker: the
    float32 kernel (scalar rule, pinned against the oracle above) on the same bf16-representable values, and ATen."""
    g = torch.Generator().manual_seed(11)
    n, h, w, c = 3, 13, 10, 16
    x = torch.randn(n, h, w, c, generator=g).relu()
    x[torch.rand(n, h, w, c, generator=g) < 0.05] = float("nan")
    x[torch.rand(n, h, w, c, generator=g) < 0.05] = float("-inf")
    x[torch.rand(n, h, w, c, generator=g) < 0.05] = -0.0
    x[1, 4:9, 2:7, :] = float("nan")                      # windows made of NaN only: the LAST tap is the argmax
    x[2, 0:5, 0:5, :] = float("-inf")                     # windows made of -inf only: the FIRST tap in the image
    xb = x.to(torch.bfloat16)
    yb, ab = cabi.maxpool_fwd(xb.to(DEV).permute(0, 3, 1, 2))
    yf, af = cabi.maxpool_fwd(xb.float().to(DEV).permute(0, 3, 1, 2))
    assert torch.equal(ab, af)
    assert torch.equal(torch.isnan(yb), torch.isnan(yf))
    assert torch.equal(torch.nan_to_num(yb.float(), nan=7.0), torch.nan_to_num(yf, nan=7.0))
    # ATen's contiguous-layout kernel: its channels-last kernel starts an all -inf window at a different index
    yr, idx = torch.ops.aten.max_pool2d_with_indices(xb.to(DEV).permute(0, 3, 1, 2).contiguous(), [3, 3], [2, 2], [1, 1],
                                                     [1, 1], False)
    assert torch.equal(torch.isnan(yb), torch.isnan(yr))
    oy = torch.arange(yb.shape[2], device=DEV).view(1, 1, -1, 1)
    ox = torch.arange(yb.shape[3], device=DEV).view(1, 1, 1, -1)
    pos = (idx // w - (2 * oy - 1)) * 3 + (idx % w - (2 * ox - 1))
    assert torch.equal(ab.permute(0, 3, 1, 2).long(), pos)


@pytest.mark.parametrize("shape", [(2, 9, 7, 8), (3, 12, 12, 64), (1, 5, 6, 16), (4, 56, 56, 64), (2, 1, 1, 8)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_stem_bn_relu_maxpool_in_one_pass_equals_the_two_kernels(shape, dtype):
    """b200_bn_relu_maxpool3x3s2_fwd pools over relu(x*scale + shift) without storing it: values AND argmax must have the
    bits of b200_bn_fwd_train(act=RELU) followed by b200_maxpool3x3s2_fwd (which are checked against the oracle above),
    including windows cut by the image border and channels whose activations are all clipped to zero (ties)."""
    n, h, w, c = shape
    g = torch.Generator().manual_seed(h * 100 + w)
    x = (torch.randn(n, h, w, c, generator=g) * 2.0 - 0.5).to(dtype).to(DEV).permute(0, 3, 1, 2)
    gamma = torch.rand(c, generator=g).add(0.5).to(DEV)
    beta = (torch.randn(c, generator=g) * 0.5).to(DEV)
    beta[0] = -50.0                                           # a channel that the ReLU zeroes everywhere: all ties
    if n * h * w < 2:
        pytest.skip("training statistics need more than one value per channel")
    rm, rv = torch.zeros(c, device=DEV), torch.ones(c, device=DEV)
    a, saved, _ = cabi.bn_fwd_train(x, None, gamma, beta, rm, rv, None, 1e-5, 0.1, cabi.ACT_RELU)
    y2, arg2 = cabi.maxpool_fwd(a)
    rm2, rv2 = torch.zeros(c, device=DEV), torch.ones(c, device=DEV)
    saved1 = cabi.bn_stats(x, gamma, beta, rm2, rv2, None, 1e-5, 0.1)
    y1, arg1 = cabi.bn_relu_maxpool_fwd(x, saved1)
    assert torch.equal(saved1, saved) and torch.equal(rm, rm2) and torch.equal(rv, rv2)
    assert torch.equal(y1, y2) and torch.equal(arg1, arg2)
    assert float(y1[:, 0].abs().max()) == 0.0 and int(arg1[..., 0].max()) <= 4   # ties resolve to the first tap inside


def test_maxpool_full_size_against_aten():
    """BASELINE size [256,64,112,112] bf16: the library kernel the reference would have run is the
    checker here (max values and routed gradients must agree exactly / to one bf16 rounding)."""
    g = torch.Generator(device=DEV).manual_seed(0)
    x = torch.randn(256, 112, 112, 64, device=DEV, generator=g).relu().to(torch.bfloat16).permute(0, 3, 1, 2)
    y, arg = cabi.maxpool_fwd(x)
    yr, idx = torch.ops.aten.max_pool2d_with_indices(x, [3, 3], [2, 2], [1, 1], [1, 1], False)
    assert torch.equal(y, yr)
    dy = torch.randn(256, 56, 56, 64, device=DEV, generator=g).to(torch.bfloat16).permute(0, 3, 1, 2)
    dx = cabi.maxpool_bwd(dy, arg, x.shape)
    dxr = torch.ops.aten.max_pool2d_with_indices_backward(dy, x, [3, 3], [2, 2], [1, 1], [1, 1], False, idx)
    assert (dx.float() - dxr.float()).abs().max().item() <= dxr.float().abs().max().item() * 2 ** -7


@pytest.mark.parametrize("sh,sw,oh,ow", [(90, 70, 64, 48), (256, 256, 224, 224), (500, 375, 448, 448),
                                         (100, 120, 224, 224), (33, 57, 8, 10), (224, 224, 224, 224),
                                         (300, 480, 96, 96)])
def test_input_stage_space_to_depth_layout_is_a_reindexing(sh, sw, oh, ow):
    """The stem layout has its own kernel (RGBx words in shared memory, one cell per thread); it must be the oracle's
    image re-indexed, bit for bit, and identical to the generic kernel's output for that layout."""
    rng = np.random.default_rng(3 + sh)
    n = 6
    images = rng.integers(0, 256, (n, sh, sw, 3), dtype=np.uint8)
    from imagent_distributed_training_pytorch_with_slurm_b200.data import draw_batch_params
    boxes, flips = draw_batch_params(n, sh, sw, torch.Generator().manual_seed(1))
    boxes, flips = boxes.clone(), flips.clone()
    boxes[0] = torch.tensor([0, 0, sh, sw])                  # whole image (first bytes of the batch)
    boxes[1] = torch.tensor([sh - 1, sw - 1, 1, 1])          # single pixel, pure up-scaling
    boxes[2] = torch.tensor([0, sw // 2, sh, 1])             # one column
    boxes[n - 1] = torch.tensor([sh // 2, sw // 3, sh - sh // 2, sw - sw // 3])   # ends at the last byte of the batch
    flips[0], flips[1], flips[n - 1] = 0, 1, 1
    mean, std = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)
    ref = oin.input_stage_oracle(images, boxes.numpy(), flips.numpy().astype(bool), oh, ow, mean, std)
    want = oin.space_to_depth_stem_oracle(ref)
    # the images live at an odd byte offset inside a larger allocation: exercises the unaligned word loads
    backing = torch.zeros(images.size + 7, dtype=torch.uint8, device=DEV)
    for shift in (0, 3):
        dev_images = backing[shift:shift + images.size].view(n, sh, sw, 3)
        dev_images.copy_(torch.from_numpy(images))
        for dtype in (torch.float32, torch.bfloat16):
            outs = []
            for fast in (1, 0):
                cabi.tune_set("input_s2d_fast", fast)
                try:
                    out = cabi.input_stage(dev_images, boxes.to(DEV), flips.to(DEV), oh, ow, mean, std, dtype,
                                           cabi.LAYOUT_S2D_STEM)
                finally:
                    cabi.tune_set("input_s2d_fast", 1)
                assert out.shape == (n, 16, (oh + 6) // 2, (ow + 6) // 2) and out._b200_s2d_stem == (oh, ow)
                outs.append(out)
            got = outs[0].float().cpu().permute(0, 2, 3, 1).numpy()
            assert np.array_equal(got, want if dtype == torch.float32 else _bf16_round(want))
            assert torch.equal(outs[0], outs[1])


# ------------------------------------------------------------------------------------------------ K5
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_pack_grads_matches_per_tensor_copies(dtype):
    g = torch.Generator().manual_seed(0)
    shapes = [(64, 3, 7, 7), (64, 64, 1, 1), (128, 64, 3, 3), (1000, 512), (7,), (13, 5, 3, 3)] + [(8, 8, 1, 1)] * 70
    tensors, offsets, off = [], [], 0
    for shp in shapes:
        t = torch.randn(*shp, generator=g).to(dtype).to(DEV)
        if len(shp) == 4:
            t = t.contiguous(memory_format=torch.channels_last)
        tensors.append(t)
        offsets.append(off)
        off += (t.numel() + 31) // 32 * 32
    dst = torch.full((off,), 7.0, device=DEV)
    cabi.pack_grads(tensors, offsets, dst, 0.5)
    want = torch.full((off,), 7.0, device=DEV)
    for t, o in zip(tensors, offsets):
        flat = t.permute(0, 2, 3, 1).reshape(-1) if t.dim() == 4 else t.reshape(-1)
        want[o:o + t.numel()] = flat.float() * 0.5
    assert torch.equal(dst, want)   # cast + power-of-two scale: exact; padding between slots untouched


# ------------------------------------------------------------------------------------------------ K2 dual
@pytest.mark.parametrize("m,c", [(210, 64), (4099, 256), (12544, 2048), (50176, 512)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("two_branches", [False, True])
def test_bn_dual_matches_separate_kernels_and_oracle(m, c, dtype, two_branches):
    """y = relu(bn(x) + bn2(x2)) in the fused form against (a) the oracle composed of a plain BatchNorm
    followed by bn+add+relu, (b) the separate kernels."""
    rng = np.random.default_rng(m + c)
    def arr(scale=1.0, shift=0.0):
        return (rng.standard_normal((m, c)) * scale + shift).astype(np.float32)
    x, x2, dy, dyb = arr(1.5, 0.3), arr(0.7, -1.0), arr(), arr()
    g1, b1 = rng.uniform(0.5, 1.5, c).astype(np.float32), (rng.standard_normal(c) * 0.3).astype(np.float32)
    g2, b2 = rng.uniform(0.5, 1.5, c).astype(np.float32), (rng.standard_normal(c) * 0.3).astype(np.float32)
    T = lambda a: _rows_t(a, dtype)
    xd, x2d, dyd, dybd = T(x), T(x2), T(dy), T(dyb)
    xs, x2s = xd.float().cpu().numpy(), x2d.float().cpu().numpy()
    dys = dyd.float().cpu().numpy() + (dybd.float().cpu().numpy() if two_branches else 0.0)
    dev = lambda a: torch.from_numpy(a).to(DEV)
    zeros, ones = (lambda: torch.zeros(c, device=DEV)), (lambda: torch.ones(c, device=DEV))
    if not cabi.bn_dual_supported(torch.empty(1, c, 1, 1, dtype=dtype)):
        pytest.skip("outside the bulk-async kernels (C/vec > 256): the model uses the separate entry points")
    # fused
    saved2 = cabi.bn_stats(x2d, dev(g2), dev(b2), zeros(), ones(), None, 1e-5, 0.1)
    y, saved, mask = cabi.bn_fwd_train_dual(xd, x2d, saved2, dev(g1), dev(b1), zeros(), ones(), None, 1e-5, 0.1)
    dg1, db1, dg2, db2 = (torch.empty(c, device=DEV) for _ in range(4))
    dx, dx2 = cabi.bn_bwd_dual(dyd, dybd if two_branches else None, xd, x2d, mask, dev(g1), dev(g2), saved, saved2, dg1,
                               db1, dg2, db2)
    # oracle
    z_o, mean2, invstd2, _, _, _ = obn.bn_train_fwd_oracle(x2s, g2, b2, np.zeros(c), np.ones(c), act=obn.ACT_NONE)
    y_o, mean1, invstd1, _, _, pos = obn.bn_train_fwd_oracle(xs, g1, b1, np.zeros(c), np.ones(c), act=obn.ACT_ADD_RELU,
                                                            z=z_o)
    dx_o, dz_o, dg1_o, db1_o = obn.bn_train_bwd_oracle(dys, xs, g1, mean1, invstd1, obn.ACT_ADD_RELU, pos)
    dx2_o, _, dg2_o, db2_o = obn.bn_train_bwd_oracle(dz_o, x2s, g2, mean2, invstd2, obn.ACT_NONE)
    pre = (xs.astype(np.float64) - mean1) * invstd1 * g1 + b1 + (x2s.astype(np.float64) - mean2) * invstd2 * g2 + b2
    safe = np.abs(pre) > 1e-3
    bf = dtype == torch.bfloat16
    tol = (lambda ref: np.abs(ref).max() * 2 ** -7) if bf else (lambda ref: 1e-4 * (np.abs(ref).max() + 1))
    assert np.abs(y.float().cpu().numpy() - y_o).max() <= tol(y_o)
    assert (np.abs(dx.float().cpu().numpy() - dx_o)[safe]).max() <= tol(dx_o)
    assert (np.abs(dx2.float().cpu().numpy() - dx2_o)[safe]).max() <= tol(dx2_o)
    sc = np.abs(dg1_o).max() + np.abs(db1_o).max() + np.abs(dg2_o).max()
    for got, want in ((dg1, dg1_o), (db1, db1_o), (dg2, dg2_o), (db2, db2_o)):
        assert np.allclose(got.cpu().numpy(), want, rtol=2e-3 if bf else 1e-4, atol=(2e-3 if bf else 1e-4) * sc + 1e-3)
    # separate kernels give the same forward output bit for bit in float32 (same fma sequence on z = bn2(x2)
    # is not guaranteed: z is rounded to the activation dtype when materialised), so compare loosely
    z_k, _, _ = cabi.bn_fwd_train(x2d, None, dev(g2), dev(b2), zeros(), ones(), None, 1e-5, 0.1, cabi.ACT_NONE)
    y_k, _, _ = cabi.bn_fwd_train(xd, z_k, dev(g1), dev(b1), zeros(), ones(), None, 1e-5, 0.1, cabi.ACT_ADD_RELU)
    assert (y.float() - y_k.float()).abs().max().item() <= tol(y_o) * 2


# ------------------------------------------------------------------------------------------------ guard bands
def _guarded(nbytes, dev=DEV, pad=4096):
    """A byte buffer with 0xA5 guard bands on both sides; returns (whole, start offset)."""
    whole = torch.full((nbytes + 2 * pad,), 0xA5, dtype=torch.uint8, device=dev)
    return whole, pad


def _guards_intact(whole, pad, nbytes):
    return bool((whole[:pad] == 0xA5).all()) and bool((whole[pad + nbytes:] == 0xA5).all())


@pytest.mark.parametrize("m,c", [(4099, 64), (777, 256), (301, 2048), (50177, 8)])
@pytest.mark.parametrize("act", [cabi.ACT_RELU, cabi.ACT_ADD_RELU])
def test_bn_kernels_write_only_inside_their_outputs(m, c, act):
    """compute-sanitizer is not available on the GPU pool, so out-of-bounds writes are looked for with
    guard bands: every output of the (bulk-async) BatchNorm kernels is carved out of a larger buffer of
    0xA5 bytes, through the raw C ABI, for ragged row counts."""
    lib = cabi.load()
    g = torch.Generator(device=DEV).manual_seed(m)
    x = torch.randn(m, c, device=DEV, generator=g).to(torch.bfloat16)
    z = torch.randn(m, c, device=DEV, generator=g).to(torch.bfloat16)
    dy = torch.randn(m, c, device=DEV, generator=g).to(torch.bfloat16)
    gamma, beta = torch.ones(c, device=DEV), torch.zeros(c, device=DEV)
    rm, rv = torch.zeros(c, device=DEV), torch.ones(c, device=DEV)
    nb = m * c * 2
    mb = m * (c // 8) + 16
    bufs = {k: _guarded(n) for k, n in (("y", nb), ("mask", mb), ("saved", 16 * c), ("dx", nb), ("dz", nb),
                                         ("dg", 4 * c), ("db", 4 * c))}
    ptr = lambda k: bufs[k][0].data_ptr() + bufs[k][1]
    ws = torch.empty(lib.b200_bn_workspace_bytes(m, c, cabi.BF16), dtype=torch.uint8, device=DEV)
    st = torch.cuda.current_stream().cuda_stream
    rc = lib.b200_bn_fwd_train(x.data_ptr(), z.data_ptr() if act == 2 else None, ptr("y"), ptr("mask") if act == 2 else None,
                               gamma.data_ptr(), beta.data_ptr(), rm.data_ptr(), rv.data_ptr(), None, ptr("saved"),
                               ws.data_ptr(), ws.numel(), m, c, 1e-5, 0.1, act, cabi.BF16, st)
    assert rc == 0
    rc = lib.b200_bn_bwd(dy.data_ptr(), None, x.data_ptr(), ptr("mask") if act == 2 else None, gamma.data_ptr(),
                         ptr("saved"), ptr("dx"), ptr("dz") if act == 2 else None, ptr("dg"), ptr("db"), ws.data_ptr(),
                         ws.numel(), m, c, act, cabi.BF16, st)
    assert rc == 0
    torch.cuda.synchronize()
    sizes = {"y": nb, "mask": mb, "saved": 16 * c, "dx": nb, "dz": nb, "dg": 4 * c, "db": 4 * c}
    for k, (whole, pad) in bufs.items():
        assert _guards_intact(whole, pad, sizes[k]), k
    # and the results carved out of the guarded buffers are the ones the wrappers give
    y2, saved2, _ = cabi.bn_fwd_train(x, z if act == 2 else None, gamma, beta, torch.zeros(c, device=DEV),
                                      torch.ones(c, device=DEV), None, 1e-5, 0.1, act)
    y = bufs["y"][0][4096:4096 + nb].view(torch.bfloat16).view(m, c)
    assert torch.equal(y, y2)


def test_maxpool_and_pack_write_only_inside_their_outputs():
    lib = cabi.load()
    n, c, h, w = 3, 24, 13, 9
    oh, ow = (h - 1) // 2 + 1, (w - 1) // 2 + 1
    g = torch.Generator(device=DEV).manual_seed(0)
    x = torch.randn(n, h, w, c, device=DEV, generator=g).to(torch.bfloat16)
    dy = torch.randn(n, oh, ow, c, device=DEV, generator=g).to(torch.bfloat16)
    sizes = {"y": n * oh * ow * c * 2, "arg": n * oh * ow * c, "dx": n * h * w * c * 2}
    bufs = {k: _guarded(v) for k, v in sizes.items()}
    ptr = lambda k: bufs[k][0].data_ptr() + bufs[k][1]
    st = torch.cuda.current_stream().cuda_stream
    assert lib.b200_maxpool3x3s2_fwd(x.data_ptr(), ptr("y"), ptr("arg"), n, c, h, w, cabi.BF16, st) == 0
    assert lib.b200_maxpool3x3s2_bwd(dy.data_ptr(), None, ptr("arg"), ptr("dx"), n, c, h, w, cabi.BF16, st) == 0
    torch.cuda.synchronize()
    for k, (whole, pad) in bufs.items():
        assert _guards_intact(whole, pad, sizes[k]), k
    # pack: slots separated by untouched padding
    ts = [torch.randn(n_, device=DEV).to(torch.bfloat16) for n_ in (7, 64, 1000, 9408)]
    offs, off = [], 32
    for t in ts:
        offs.append(off)
        off += (t.numel() + 31) // 32 * 32 + 32
    dst = torch.full((off,), -7.0, device=DEV)
    cabi.pack_grads(ts, offs, dst)
    touched = torch.zeros(off, dtype=torch.bool, device=DEV)
    for t, o in zip(ts, offs):
        touched[o:o + t.numel()] = True
        assert torch.equal(dst[o:o + t.numel()], t.float())
    assert bool((dst[~touched] == -7.0).all())


@pytest.mark.parametrize("layout", [cabi.LAYOUT_S2D_STEM, cabi.LAYOUT_NHWC])
def test_input_stage_writes_only_inside_its_output_and_workspace(layout):
    """Guard bands around the output and the coefficient workspace of the input stage (both kernels of the stem
    layout write whole cells / tables with vector stores)."""
    import ctypes

    lib = cabi.load()
    n, sh, sw, oh, ow = 5, 61, 83, 40, 36
    g = torch.Generator().manual_seed(5)
    images = torch.randint(0, 256, (n, sh, sw, 3), dtype=torch.uint8, generator=g).to(DEV)
    from imagent_distributed_training_pytorch_with_slurm_b200.data import draw_batch_params
    boxes, flips = draw_batch_params(n, sh, sw, g)
    boxes, flips = boxes.to(DEV), flips.to(DEV)
    out_elems = n * ((oh + 6) // 2) * ((ow + 6) // 2) * 16 if layout == cabi.LAYOUT_S2D_STEM else n * oh * ow * 3
    ws_bytes = lib.b200_input_workspace_bytes(n, sh, sw, oh, ow)
    sizes = {"out": out_elems * 2, "ws": ws_bytes}
    bufs = {k: _guarded(v) for k, v in sizes.items()}
    ptr = lambda k: bufs[k][0].data_ptr() + bufs[k][1]
    f3 = ctypes.c_float * 3
    rc = lib.b200_input_crop_resize_flip_norm(images.data_ptr(), boxes.data_ptr(), flips.data_ptr(), ptr("out"), ptr("ws"),
                                              ws_bytes, n, sh, sw, oh, ow, f3(.485, .456, .406), f3(.229, .224, .225),
                                              cabi.BF16, layout, torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    for k, (whole, pad) in bufs.items():
        assert _guards_intact(whole, pad, sizes[k]), k
    # and the guarded call produced the same image as the wrapper
    want = cabi.input_stage(images, boxes, flips, oh, ow, (.485, .456, .406), (.229, .224, .225), torch.bfloat16, layout)
    got = bufs["out"][0][bufs["out"][1]:bufs["out"][1] + sizes["out"]].view(torch.bfloat16)
    ref = want.permute(0, 2, 3, 1).contiguous().view(-1) if layout != cabi.LAYOUT_NCHW else want.view(-1)
    assert torch.equal(got, ref)
