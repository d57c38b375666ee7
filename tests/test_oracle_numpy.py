"""The torch oracle against a second, independent restatement of the same reference lines written op by op in NumPy
float16 (oracle/numpy_f16_ref.py: no torch, no autograd, hand-written gradients).  Neither has met TensorFlow (parity
unpinned); agreement bit for bit means the fp16 model of the path does not rest on one formulation only."""
import numpy as np
import pytest
import torch

from oracle import numpy_f16_ref as N
from oracle import serann_oracle as O

H = np.float16


def _t(a):
    return torch.tensor(np.asarray(a, dtype=np.float32))


def ulps(a, b):
    """distance in fp16 units in the last place of the larger magnitude (NaN == NaN)"""
    a = np.asarray(a, dtype=np.float32)
    b = np.asarray(b, dtype=np.float32)
    both_nan = np.isnan(a) & np.isnan(b)
    mag = np.maximum(np.abs(a), np.abs(b))
    ulp = np.where(mag > 0, 2.0 ** (np.floor(np.log2(np.maximum(mag, 6.1e-5))) - 10), 6e-8)
    d = np.where(both_nan, 0.0, np.abs(a - b) / ulp)
    return d


@pytest.mark.parametrize('rows,C,seed', [(750, 7, 0), (6000, 16, 1), (48 * 676, 8, 2), (70000, 3, 3), (65519, 2, 4), (65520, 2, 5)])
def test_bn_train_forward_bit_equal(rows, C, seed):
    r = np.random.default_rng(seed)
    x = (r.standard_normal((rows, C)) * 1.7 + 0.4).astype(H)
    gamma = (1 + 0.2 * r.standard_normal(C)).astype(H)
    beta = (0.1 * r.standard_normal(C)).astype(H)
    mm = (0.1 * r.standard_normal(C)).astype(H)
    mv = np.abs(1 + 0.2 * r.standard_normal(C)).astype(H)
    y, nm, nv, _ = N.bn_train(x, gamma, beta, mm, mv)
    tm, tv = _t(mm), _t(mv)
    yo = O._bn_train(_t(x), _t(gamma), _t(beta), tm, tv).numpy().astype(H)
    assert np.array_equal(y, yo)
    assert np.array_equal(nm, tm.numpy().astype(H))
    assert np.array_equal(nv, tv.numpy().astype(H), equal_nan=True)
    if rows >= 65520:                              # F6: n cast to fp16 is inf from 65520 on (65519 still rounds to 65504), inf / inf = NaN
        assert np.isnan(nv).all() and np.isfinite(nm).all()
    else:
        assert np.isfinite(nv).all()
    yi = N.bn_infer(x, gamma, beta, mm, mv)
    assert np.array_equal(yi, O._bn_infer(_t(x), _t(gamma), _t(beta), _t(mm), _t(mv)).numpy().astype(H))


def test_moving_average_is_the_product_form_and_stagnates_in_fp16():
    """x * m + v * (1 - m) in fp16: a value within half an ulp of the fixed point no longer moves."""
    x = H(1.0)
    for _ in range(2000):
        x = N.moving_average_update(np.array([x], H), np.array([0.25], H))[0]
    assert abs(float(x) - 0.25) < 0.02
    assert N.moving_average_update(np.array([x], H), np.array([0.25], H))[0] == x
    assert float(N.BN_MOMENTUM) == 0.990234375 and float(H(1.0) - N.BN_MOMENTUM) == 0.009765625


def test_bn_train_backward_within_two_ulps_of_the_autograd_graph():
    """The hand-written TF gradient graph rounds the three contributions to dx separately (Mul grad, Cast grad, AddN);
    the torch oracle (and the CUDA kernel) add them in fp32 and round once: at most a couple of fp16 ulps apart."""
    r = np.random.default_rng(11)
    x = (r.standard_normal((512, 12)) * 1.3 + 0.2).astype(H)
    gamma = (1 + 0.2 * r.standard_normal(12)).astype(H)
    beta = (0.1 * r.standard_normal(12)).astype(H)
    dy = (r.standard_normal((512, 12)) * 1e-2).astype(H)
    y, _, _, saved = N.bn_train(x, gamma, beta, np.zeros(12, H), np.ones(12, H))
    dx, dg, db = N.bn_train_backward(dy, saved)
    xt = O.q(_t(x).requires_grad_(True))
    gt, bt = _t(gamma).requires_grad_(True), _t(beta).requires_grad_(True)
    leaf = xt
    leaf.retain_grad()
    yo, _, _ = O._bn_train_graph(leaf, gt, bt)
    yo.backward(_t(dy))
    dxo = leaf.grad.half().float().numpy()
    scale = np.abs(dxo).max()
    assert np.abs(dx.astype(np.float32) - dxo).max() <= 3 * scale * 2.0 ** -10
    assert ulps(dg, gt.grad.half().numpy()).max() <= 4
    assert ulps(db, bt.grad.half().numpy()).max() <= 2


@pytest.mark.parametrize('seed', [0, 1, 2])
def test_adam_bit_equal_over_steps(seed):
    """fp16 ApplyAdam incl. |g| < 7.8e-3 (v flushes to zero) and large gradients."""
    from serann_b200 import frontend
    r = np.random.default_rng(seed)
    ir = frontend.compile_source("con = X_layer\nloss_balance = 0.5")
    ind = O.Individual(ir, O.init_weights(ir, r))
    li, pi, p = list(ind.trainable())[0]
    p_np = p.detach().numpy().astype(H)
    m_np, v_np = np.zeros_like(p_np), np.zeros_like(p_np)
    for t in range(1, 9):
        scale = [1e-4, 1e-3, 5e-3, 1e-2, 0.1, 1.0, 30.0, 1e-6][t - 1]
        grads = {}
        for (lj, pj, q_) in ind.trainable():
            g = (r.standard_normal(tuple(q_.shape)) * scale).astype(np.float32)
            q_.grad = torch.tensor(g)
            grads[(lj, pj)] = g
        O.adam_step(ind, t)
        p_np, m_np, v_np = N.adam_step(p_np, m_np, v_np, grads[(li, pi)].astype(H), t)
        assert np.array_equal(p_np, p.detach().numpy().astype(H)), t
        assert np.array_equal(m_np, ind.m[li][pi].numpy()), t
        assert np.array_equal(v_np, ind.v[li][pi].numpy()), t
        if t <= 2:                                      # |g| well below 7.8e-3: (1 - beta2) * g^2 underflows, v stays zero
            assert float(np.abs(v_np).max()) == 0.0


def test_adam_scalars_match_the_engine_formula():
    for t in (1, 2, 10, 380, 2047, 2049):
        alpha, omb1, omb2, eps = N.adam_scalars(t)
        assert float(omb1) == float(H(1) - H(0.9)) and float(omb2) == float(H(1) - H(0.999))
        assert np.isfinite(alpha) and float(eps) == float(H(1e-4))


@pytest.mark.parametrize('lb', [0.3, 0.9951, 1.7])
def test_losses_and_logit_gradients(lb):
    r = np.random.default_rng(int(lb * 1000))
    B, ncls, G = 750, 10, 100
    logits = (r.standard_normal((B, ncls + G)) * 2.5).astype(H)
    logits[5, :ncls] = logits[5, 0]                     # an exact tie: first index wins
    labels = r.integers(0, ncls, B)
    bits = r.integers(0, 2, (B, G))
    out = N.losses_and_grads(logits, labels, bits, ncls, lb)
    z = _t(logits).requires_grad_(True)
    zq = O.q(z)
    ce, mse, correct, s = O.losses(zq, torch.tensor(labels), _t(bits), ncls)
    total, lc, lr_ = O.weighted_loss(ce, mse, lb)
    total.backward()
    assert np.array_equal(out['ce'], ce.detach().numpy().astype(H))
    assert np.array_equal(out['mse'], mse.detach().numpy().astype(H))
    assert np.array_equal(out['probs'], s.detach().numpy().astype(H))
    assert out['correct'] == int(correct.sum())
    assert float(out['total']) == float(total.detach())
    g = z.grad.numpy().astype(H)
    d = ulps(out['dlogits'], g)
    assert (d > 0).mean() < 1e-3 and d.max() <= 2       # fp32 exp / log-sum-exp order: isolated last-bit differences only
    assert np.array_equal(out['dlogits'][:, ncls:], g[:, ncls:])     # the replication head is pure fp16 op-by-op: exact


def test_f16_mean_matches_and_loses_integer_resolution_above_2048():
    a, b = N.F16Mean(), O.F16Mean()
    for c in [750, 749, 12, 750, 750, 3, 750]:
        a.update(c, 750)
        b.update(c, 750)
    assert a.result() == b.result()
    from serann_b200.worker import F16Mean as W
    w = W()
    for c in [750, 749, 12, 750, 750, 3, 750]:
        w.update(float(c), 750)
    assert w.result() == a.result()
    assert float(a.count) != 7 * 750                   # 5250 is not representable in fp16


def test_centred_bn_backward_is_the_accurate_one_on_ill_conditioned_channels():
    """Why the kernels (and the oracle's default) use the centred closed form of the BatchNorm gradient: on a channel
    with |mean| / std in the hundreds the op-by-op graph rounds sum(dy) to fp16 and multiplies by the mean afterwards;
    the cancellation noise swamps d gamma and is a visible share of dx.  Against a float64 evaluation of the same
    derivative the centred form is exact to fp32, the graph form is not; on well-conditioned channels the two agree."""
    r = np.random.default_rng(4)
    n, C = 750, 6
    x = (5.0 + 0.02 * r.standard_normal((n, C))).astype(H)          # mean / std = 250
    x[:, 3:] = (0.3 + 1.1 * r.standard_normal((n, 3))).astype(H)     # well conditioned
    gamma = np.ones(C, H)
    beta = np.zeros(C, H)
    dy = (1e-3 * (0.5 + r.standard_normal((n, C)))).astype(H)        # sum(dy) is not small

    def run(mode):
        O.BN_BACKWARD = mode
        try:
            leaf, g = _t(x).requires_grad_(True), _t(gamma).requires_grad_(True)
            y = O._bn_train(leaf, g, _t(beta), torch.zeros(C), torch.ones(C))
            y.backward(_t(dy))
            return leaf.grad.numpy().astype(np.float64), g.grad.numpy().astype(np.float64)
        finally:
            O.BN_BACKWARD = 'centered'
    xd, dyd = x.astype(np.float64), dy.astype(np.float64)
    mean = xd.mean(0)
    mean16 = mean.astype(np.float32).astype(H).astype(np.float64)
    var16 = xd.var(0).astype(np.float32).astype(H).astype(np.float64)
    inv_raw = (1.0 / np.sqrt((var16 + float(H(1e-3))).astype(H).astype(np.float64))).astype(H).astype(np.float64)
    s2 = (dyd * (xd - mean16)).sum(0)
    dx_true = inv_raw * (dyd - dyd.mean(0)) - (xd - mean) * inv_raw ** 3 * s2 / n
    dg_true = s2 * inv_raw
    (dx_c, dg_c), (dx_g, dg_g) = run('centered'), run('graph')
    err_x = lambda d: np.abs(d - dx_true).max(0) / np.abs(dx_true).max(0)
    err_g = lambda d: np.abs(d - dg_true) / np.abs(dg_true)
    assert (err_x(dx_c) < 1e-5).all() and (err_g(dg_c) < 1e-5).all()
    assert (err_x(dx_g)[3:] < 1e-3).all() and (err_g(dg_g)[3:] < 5e-3).all()         # well conditioned: the forms agree
    assert (err_g(dg_g)[:3] > 0.05).all() and (err_x(dx_g)[:3] > 1e-3).all()          # ill conditioned: the graph form is noise


# --------------------------------------------------------------------------- Dense / Conv1D / Conv2D, pooling
def _layer(kind, units, kernel=(1, 1), strides=(1, 1), act=0, out_shape=None):
    from types import SimpleNamespace
    return SimpleNamespace(kind=kind, units=units, kernel=kernel, strides=strides, act=act, out_shape=out_shape)


@pytest.mark.parametrize('case', [
    dict(x=(4, 9, 8, 5), kernel=(3, 3), strides=(1, 1), units=7, act='relu'),        # Conv2D
    dict(x=(3, 12, 10, 8), kernel=(5, 4), strides=(2, 3), units=6, act='linear'),    # strided, non-square window
    dict(x=(5, 40, 3), kernel=(1, 7), strides=(1, 2), units=4, act='sigmoid'),       # Conv1D = 1 x k convolution
    dict(x=(6, 17), kernel=(1, 1), strides=(1, 1), units=9, act='relu'),             # Dense
    dict(x=(2, 11, 6), kernel=(1, 1), strides=(1, 1), units=5, act='linear'),        # Dense on the last axis of a 3-D tensor
])
def test_conv_like_forward_and_activation_grad(case):
    """Kernel layouts (HWIO, Conv1D as H = 1), strides and the three roundings of a layer, through im2col + matmul in NumPy
    against torch's conv2d in the other restatement: equal except for rare last-bit ties of the fp32 summation order."""
    from serann_b200.frontend import KIND_CONV, KIND_DENSE, ACT_NONE, ACT_RELU, ACT_SIGMOID
    r = np.random.default_rng(11)
    x = r.standard_normal(case['x']).astype(H)
    kh, kw = case['kernel']
    cin = case['x'][-1]
    dense = kh * kw == 1
    wshape = (cin, case['units']) if dense else ((kw, cin, case['units']) if len(case['x']) == 3 else (kh, kw, cin, case['units']))
    w = (0.3 * r.standard_normal(wshape)).astype(H)
    b = (0.1 * r.standard_normal(case['units'])).astype(H)
    z, y = N.conv_like_forward(x, w, b, case['kernel'], case['strides'], case['act'])
    act = {'linear': ACT_NONE, 'relu': ACT_RELU, 'sigmoid': ACT_SIGMOID}[case['act']]
    l = _layer(KIND_DENSE if dense else KIND_CONV, case['units'], case['kernel'], case['strides'], act, out_shape=y.shape[1:])
    xt = _t(x).requires_grad_(True)
    yo = O._conv_like(l, xt, _t(w), _t(b))
    d = ulps(y, yo.detach().numpy())
    assert y.shape == tuple(yo.shape)
    assert d.max() <= 1 and (d > 0).mean() < 2e-3, (d.max(), (d > 0).mean())
    # gradient of the activation alone (ReluGrad / SigmoidGrad), on the torch side through autograd of the same node
    g = r.standard_normal(y.shape).astype(H)
    zt = _t(z).requires_grad_(True)
    yt = torch.relu(zt) if case['act'] == 'relu' else (O.sigmoid_q(zt) if case['act'] == 'sigmoid' else zt * 1.0)
    yt.backward(_t(g))
    want = O.f16(zt.grad).numpy()
    got = N.activation_backward(g, yt.detach().numpy().astype(H), case['act']).astype(np.float32)
    assert np.array_equal(got, want)


@pytest.mark.parametrize('pool,strides', [((2, 2), (2, 2)), ((2, 3), (2, 2)), ((3, 3), (1, 1))])
def test_maxpool_first_maximum_rule_bit_equal(pool, strides):
    """Ties go to the first element of the window in scan order, NaN beats a number, overlapping windows add in window order."""
    r = np.random.default_rng(5)
    x = np.round(r.standard_normal((2, 9, 8, 3)) * 2).astype(H) / H(2)          # many exact ties
    x[0, 0, 0, 0] = np.nan
    x[1, 4, 4, 2] = np.nan
    y, arg = N.maxpool_forward(x, pool, strides)
    xt = _t(x).requires_grad_(True)
    yt = O._MaxPoolFirst.apply(xt, pool[0], pool[1], strides[0], strides[1])
    assert np.array_equal(y.astype(np.float32), yt.detach().numpy(), equal_nan=True)
    g = r.standard_normal(y.shape).astype(H)
    yt.backward(_t(g))
    dx = N.maxpool_backward(g, arg, x.shape, pool, strides)
    assert np.array_equal(dx.astype(np.float32), O.f16(xt.grad).numpy(), equal_nan=True)
