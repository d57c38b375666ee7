"""The CPU oracle against closed-form expectations and the committed fixtures (it has no TF to be checked against:
parity is unpinned, see oracle/serann_oracle.py)."""
import pathlib

import numpy as np
import torch

from oracle import serann_oracle as O
from serann_b200 import frontend

GOLDEN = pathlib.Path(__file__).resolve().parent / 'golden'


def test_readout_is_round_half_even_of_clip():
    p = np.array([[0.5, 0.501, 0.499, 1.7, -0.3, np.nan]], dtype=np.float16)
    bits, mut = O.readout(p, np.array([[0, 0, 0, 0, 0, 0]]))
    assert bits[0, :5].tolist() == [0, 1, 0, 1, 0] and np.isnan(bits[0, 5])
    assert mut[0] == 3          # two set bits + the NaN locus (nan != 0)


def test_bn_moving_variance_nan_above_fp16_range():
    """sample_size cast to fp16 is inf for B * spatial > 65504 -> Bessel factor NaN."""
    x = torch.randn(70000, 2).half().float()
    mm, mv = torch.zeros(2), torch.ones(2)
    y = O._bn_train(x, torch.ones(2), torch.zeros(2), mm, mv)
    assert torch.isfinite(y).all() and torch.isnan(mv).all() and torch.isfinite(mm).all()
    x = torch.randn(750, 2).half().float()
    mm, mv = torch.zeros(2), torch.ones(2)
    O._bn_train(x, torch.ones(2), torch.zeros(2), mm, mv)
    assert torch.isfinite(mv).all()
    # momentum 0.99 in fp16: one update moves the variance by about 1 % of the gap
    assert abs(float(mv[0]) - (1 - 0.01 * (1 - float(x[:, 0].var())))) < 2e-3


def test_adam_v_underflows_for_small_gradients():
    """(1 - beta2) * g^2 flushes to zero in fp16 for small g, so the step is alpha * m / eps."""
    ir = frontend.compile_source("con = X_layer\nloss_balance = 0.5")
    ind = O.Individual(ir, O.init_weights(ir, np.random.default_rng(0)))
    for _, _, p in ind.trainable():
        p.grad = torch.full_like(p, 1e-3)
    w0 = ind.params[-1][0].detach().clone()
    O.adam_step(ind, 1)
    assert float(ind.v[-1][0].abs().max()) == 0.0
    step = (w0 - ind.params[-1][0].detach()).abs().max()
    assert 5e-4 < float(step) < 5e-2


def test_oracle_gradients_match_fp32_autograd_on_a_small_net():
    """With the rounding nodes in place the gradients stay within fp16 resolution of a plain fp32 graph."""
    src = ("X_layer = Conv2D(filters=4, kernel_size=5, strides=2)(X_layer)\nX_layer = MaxPool2D(pool_size=2)(X_layer)\n"
           "g_layer = Dense(units=6, activation='relu')(g_layer)\n"
           "con = concatenate([Reshape((1, -1))(X_layer), Reshape((1, -1))(g_layer)])\ncon = Dense(units=16, activation='sigmoid')(con)\nloss_balance = 0.6")
    ir = frontend.compile_source(src)
    r = np.random.default_rng(1)
    w = O.init_weights(ir, r)
    ind = O.Individual(ir, w)
    x = r.random((12, 28, 28), dtype=np.float32)
    g = r.integers(0, 2, (12, 100))
    y = torch.tensor(r.integers(0, 10, 12))
    xi, gi = O.prepare_inputs(x, g)
    O.train_step(ind, xi, gi, y, torch.tensor(g.astype(np.float32)), 1, apply_adam=False)
    # plain fp32 reference of the same network
    import torch.nn.functional as F
    ws = [[torch.tensor(a.astype(np.float32), requires_grad=True) for a in l] for l in w]
    h = F.conv2d(xi.permute(0, 3, 1, 2), ws[0][0].permute(3, 2, 0, 1), ws[0][1], stride=2)
    h = F.max_pool2d(h, 2).permute(0, 2, 3, 1).reshape(12, -1)
    gg = torch.relu(gi @ ws[3][0] + ws[3][1]).reshape(12, -1)
    con = torch.sigmoid(torch.cat([h, gg], 1) @ ws[6][0] + ws[6][1])
    z = con @ ws[8][0] + ws[8][1]
    loss = 0.6 * F.cross_entropy(z[:, :10], y) + 0.4 * ((torch.sigmoid(z[:, 10:]) - torch.tensor(g.astype(np.float32))) ** 2).mean()
    loss.backward()
    for li in (0, 3, 6, 8):
        a, b = ind.params[li][0].grad, ws[li][0].grad
        assert float((a - b).abs().max() / b.abs().max()) < 3e-2, li


def test_golden_fixture_reproduces():
    """tests/golden/make_golden.py wrote these with this oracle; a drift in the restatement shows up here."""
    from make_golden import build_case
    z = np.load(GOLDEN / 'oracle_small.npz')
    case = build_case()
    assert np.array_equal(case['logits_step0'], z['logits_step0'])
    assert np.allclose(case['losses'], z['losses'], rtol=0, atol=0)
    assert np.array_equal(case['offspring'], z['offspring'])
