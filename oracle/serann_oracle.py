"""CPU oracle for the SeRANN population-evaluation path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and the CPU-baseline legs of
``bench.py`` may import this module; the product package never does (and fails
loudly when its CUDA library is missing).

PARITY UNPINNED.  The arithmetic of this path lives in third-party code that
is not vendored in the reference and cannot be installed here: tensorflow==2.6.0,
keras==2.6, tensorflow-probability==0.13.0 (reference ``requirements.txt:1-3``).
The reference holds no tests, golden vectors or fixtures for the path
(SURVEY.md section 4).  This file therefore restates the published algorithms
(Keras layer semantics, TF's Adam kernel, the reference's own BatchNormalizationF16)
from a reading of the reference's call sites; every place where TF behaviour was
inferred rather than observed is marked [TF2.6].  The only reference-held known
answers (layer shapes and parameter counts in
``synthetic_serann_generator/tutorial.ipynb:3606-3643``) are checked in
``tests/test_frontend.py``.

Numerics model [TF2.6]: ``floatx=float16`` (reference
``evolutionary_experiment/logic/experiment_worker.py:36-37``) - every tensor and
variable is fp16, matmul/conv accumulate in fp32 and round once, elementwise ops
round after each op.  Implemented as fp32 torch ops between explicit
round-to-fp16 nodes ``q`` whose backward rounds the gradient too.

Reference lines followed:
  model contract          common/logic.py:17-35
  BatchNormalizationF16   common/BatchNormalizationF16.py:81-153
  losses / optimizer      evolutionary_experiment/logic/experiment_worker.py:78-80,269-282
  fit / validation split  experiment_worker.py:101-128
  metrics                 experiment_worker.py:169-195
  replicate               experiment_worker.py:130-167
  threshold / mutations   evolutionary_experiment/logic/experiment.py:210,229
  autoencoder decode      evolutionary_experiment/logic/ribosomal_autoencoder.py:119-126,
                          ribosomal_autoencoder/model.py:35-48
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch
import torch.nn.functional as F

KIND_DENSE, KIND_CONV, KIND_MAXPOOL, KIND_BN, KIND_CONCAT, KIND_RESHAPE = range(6)
ACT_NONE, ACT_RELU, ACT_SIGMOID = 0, 1, 2

BN_EPS = 1e-3            # BatchNormalizationF16.py:13
BN_MOMENTUM = 0.99       # BatchNormalizationF16.py:12
ADAM_LR, ADAM_B1, ADAM_B2, ADAM_EPS = 1e-3, 0.9, 0.999, 1e-4   # experiment_worker.py:80 (Keras defaults for b1, b2)


def f16(x) -> torch.Tensor:
    """Scalar/array -> value rounded to fp16, held in fp32."""
    return torch.as_tensor(x, dtype=torch.float32).half().float()


class _Q(torch.autograd.Function):
    """Round to fp16 in both directions (a tensor boundary in an all-fp16 TF graph)."""

    @staticmethod
    def forward(ctx, x):
        return x.half().float()

    @staticmethod
    def backward(ctx, g):
        return g.half().float()


q = _Q.apply


class _SigmoidQ(torch.autograd.Function):
    """fp16 sigmoid with TF's SigmoidGrad ``dy * y * (1 - y)`` evaluated op by op in fp16 [TF2.6]."""

    @staticmethod
    def forward(ctx, x):
        y = torch.sigmoid(x).half().float()
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        a = (dy * y).half().float()
        b = (1.0 - y).half().float()
        return (a * b).half().float()


sigmoid_q = _SigmoidQ.apply


class _SqDiffQ(torch.autograd.Function):
    """tf.math.squared_difference in fp16 with TF's registered gradient ``scalar_mul(2.0, grad) * (x - y)``, each op
    rounding to fp16 [TF2.6 math_grad._SquaredDifferenceGrad]."""

    @staticmethod
    def forward(ctx, x, y):
        ctx.save_for_backward(x, y)
        return ((x - y) ** 2).half().float()

    @staticmethod
    def backward(ctx, g):
        x, y = ctx.saved_tensors
        g2 = (2.0 * g.half().float()).half().float()
        d = (x - y).half().float()
        return (g2 * d).half().float(), None


def _first_argmax(a: torch.Tensor) -> torch.Tensor:
    """argmax over the last axis, first maximum wins; an all-NaN row gives 0 (as tf.argmax)."""
    a = a.float()
    best = a[:, 0].clone()
    arg = torch.zeros(a.shape[0], dtype=torch.long)
    for j in range(1, a.shape[1]):
        better = a[:, j] > best
        best = torch.where(better, a[:, j], best)
        arg = torch.where(better, torch.full_like(arg, j), arg)
    return arg


class _MaxPoolFirst(torch.autograd.Function):
    """Max pooling whose gradient goes to the first maximum of each window in row-major scan order
    and is summed per input element in window order [TF2.6: MaxPoolGrad tie rule unverified]."""

    @staticmethod
    def forward(ctx, x, ph, pw, sh, sw):      # x: (B, H, W, C)
        B, H, W, C = x.shape
        Ho, Wo = (H - ph) // sh + 1, (W - pw) // sw + 1
        best = torch.full((B, Ho, Wo, C), -float('inf'))
        arg = torch.zeros((B, Ho, Wo, C), dtype=torch.long)
        for i in range(ph):
            for j in range(pw):
                v = x[:, i:i + (Ho - 1) * sh + 1:sh, j:j + (Wo - 1) * sw + 1:sw, :]
                better = (v > best) | (torch.isnan(v) & ~torch.isnan(best))
                best = torch.where(better, v, best)
                arg = torch.where(better, torch.full_like(arg, i * pw + j), arg)
        ctx.save_for_backward(arg)
        ctx.geom = (x.shape, ph, pw, sh, sw)
        return best

    @staticmethod
    def backward(ctx, g):
        (arg,) = ctx.saved_tensors
        shape, ph, pw, sh, sw = ctx.geom
        B, H, W, C = shape
        Ho, Wo = arg.shape[1:3]
        dx = torch.zeros(shape)
        for i in range(ph):
            for j in range(pw):
                sel = (arg == i * pw + j).float() * g
                dx[:, i:i + (Ho - 1) * sh + 1:sh, j:j + (Wo - 1) * sw + 1:sw, :] += sel
        return dx, None, None, None, None


# --------------------------------------------------------------------------- parameters

def glorot_uniform(shape, fan_in, fan_out, rng: np.random.Generator) -> np.ndarray:
    limit = math.sqrt(6.0 / (fan_in + fan_out))
    return rng.uniform(-limit, limit, size=shape).astype(np.float16)


def init_weights(ir, rng: np.random.Generator) -> List[List[np.ndarray]]:
    """Keras default initialisation per layer, fp16, Keras layouts [TF2.6].

    The fused head is initialised as the two separate Dense layers it stands for
    (common/logic.py:29-33): columns [0, n_classes) with fan_out n_classes, the
    rest with fan_out genotype_size.
    """
    out = []
    for l in ir.layers:
        cin = l.in_shape[-1] if l.in_shape else 0
        if l.kind == KIND_DENSE and l.role == 'heads':
            a = glorot_uniform((cin, ir.n_classes), cin, ir.n_classes, rng)
            b = glorot_uniform((cin, ir.genotype_size), cin, ir.genotype_size, rng)
            out.append([np.concatenate([a, b], axis=1), np.zeros(l.units, np.float16)])
        elif l.kind == KIND_DENSE:
            out.append([glorot_uniform((cin, l.units), cin, l.units, rng), np.zeros(l.units, np.float16)])
        elif l.kind == KIND_CONV:
            kh, kw = l.kernel
            rf = kh * kw
            shape = dict(l.param_shapes())['kernel']
            out.append([glorot_uniform(shape, cin * rf, l.units * rf, rng), np.zeros(l.units, np.float16)])
        elif l.kind == KIND_BN:
            c = l.out_shape[-1]
            out.append([np.ones(c, np.float16), np.zeros(c, np.float16),
                        np.zeros(c, np.float16), np.ones(c, np.float16)])
        else:
            out.append([])
    return out


class Individual:
    """One SeRANN on the CPU: fp16-valued fp32 tensors plus fp16 Adam slots."""

    def __init__(self, ir, weights: List[List[np.ndarray]]):
        self.ir = ir
        self.params: List[List[torch.Tensor]] = []
        for l, ws in zip(ir.layers, weights):
            ts = [torch.tensor(np.asarray(w, dtype=np.float16).astype(np.float32)) for w in ws]
            n_train = 2 if l.kind in (KIND_DENSE, KIND_CONV, KIND_BN) else 0
            for t in ts[:n_train]:
                t.requires_grad_(True)
            self.params.append(ts)
        self.m = [[torch.zeros_like(t, dtype=torch.float16) for t in ts[:2]] for ts in self.params]
        self.v = [[torch.zeros_like(t, dtype=torch.float16) for t in ts[:2]] for ts in self.params]

    def trainable(self):
        for li, (l, ts) in enumerate(zip(self.ir.layers, self.params)):
            if l.kind in (KIND_DENSE, KIND_CONV, KIND_BN):
                for pi in range(2):
                    yield li, pi, ts[pi]

    def export(self) -> List[List[np.ndarray]]:
        return [[t.detach().numpy().astype(np.float16) for t in ts] for ts in self.params]


# --------------------------------------------------------------------------- layers

def _conv_like(l, x, w, b):
    """Dense on the last axis / Conv1D / Conv2D, 'valid' padding; x: (B, ..., C) fp16-valued."""
    if l.kind == KIND_DENSE:
        z = q(x @ w)                                   # fp32 accumulate, one rounding [TF2.6]
    else:
        kh, kw = l.kernel
        sh, sw = l.strides
        cin = x.shape[-1]
        x4 = x.reshape(x.shape[0], 1, x.shape[1], cin) if x.dim() == 3 else x       # Conv1D: H = 1
        w4 = w.reshape(kh, kw, cin, l.units).permute(3, 2, 0, 1)                    # HWIO -> OIHW
        z4 = F.conv2d(x4.permute(0, 3, 1, 2), w4, stride=(sh, sw)).permute(0, 2, 3, 1)
        z = q(z4.reshape((x.shape[0],) + tuple(l.out_shape)))
    z = q(z + b)                                       # separate BiasAdd op in fp16 [TF2.6]
    if l.act == ACT_RELU:
        return torch.relu(z)
    if l.act == ACT_SIGMOID:
        return sigmoid_q(z)
    return z


def _bn_scale_shift(mean16, var16, gamma, beta):
    """tf.nn.batch_normalization in fp16: inv = rsqrt(var + eps) * gamma; shift = beta - mean * inv."""
    inv = q(torch.rsqrt(q(var16 + f16(BN_EPS))))
    inv = q(inv * gamma)
    shift = q(beta - q(mean16 * inv))
    return inv, shift


class _BNTrainCentered(torch.autograd.Function):
    """Training-mode normalisation with the gradient in its centred closed form

        dx = inv (dy - mean(dy)) - (x - mean32) gamma inv_raw^3 mean(dy (x - mean16)),  dgamma = inv_raw sum dy (x - mean16),
        dbeta = sum dy

    (reductions in fp32, one rounding of dx at the producer's boundary).  This is what differentiating the forward
    graph gives - the variance is taken around the fp32 mean, the shift uses the fp16 mean - but with the two terms
    of d inv = sum(dy x) - mean16 sum(dy) combined before any rounding.  TF's op-by-op graph (and ``_bn_train_graph``
    below) rounds sum(dy) to fp16 first and multiplies by mean16 afterwards, which on channels whose spread is small
    against the fp16 spacing of their mean (|mean| / std in the hundreds: features of near-constant inputs) leaves
    cancellation noise of relative size 2^-11 |mean| / std in every upstream gradient.  The kernels and this oracle use
    the centred form; the deviation from the TF graph is documented in DESIGN.md section 2 and measured in
    tests/test_oracle_numpy.py."""

    @staticmethod
    def forward(ctx, x, gamma, beta):
        dims = tuple(range(x.dim() - 1))
        n = 1
        for d in dims:
            n *= x.shape[d]
        mean32 = x.mean(dims)
        var32 = ((x - mean32) ** 2).mean(dims)
        mean16, var16 = f16(mean32), f16(var32)
        inv_raw = f16(torch.rsqrt(f16(var16 + f16(BN_EPS))))
        inv = f16(inv_raw * gamma)
        shift = f16(beta - f16(mean16 * inv))
        y = f16(f16(x * inv) + shift)
        ctx.save_for_backward(x, mean32, mean16, inv_raw, inv, gamma)
        ctx.n = n
        ctx.mark_non_differentiable(mean16, var16)
        return y, mean16, var16

    @staticmethod
    def backward(ctx, gy, _gm, _gv):
        x, mean32, mean16, inv_raw, inv, gamma = ctx.saved_tensors
        dims = tuple(range(x.dim() - 1))
        g = gy.half().float()
        s1 = g.double().sum(dims)
        s2 = (g.double() * (x.double() - mean16.double())).sum(dims)
        coef = gamma.double() * inv_raw.double() ** 3 * s2 / ctx.n
        dx = inv.double() * (g.double() - s1 / ctx.n) - (x.double() - mean32.double()) * coef
        return dx.float(), (s2 * inv_raw.double()).float(), s1.float()


def _bn_train_graph(x, gamma, beta):
    """The same forward as a plain autograd graph (TF's op structure with one rounding per tensor boundary): kept for
    the cross-check against the NumPy op-by-op restatement."""
    dims = tuple(range(x.dim() - 1))
    mean32 = x.mean(dims)
    var32 = ((x - mean32.detach()) ** 2).mean(dims)
    mean16, var16 = q(mean32), q(var32)
    inv, shift = _bn_scale_shift(mean16, var16, gamma, beta)
    return q(q(x * inv) + shift), mean16.detach(), var16.detach()


BN_BACKWARD = 'centered'        # 'graph' differentiates the op graph instead (see _BNTrainCentered)


def _bn_train(x, gamma, beta, mov_mean, mov_var):
    """BatchNormalizationF16.call, training branch (BatchNormalizationF16.py:130-153).

    moments in fp32 for fp16 input then cast back [TF2.6]; the value fed to the
    moving variance is var * n / (n - (1 + eps)) with n cast to fp16 first, which
    is inf/inf = NaN once B * spatial >= 65520.
    """
    dims = tuple(range(x.dim() - 1))
    if BN_BACKWARD == 'centered':
        y, mean16, var16 = _BNTrainCentered.apply(x, gamma, beta)
    else:
        y, mean16, var16 = _bn_train_graph(x, gamma, beta)
    with torch.no_grad():
        n = 1
        for d in dims:
            n *= x.shape[d]
        n16 = f16(float(n))                                          # inf above 65504
        bessel = f16(n16 / f16(n16 - f16(1.0 + BN_EPS)))             # NaN when n16 is inf
        var_unbiased = f16(var16 * bessel)
        # [TF2.6] keras/backend.py moving_average_update, TF2 branch: x.assign(x * momentum + value * (1 - momentum))
        # with momentum cast to the variable's dtype (fp16(0.99) = 0.990234375, 1 - that = 0.009765625 exactly).  TF <= 2.4
        # used assign_sub((x - value) * (1 - momentum)) instead; the two differ in decay and in rounding order.
        mom = f16(BN_MOMENTUM)
        one_minus = f16(1.0 - mom)
        mov_mean.copy_(f16(f16(mov_mean * mom) + f16(mean16 * one_minus)))
        mov_var.copy_(f16(f16(mov_var * mom) + f16(var_unbiased * one_minus)))
    return y


def _bn_infer(x, gamma, beta, mov_mean, mov_var):
    inv, shift = _bn_scale_shift(mov_mean, mov_var, gamma, beta)
    return q(q(x * inv) + shift)


def forward(ind: Individual, x_img: torch.Tensor, g_in: torch.Tensor, training: bool,
            tape: Optional[Dict[int, torch.Tensor]] = None) -> torch.Tensor:
    """Returns the fused head logits (B, n_classes + genotype_size), fp16-valued.

    ``tape`` (optional) receives every layer output by tensor id, with ``retain_grad`` set, so the
    parity tests can compare activations and activation gradients layer by layer."""
    t: Dict[int, torch.Tensor] = {0: x_img, 1: g_in}
    for i, (l, ps) in enumerate(zip(ind.ir.layers, ind.params)):
        x = t[l.inputs[0]]
        if l.kind in (KIND_DENSE, KIND_CONV):
            y = _conv_like(l, x, ps[0], ps[1])
        elif l.kind == KIND_MAXPOOL:
            y = _MaxPoolFirst.apply(x, l.kernel[0], l.kernel[1], l.strides[0], l.strides[1])
        elif l.kind == KIND_BN:
            y = (_bn_train if training else _bn_infer)(x, *ps)
        elif l.kind == KIND_CONCAT:
            y = torch.cat([t[j] for j in l.inputs], dim=-1)
        elif l.kind == KIND_RESHAPE:
            y = x.reshape((x.shape[0],) + tuple(l.out_shape))
        else:
            raise ValueError(l.kind)
        t[2 + i] = y
        if tape is not None:
            if y.requires_grad:
                y.retain_grad()
            tape[2 + i] = y
    return t[1 + len(ind.ir.layers)].reshape(x_img.shape[0], -1)


# --------------------------------------------------------------------------- loss, metrics, optimizer

def losses(logits: torch.Tensor, labels: torch.Tensor, bits: torch.Tensor, n_classes: int):
    """Per-sample categorical cross-entropy (from logits [TF2.6: Keras uses the cached logits of a
    softmax Activation]) and per-sample MSE of the sigmoid output (experiment_worker.py:276-277)."""
    zc, zr = logits[:, :n_classes], logits[:, n_classes:]
    ce = q(torch.logsumexp(zc, dim=1) - zc.gather(1, labels[:, None]).squeeze(1))
    s = sigmoid_q(zr)
    mse = q(_SqDiffQ.apply(s, bits).mean(dim=1))
    # categorical_accuracy compares argmax(y_true) with argmax(y_pred) where y_pred is the model's fp16 softmax
    # OUTPUT [TF2.6 keras/metrics.py]: near-equal top probabilities can round to a tie, the first index wins
    with torch.no_grad():
        zmax = zc.max(dim=1, keepdim=True).values
        e = torch.exp(zc - zmax)
        y_pred = (e / e.sum(dim=1, keepdim=True)).half()
        correct = (_first_argmax(y_pred) == labels)
    return ce, mse, correct, s


def weighted_loss(ce, mse, loss_balance: float):
    """loss_weights lb and 1 - lb (experiment_worker.py:278-279), sum_over_batch_size reduction."""
    lc, lr = q(ce.mean()), q(mse.mean())
    w_c, w_r = f16(loss_balance), f16(1.0 - f16(loss_balance))
    return q(q(w_c * lc) + q(w_r * lr)), lc, lr


def adam_step(ind: Individual, t: int):
    """TF's ApplyAdam kernel with T = half: every scalar and intermediate is fp16 [TF2.6]."""
    H = torch.float16
    lr, b1, b2, eps = (torch.tensor(v, dtype=H) for v in (ADAM_LR, ADAM_B1, ADAM_B2, ADAM_EPS))
    one = torch.tensor(1.0, dtype=H)
    step = torch.tensor(float(t), dtype=H)
    b1p = torch.pow(b1.float(), step.float()).to(H)
    b2p = torch.pow(b2.float(), step.float()).to(H)
    alpha = (lr * torch.sqrt(one - b2p)) / (one - b1p)          # fp16 op by op
    for li, pi, p in ind.trainable():
        g = p.grad.to(H) if p.grad is not None else torch.zeros_like(p, dtype=H)
        m, v = ind.m[li][pi], ind.v[li][pi]
        m += (one - b1) * (g - m)
        v += (one - b2) * (g * g - v)
        upd = (alpha * m) / (eps + torch.sqrt(v))
        with torch.no_grad():
            p.copy_((p.to(H) - upd).float())
        p.grad = None


class F16Mean:
    """Keras ``Mean`` metric with fp16 ``total`` and ``count`` variables [TF2.6]."""

    def __init__(self):
        self.total = np.float16(0)
        self.count = np.float16(0)

    def update(self, value_sum, n):
        self.total = np.float16(self.total + np.float16(value_sum))
        self.count = np.float16(self.count + np.float16(n))

    def result(self) -> float:
        with np.errstate(all='ignore'):
            return float(np.float16(self.total / self.count)) if self.count != 0 else 0.0


def prepare_inputs(x: np.ndarray, g_bits: np.ndarray):
    """Model-boundary casts: images float32 -> fp16 [TF2.6]; genotype ints -> fp16 (B, 100, 1)."""
    xi = torch.tensor(np.asarray(x, dtype=np.float32)).half().float()
    if xi.dim() == 3:
        xi = xi[..., None]
    gi = torch.tensor(np.asarray(g_bits, dtype=np.float32)).half().float()[..., None]
    return xi, gi


def train_step(ind: Individual, xi, gi, labels, bits, t: int, tape=None, apply_adam: bool = True):
    logits = forward(ind, xi, gi, training=True, tape=tape)
    ce, mse, correct, _ = losses(logits, labels, bits, ind.ir.n_classes)
    total, lc, lr = weighted_loss(ce, mse, ind.ir.loss_balance)
    total.backward()
    if apply_adam:
        adam_step(ind, t)
    return {'loss': float(total.detach()), 'ce': float(lc.detach()), 'mse': float(lr.detach()),
            'correct': int(correct.sum()), 'logits': logits.detach()}


def evaluate(ind: Individual, x, g_bits, labels, chunk: int):
    """Inference-mode pass with Keras' per-batch fp16 metric accumulation (batch = ``chunk``)."""
    acc, mse_m = F16Mean(), F16Mean()
    labels_t = torch.tensor(np.asarray(labels, dtype=np.int64))
    bits_t = torch.tensor(np.asarray(g_bits, dtype=np.float32))
    with torch.no_grad():
        for s in range(0, len(x), chunk):
            xi, gi = prepare_inputs(x[s:s + chunk], g_bits[s:s + chunk])
            logits = forward(ind, xi, gi, training=False)
            ce, mse, correct, _ = losses(logits, labels_t[s:s + chunk], bits_t[s:s + chunk], ind.ir.n_classes)
            n = len(xi)
            acc.update(float(correct.sum()), n)
            mse_m.update(float(np.float16(float(q(mse.mean())) * n)), n)
    return acc.result(), mse_m.result()


def fit(inds: Sequence[Individual], x_train, y_train, g_train, perms: Sequence[np.ndarray],
        batch_size: int, n_val: int, record_steps: int = 0):
    """``Model.fit(batch_size, epochs, validation_split)`` (experiment_worker.py:113-118).

    The last ``n_val`` rows are the validation set, unshuffled; ``perms[e]`` is the
    row order of epoch ``e`` over the first ``len(x) - n_val`` rows (Keras shuffles
    with an unseeded RNG; the order is an explicit input here).  One shared Adam
    step counter.  Returns per-individual dicts of the last epoch's metrics.
    """
    n_fit = len(x_train) - n_val
    labels_all = torch.tensor(np.asarray(y_train, dtype=np.int64))
    bits_all = torch.tensor(np.asarray(g_train, dtype=np.float32))
    t = 0
    out = [dict(steps=[]) for _ in inds]
    for perm in perms:
        accs = [F16Mean() for _ in inds]
        for s in range(0, n_fit, batch_size):
            idx = np.asarray(perm[s:s + batch_size])
            xi, gi = prepare_inputs(x_train[idx], g_train[idx])
            t += 1
            for k, ind in enumerate(inds):
                r = train_step(ind, xi, gi, labels_all[idx], bits_all[idx], t)
                accs[k].update(r['correct'], len(idx))
                if t <= record_steps:
                    out[k]['steps'].append({kk: r[kk] for kk in ('loss', 'ce', 'mse', 'correct')})
        for k, ind in enumerate(inds):
            out[k]['classification_training_accuracy'] = accs[k].result()
            if n_val:
                va, vm = evaluate(ind, x_train[n_fit:], g_train[n_fit:], y_train[n_fit:], batch_size)
                out[k]['classification_validation_accuracy'] = va
                out[k]['replication_mse'] = vm
    return out


def replicate(inds: Sequence[Individual], genotypes: np.ndarray, x_test: np.ndarray, R: int) -> List[np.ndarray]:
    """experiment_worker.py:130-167: row ``i*R + r`` pairs individual i's own genotype with test
    image ``(i*R + r) // ceil(N*R / len(x_test))``; inference mode; returns (R, 100) fp16 per individual."""
    n = len(inds)
    repeat = int(np.ceil(n * R / len(x_test)))
    out = []
    with torch.no_grad():
        for i, ind in enumerate(inds):
            rows = (np.arange(i * R, (i + 1) * R) // repeat)
            xi, gi = prepare_inputs(x_test[rows], np.repeat(genotypes[i][None], R, axis=0))
            logits = forward(ind, xi, gi, training=False)
            s = sigmoid_q(logits[:, ind.ir.n_classes:])
            out.append(s.numpy().astype(np.float16))
    return out


def readout(probs: np.ndarray, parent_bits: np.ndarray):
    """np.round(np.clip(p, 0, 1)) (experiment.py:210; half-to-even, so bit = p > 0.5) and the
    per-offspring mutation count vs the parent (experiment.py:229)."""
    p = np.asarray(probs, dtype=np.float16)
    bits = np.round(np.clip(p, 0, 1))
    parent = np.asarray(parent_bits).astype(bits.dtype)
    mutations = (bits != parent).sum(axis=-1)
    return bits, mutations.astype(np.int32)


# --------------------------------------------------------------------------- ribosomal autoencoder decode

def _bn_affine_f32(x, w, name):
    """tf.nn.batch_normalization, inference: x * inv + (beta - mean * inv), inv = rsqrt(var + eps) * gamma."""
    f = lambda k: torch.tensor(np.asarray(w[f'{name}_{k}'], dtype=np.float32))
    inv = (1.0 / torch.sqrt(f('var') + np.float32(1e-3))) * f('gamma')
    return x * inv + (f('beta') - f('mean') * inv)


def ae_decode(genotypes: np.ndarray, w: Dict[str, np.ndarray], max_tokens: int, vocab: int) -> np.ndarray:
    """fp32 ``generative_net`` in inference mode + argmax (ribosomal_autoencoder.py:122-124;
    model.py:35-48): one_hot -> Conv1D(32, 5) -> BN -> Flatten -> Dense -> Reshape -> BN -> argmax.
    Keras BatchNormalization defaults: eps 1e-3.  ``np.eye(2)[nan.astype(int)]`` raises for NaN
    genotypes in the reference; here a ValueError.
    """
    gen = np.asarray(genotypes)
    if np.isnan(gen.astype(np.float64)).any():
        raise ValueError('NaN genotype')
    with torch.no_grad():
        oh = torch.tensor(np.eye(2, dtype=np.float32)[gen.astype(int)])               # (n, 100, 2)
        k = torch.tensor(w['conv_kernel'], dtype=torch.float32)                       # (5, 2, 32)
        h = F.conv1d(oh.permute(0, 2, 1), k.permute(2, 1, 0)) + torch.tensor(w['conv_bias'])[None, :, None]
        h = h.permute(0, 2, 1)                                                        # (n, 96, 32)
        h = _bn_affine_f32(h, w, 'bn1')
        z = h.reshape(len(gen), -1) @ torch.tensor(w['dense_kernel']) + torch.tensor(w['dense_bias'])
        z = z.reshape(len(gen), max_tokens, vocab)
        z = _bn_affine_f32(z, w, 'bn2')
        return z.argmax(dim=-1).numpy().astype(np.int64), z.numpy()
