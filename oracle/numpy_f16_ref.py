"""Second CPU restatement of the fp16 arithmetic of the SeRANN path, op by op in NumPy float16.
TEST INFRASTRUCTURE ONLY (same rule as ``serann_oracle.py``: only ``tests/`` may import it).

PARITY UNPINNED - this file does not change that: TensorFlow 2.6 / Keras 2.6 are not installable
here.  Its purpose is narrower: ``serann_oracle.py`` expresses the path as torch fp32 ops between
rounding nodes with autograd; this file restates the same reference lines a second time with *no
torch and no autograd* - every tensor is a ``numpy.float16`` array, every TF op is one NumPy op on
float16 (NumPy computes a float16 op in float32 and rounds once, which is what an Eigen::half /
CUDA __half op does), and the gradients of BatchNormalization and of the two losses are written out
by hand from TF's registered gradient functions.  ``tests/test_oracle_numpy.py`` asserts that the two
restatements agree bit for bit on seeded inputs, so a slip in either reading shows up as a mismatch.

Reference lines and the TF/Keras code each function follows ([TF2.6] = read from TF/Keras sources
from memory, to be re-checked against the named file when an install exists):

  bn_train / bn_infer   common/BatchNormalizationF16.py:81-153;
                        [TF2.6] keras/backend.py _regular_normalize_batch_in_training, moving_average_update;
                        tensorflow/python/ops/nn_impl.py moments (fp16 input -> fp32 moments -> cast back),
                        batch_normalization (rsqrt(var + eps) * scale; x * inv + (offset - mean * inv))
  bn_train_backward     [TF2.6] gradients of the ops above, one rounding per gradient op
  adam_step             experiment_worker.py:78-80; [TF2.6] keras/optimizer_v2/adam.py _prepare_local +
                        tensorflow/core/kernels/training_ops_gpu.cu.cc ApplyAdamKernel<Eigen::half>
  losses_and_grads      experiment_worker.py:269-282; [TF2.6] keras/backend.py categorical_crossentropy (cached logits
                        of a softmax Activation), nn_ops.softmax_cross_entropy_with_logits_v2 (fp16 logits are cast to
                        fp32, cost cast back), keras/losses.py mean_squared_error, math_grad._SquaredDifferenceGrad,
                        SigmoidGrad, keras/utils/losses_utils.py sum_over_batch_size
  conv_like_forward     common/logic.py:17-35 (Keras layers as built); [TF2.6] conv_ops / matmul_op (fp32 accumulation, one
                        rounding), bias_op, ReluGrad / SigmoidGrad (activation_backward)
  maxpool_forward/bwd   [TF2.6] maxpooling_op (first maximum of a window receives the gradient - tie rule unverified)
  F16Mean / accuracy    experiment_worker.py:169-195,280; [TF2.6] keras/metrics.py Mean with floatx variables,
                        categorical_accuracy = argmax(y_true) == argmax(y_pred) on the fp16 softmax output
"""
from __future__ import annotations

import numpy as np

H = np.float16
F = np.float32

BN_EPS = H(1e-3)                    # BatchNormalizationF16.py:13 as an fp16 graph constant
BN_MOMENTUM = H(0.99)               # BatchNormalizationF16.py:12, tf.cast(momentum, x.dtype)
ONE = H(1.0)


def _sum32(a: np.ndarray, axis=None) -> np.ndarray:
    """A reduction TF runs with fp32 accumulators on fp16/fp32 data; accumulated in float64 here so that the
    result does not depend on NumPy's summation order, then rounded to fp32 once."""
    return np.asarray(np.sum(a.astype(np.float64), axis=axis)).astype(F)


# ------------------------------------------------------------------ BatchNormalizationF16

def bn_scale_shift(mean16, var16, gamma, beta):
    """tf.nn.batch_normalization's coefficients, every op in fp16."""
    with np.errstate(all='ignore'):
        u = (var16 + BN_EPS).astype(H)
        inv_raw = (F(1.0) / np.sqrt(u.astype(F))).astype(H)            # Rsqrt<half>: fp32 inside, one rounding
        inv = (inv_raw * gamma).astype(H)
        shift = (beta - (mean16 * inv).astype(H)).astype(H)
    return u, inv_raw, inv, shift


def bn_moments(x16: np.ndarray):
    """nn_impl.moments on fp16 input: cast to fp32, mean, mean of squared differences, cast back."""
    axes = tuple(range(x16.ndim - 1))
    n = int(np.prod([x16.shape[a] for a in axes]))
    x32 = x16.astype(F)
    mean32 = (_sum32(x32, axes).astype(np.float64) / n).astype(F)
    d = x32.astype(np.float64) - mean32.astype(np.float64)
    var32 = (np.sum(d * d, axis=axes) / n).astype(F)
    return mean32, var32, n


def moving_average_update(x16, value16, momentum16=BN_MOMENTUM):
    """[TF2.6] keras/backend.py moving_average_update, TF2 branch: x.assign(x * momentum + value * (1 - momentum))
    with momentum and value cast to the variable's dtype."""
    with np.errstate(all='ignore'):
        one_minus = (ONE - momentum16).astype(H)
        return ((x16 * momentum16).astype(H) + (value16 * one_minus).astype(H)).astype(H)


def bn_train(x16, gamma, beta, mov_mean, mov_var):
    """Training branch of BatchNormalizationF16.call.  Returns y, new moving mean / variance and the saved values the
    backward needs."""
    mean32, var32, n = bn_moments(x16)
    mean16, var16 = mean32.astype(H), var32.astype(H)
    u, inv_raw, inv, shift = bn_scale_shift(mean16, var16, gamma, beta)
    with np.errstate(all='ignore'):
        y = ((x16 * inv).astype(H) + shift).astype(H)
        n16 = H(n)                                                     # K.cast(sample_size, fp16): inf above 65504
        bessel = (n16 / (n16 - H(1.0 + 1e-3)).astype(H)).astype(H)     # inf / inf = NaN
        var_unbiased = (var16 * bessel).astype(H)
    new_mean = moving_average_update(mov_mean, mean16)
    new_var = moving_average_update(mov_var, var_unbiased)
    saved = dict(x=x16, mean32=mean32, mean16=mean16, u=u, inv_raw=inv_raw, inv=inv, gamma=gamma, n=n)
    return y, new_mean, new_var, saved


def bn_infer(x16, gamma, beta, mov_mean, mov_var):
    _, _, inv, shift = bn_scale_shift(mov_mean, mov_var, gamma, beta)
    with np.errstate(all='ignore'):
        return ((x16 * inv).astype(H) + shift).astype(H)


def bn_train_backward(dy16, s):
    """Gradients of y = x * inv + shift wrt x, gamma, beta through the batch moments, one fp16 rounding per
    gradient op, reductions with fp32 accumulators."""
    x16, inv, inv_raw, gamma, mean16, u, n = s['x'], s['inv'], s['inv_raw'], s['gamma'], s['mean16'], s['u'], s['n']
    axes = tuple(range(x16.ndim - 1))
    with np.errstate(all='ignore'):
        g = dy16.astype(H)
        dx = (g * inv).astype(H)                                       # Mul grad wrt x
        d_inv = _sum32((g.astype(F) * x16.astype(F)), axes).astype(H)   # Mul grad wrt inv: sum(dy * x)
        d_shift = _sum32(g.astype(F), axes).astype(H)                  # Add grad wrt shift: sum(dy)
        d_beta = d_shift
        d_minv = (-d_shift).astype(H)                                  # shift = beta - mean * inv
        d_mean16 = (d_minv * inv).astype(H)
        d_inv = (d_inv + (d_minv * mean16).astype(H)).astype(H)        # AddN of the two paths into inv
        d_gamma = (d_inv * inv_raw).astype(H)
        d_inv_raw = (d_inv * gamma).astype(H)
        # RsqrtGrad(y, dy) = -0.5 * dy * y^3
        d_u = (F(-0.5) * d_inv_raw.astype(F) * inv_raw.astype(F) ** 3).astype(H)
        d_var32 = d_u.astype(F)                                         # Add eps, Cast fp32 -> fp16: gradient casts back
        d_mean32 = d_mean16.astype(F)
        x32 = x16.astype(F)
        # variance = mean(squared_difference(x32, stop_gradient(mean))); mean = reduce_mean(x32)
        dx32 = (d_var32 / F(n)) * F(2.0) * (x32 - s['mean32']) + d_mean32 / F(n)
        dx = (dx.astype(F) + dx32.astype(H).astype(F)).astype(H)       # Cast grad (fp32 -> fp16), then AddN with the direct path
    return dx, d_gamma, d_beta


# ------------------------------------------------------------------ Adam

def adam_scalars(t: int, lr=1e-3, b1=0.9, b2=0.999, eps=1e-4):
    lr, b1, b2, eps = H(lr), H(b1), H(b2), H(eps)
    step = H(float(t))
    b1p = H(np.power(F(b1), F(step)))
    b2p = H(np.power(F(b2), F(step)))
    with np.errstate(all='ignore'):
        alpha = H(H(lr * H(np.sqrt(F(H(ONE - b2p))))) / H(ONE - b1p))
    return alpha, H(ONE - b1), H(ONE - b2), eps


def adam_step(p, m, v, g, t: int):
    """ApplyAdam with T = half on fp16 arrays (g already rounded to fp16); returns new p, m, v."""
    alpha, omb1, omb2, eps = adam_scalars(t)
    with np.errstate(all='ignore'):
        m = (m + (omb1 * (g - m).astype(H)).astype(H)).astype(H)
        v = (v + (omb2 * ((g * g).astype(H) - v).astype(H)).astype(H)).astype(H)
        upd = ((alpha * m).astype(H) / (eps + np.sqrt(v.astype(F)).astype(H)).astype(H)).astype(H)
        p = (p - upd).astype(H)
    return p, m, v


# ------------------------------------------------------------------ losses, metrics

def sigmoid16(z16):
    with np.errstate(all='ignore'):
        return (F(1.0) / (F(1.0) + np.exp(-z16.astype(F)))).astype(H)


def losses_and_grads(logits16: np.ndarray, labels: np.ndarray, bits: np.ndarray, n_classes: int, loss_balance: float):
    """Per-sample CE and MSE (fp16), the weighted total, correct-count and d total / d logits (fp16)."""
    B = logits16.shape[0]
    zc = logits16[:, :n_classes].astype(F)
    zr = logits16[:, n_classes:]
    G = zr.shape[1]
    with np.errstate(all='ignore'):
        zmax = zc.max(axis=1, keepdims=True)
        e = np.exp(zc - zmax)
        se = e.sum(axis=1, keepdims=True, dtype=F)
        lse = (zmax + np.log(se))[:, 0]
        ce = (lse - zc[np.arange(B), labels]).astype(H)                 # fp32 cost, Cast to fp16
        softmax32 = e / se
        y_pred16 = softmax32.astype(H)                                   # the model's fp16 softmax output
        correct = int((np.argmax(y_pred16, axis=1) == labels).sum())   # first maximum wins
        s = sigmoid16(zr)
        t16 = bits.astype(H)
        sq = ((s.astype(F) - t16.astype(F)) ** 2).astype(H)            # SquaredDifference<half>: one rounding
        mse = (_sum32(sq.astype(F), 1) / F(G)).astype(H)
        lc = (_sum32(ce.astype(F)) / F(B)).astype(H)                    # sum_over_batch_size
        lr_ = (_sum32(mse.astype(F)) / F(B)).astype(H)
        wc = H(loss_balance)
        wr = H(ONE - wc)
        total = H(H(wc * lc) + H(wr * lr_))
        # backward
        gc = H(wc / H(B))                                               # d total / d ce_b
        onehot = np.zeros_like(softmax32)
        onehot[np.arange(B), labels] = 1
        d_zc = (gc.astype(F) * (softmax32 - onehot)).astype(H)          # fused op's backprop in fp32, Cast to fp16
        gm = H(H(wr / H(B)) / H(G))                                     # d total / d sq_bj
        two_g = H(F(2.0) * gm)
        diff16 = (s - t16).astype(H)                                    # Sub<half>
        dy = (two_g * diff16).astype(H)
        d_zr = ((dy * s).astype(H) * (ONE - s).astype(H)).astype(H)    # SigmoidGrad<half>
    return dict(ce=ce, mse=mse, lc=lc, lr=lr_, total=total, correct=correct, probs=s,
                dlogits=np.concatenate([d_zc, d_zr], axis=1))


class F16Mean:
    """keras.metrics.Mean with fp16 total / count variables."""

    def __init__(self):
        self.total, self.count = H(0), H(0)

    def update(self, value_sum, n):
        with np.errstate(all='ignore'):
            self.total = H(self.total + H(value_sum))
            self.count = H(self.count + H(n))

    def result(self) -> float:
        with np.errstate(all='ignore'):
            return float(H(self.total / self.count)) if self.count != 0 else 0.0


# --------------------------------------------------------------------------- Dense / Conv1D / Conv2D and max pooling
# experiment_worker.py builds the layers with Keras defaults: 'valid' padding, channels last, kernels in HWIO layout
# (Conv1D: (k, in, out)), Dense on the last axis.  [TF2.6] one rounding per op: contraction (fp32 accumulation) -> fp16,
# BiasAdd -> fp16, activation -> fp16.  The contraction is an explicit im2col + matmul here (torch's conv2d in the other
# restatement), so the two agree up to the fp32 summation order: equal after the fp16 rounding except for rare last-bit ties.

def conv_like_forward(x16: np.ndarray, w16: np.ndarray, b16: np.ndarray, kernel=(1, 1), strides=(1, 1), act: str = 'linear'):
    """x16 (B, H, W, C) [Conv1D: (B, W, C), Dense: (B, ..., C)], w16 Keras layout, returns (z, y): pre-activation and output."""
    x = np.asarray(x16, dtype=H)
    kh, kw = kernel
    sh, sw = strides
    if kh == 1 and kw == 1 and w16.ndim == 2:                      # Dense on the last axis
        acc = x.astype(F).reshape(-1, x.shape[-1]) @ np.asarray(w16, dtype=H).astype(F)
        z = acc.astype(H).reshape(x.shape[:-1] + (w16.shape[-1],))
    else:
        one_d = x.ndim == 3
        x4 = x[:, None] if one_d else x
        B, Hh, Ww, C = x4.shape
        w4 = np.asarray(w16, dtype=H).reshape(kh, kw, C, -1)
        Ho, Wo = (Hh - kh) // sh + 1, (Ww - kw) // sw + 1
        cols = np.empty((B, Ho, Wo, kh, kw, C), dtype=F)
        for i in range(kh):
            for j in range(kw):
                cols[:, :, :, i, j, :] = x4[:, i:i + (Ho - 1) * sh + 1:sh, j:j + (Wo - 1) * sw + 1:sw, :].astype(F)
        acc = cols.reshape(B * Ho * Wo, kh * kw * C) @ w4.astype(F).reshape(kh * kw * C, -1)
        z = acc.astype(H).reshape(B, Ho, Wo, -1)
        if one_d:
            z = z[:, 0]
    z = (z + np.asarray(b16, dtype=H)).astype(H)                   # BiasAdd, one fp16 op
    if act == 'relu':
        y = np.where(np.isnan(z), z, np.maximum(z, H(0))).astype(H)
    elif act == 'sigmoid':
        y = sigmoid16(z)
    else:
        y = z
    return z, y


def activation_backward(dy16: np.ndarray, y16: np.ndarray, act: str) -> np.ndarray:
    """[TF2.6] ReluGrad: dy where y > 0; SigmoidGrad: dy * y * (1 - y), each product one fp16 op."""
    dy, y = np.asarray(dy16, dtype=H), np.asarray(y16, dtype=H)
    if act == 'relu':
        return np.where(y > H(0), dy, H(0)).astype(H)
    if act == 'sigmoid':
        return ((dy * y).astype(H) * (ONE - y).astype(H)).astype(H)
    return dy


def maxpool_forward(x16: np.ndarray, pool=(2, 2), strides=(2, 2)):
    """(B, H, W, C) -> (y, arg): the FIRST maximum of a window in row-major scan order wins; NaN beats a number."""
    x = np.asarray(x16, dtype=H)
    B, Hh, Ww, C = x.shape
    ph, pw = pool
    sh, sw = strides
    Ho, Wo = (Hh - ph) // sh + 1, (Ww - pw) // sw + 1
    y = np.empty((B, Ho, Wo, C), dtype=H)
    arg = np.zeros((B, Ho, Wo, C), dtype=np.int64)
    for b in range(B):
        for ho in range(Ho):
            for wo in range(Wo):
                for c in range(C):
                    best, at = H(-np.inf), 0
                    for i in range(ph):
                        for j in range(pw):
                            v = x[b, ho * sh + i, wo * sw + j, c]
                            if v > best or (np.isnan(v) and not np.isnan(best)):
                                best, at = v, i * pw + j
                    y[b, ho, wo, c] = best
                    arg[b, ho, wo, c] = at
    return y, arg


def maxpool_backward(dy16: np.ndarray, arg: np.ndarray, in_shape, pool=(2, 2), strides=(2, 2)) -> np.ndarray:
    """The gradient of a window goes to its first maximum; overlapping windows add in window scan order (fp32, rounded once)."""
    dy = np.asarray(dy16, dtype=H)
    B, Hh, Ww, C = in_shape
    ph, pw = pool
    sh, sw = strides
    dx = np.zeros(in_shape, dtype=F)
    for i in range(ph):
        for j in range(pw):
            for b in range(B):
                for ho in range(dy.shape[1]):
                    for wo in range(dy.shape[2]):
                        sel = arg[b, ho, wo] == i * pw + j
                        dx[b, ho * sh + i, wo * sw + j] += np.where(sel, dy[b, ho, wo].astype(F), F(0))
    return dx.astype(H)
