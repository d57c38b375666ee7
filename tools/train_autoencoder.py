"""Ribosomal autoencoder: PyTorch restatement of the reference's model and training loop.

Not part of the per-generation path and not part of the product library: the evolution loop only ever *decodes*
(libserann_b200's tcgen05 decode kernel).  The figshare autoencoder is not available offline, and a random-init
decoder turns offspring genotypes into invalid programs; this tool trains one on generator-sampled sources so that the
synthetic loop can run on its own decodes, and exports it in the Keras weight layouts `RibosomalAutoencoder` loads.

Restates (reference file:line):
  GeneticAutoencoder.inference_net / generative_net   ribosomal_autoencoder/model.py:17-48
  ConcreteGAE (Gumbel relaxation, prior, nelbo)        ribosomal_autoencoder/model.py:62-104
  training loop, temperature / learning-rate / KL      ribosomal_autoencoder/training.py:51-101
  schedules (log-spaced, indexed by the batch number)
  encode = argmax of the inference net's logits         model.py:101-104; helpers/tokens_to_genotypes.py:51-53
Layers are applied channels-last exactly as Keras does (Flatten order, kernel layouts), so exported weights mean the
same thing to the CUDA decoder.  Keras BatchNormalization defaults: momentum 0.99, epsilon 1e-3.

    python tools/train_autoencoder.py --steps 20000 --sources 50000 --out ae_trained.npz [--device cuda]
    python -m serann_b200.run_experiment -p params.json --assets ae_trained.npz ...      (with the data arrays added)
"""
from __future__ import annotations

import argparse
import math
import pathlib
import sys

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

BN_MOMENTUM, BN_EPS = 0.99, 1e-3


class KerasBN(nn.Module):
    """keras.layers.BatchNormalization on the last axis (moving = moving * 0.99 + batch * 0.01, biased batch variance)."""

    def __init__(self, c):
        super().__init__()
        self.gamma, self.beta = nn.Parameter(torch.ones(c)), nn.Parameter(torch.zeros(c))
        self.register_buffer('mean', torch.zeros(c))
        self.register_buffer('var', torch.ones(c))

    def forward(self, x, training: bool):
        if training:
            dims = tuple(range(x.dim() - 1))
            m, v = x.mean(dims), x.var(dims, unbiased=False)
            with torch.no_grad():
                self.mean.mul_(BN_MOMENTUM).add_(m * (1 - BN_MOMENTUM))
                self.var.mul_(BN_MOMENTUM).add_(v * (1 - BN_MOMENTUM))
        else:
            m, v = self.mean, self.var
        inv = torch.rsqrt(v + BN_EPS) * self.gamma
        return x * inv + (self.beta - m * inv)


def glorot_(w: torch.Tensor, fan_in: int, fan_out: int):
    lim = math.sqrt(6.0 / (fan_in + fan_out))
    with torch.no_grad():
        w.uniform_(-lim, lim)
    return w


class GeneticAutoencoder(nn.Module):
    """model.py:10-60.  All tensors channels-last; kernels in Keras layouts (Conv2D HWIO, Conv1D (k, in, out), Dense (in, out))."""

    def __init__(self, genotype_length=100, max_tokens=350, vocab=40, embedding=50, alphabet=2):
        super().__init__()
        self.G, self.T, self.V, self.E, self.A = genotype_length, max_tokens, vocab, embedding, alphabet
        P = lambda *s: nn.Parameter(torch.empty(*s))
        # inference net
        self.emb = nn.Parameter(torch.empty(vocab, embedding).uniform_(-0.05, 0.05))       # Embedding default init
        self.e_bn0 = KerasBN(1)
        self.e_k1 = glorot_(P(5, 5, 1, 32), 25, 25 * 32); self.e_b1 = nn.Parameter(torch.zeros(32)); self.e_bn1 = KerasBN(32)
        self.e_k2 = glorot_(P(3, 3, 32, 16), 9 * 32, 9 * 16); self.e_b2 = nn.Parameter(torch.zeros(16)); self.e_bn2 = KerasBN(16)
        self.e_k3 = glorot_(P(3, 3, 16, 16), 9 * 16, 9 * 16); self.e_b3 = nn.Parameter(torch.zeros(16)); self.e_bn3 = KerasBN(16)
        h, w = max_tokens - 4 - 2 - 2, embedding - 4 - 2 - 2
        self.e_flat = h * w * 16
        self.e_kd = glorot_(P(self.e_flat, genotype_length * alphabet), self.e_flat, genotype_length * alphabet)
        self.e_bd = nn.Parameter(torch.zeros(genotype_length * alphabet))
        # generative net
        L = genotype_length - 4
        self.d_k1 = glorot_(P(5, alphabet, 32), 5 * alphabet, 5 * 32); self.d_b1 = nn.Parameter(torch.zeros(32)); self.d_bn1 = KerasBN(32)
        self.d_kd = glorot_(P(L * 32, max_tokens * vocab), L * 32, max_tokens * vocab)
        self.d_bd = nn.Parameter(torch.zeros(max_tokens * vocab)); self.d_bn2 = KerasBN(vocab)

    @staticmethod
    def _conv2d(x, k, b):                     # x (B, H, W, C), k HWIO
        y = F.conv2d(x.permute(0, 3, 1, 2), k.permute(3, 2, 0, 1), b)
        return y.permute(0, 2, 3, 1)

    def inference_net(self, tokens: torch.Tensor, training: bool):
        x = self.emb[tokens.long()].unsqueeze(-1)                     # (B, T, E, 1)
        x = self.e_bn0(x, training)
        x = self.e_bn1(self._conv2d(x, self.e_k1, self.e_b1), training)
        x = self.e_bn2(self._conv2d(x, self.e_k2, self.e_b2), training)
        x = self.e_bn3(self._conv2d(x, self.e_k3, self.e_b3), training)
        z = x.reshape(x.shape[0], -1) @ self.e_kd + self.e_bd
        return z.reshape(-1, self.G, self.A)

    def generative_net(self, z: torch.Tensor, training: bool):
        """z (B, G, alphabet) -> log-probabilities (B, T, V)."""
        y = F.conv1d(z.permute(0, 2, 1), self.d_k1.permute(2, 1, 0), self.d_b1).permute(0, 2, 1)      # (B, L, 32)
        y = self.d_bn1(y, training)
        y = y.reshape(y.shape[0], -1) @ self.d_kd + self.d_bd                                       # Flatten is L-major
        y = self.d_bn2(y.reshape(-1, self.T, self.V), training)
        return F.log_softmax(y, dim=-1)

    def encode(self, tokens):                 # ConcreteGAE.encode
        with torch.no_grad():
            return self.inference_net(tokens, False).argmax(-1)

    def decode(self, genotype):               # GeneticAutoencoder.decode
        with torch.no_grad():
            return self.generative_net(F.one_hot(genotype.long(), self.A).float(), False).argmax(-1)


def gumbel_log_prob(x, loc, scale):
    z = (x - loc) / scale
    return -(z + torch.exp(-z)) - math.log(scale)


def concrete_loss(model: GeneticAutoencoder, tokens, temperature: float, kld_weight: float, prior_temperature=0.01, training=True):
    """ConcreteGAE.compute_loss (model.py:88-100): z ~ Gumbel(logits / t, 1 / t), softmax relaxation, categorical
    likelihood of the tokens, empirical KL against the Gumbel prior."""
    logits = model.inference_net(tokens, training)
    loc, scale = logits / temperature, 1.0 / temperature
    u = torch.rand_like(loc).clamp_(1e-20, 1.0)
    sample = loc - scale * torch.log(-torch.log(u))
    logq = gumbel_log_prob(sample, loc, scale)
    p_loc = math.log(1.0 / model.A) / prior_temperature
    logp = gumbel_log_prob(sample, torch.full_like(sample, p_loc), 1.0 / prior_temperature)
    z = F.softmax(sample, dim=-1)
    logpx = model.generative_net(z, training).gather(-1, tokens.long().unsqueeze(-1)).squeeze(-1).sum(-1)
    kl = (logq - logp).sum(dim=(1, 2))
    return {'loss': -(logpx - kld_weight * kl).mean(), 'nll': -logpx.mean(), 'kld': kl.mean()}


def schedules(total_steps: int):
    """training.py:58-64, compressed from 5e6 / 1e6 / 1e7 batches to `total_steps` with the same shapes."""
    t = np.logspace(np.log10(0.3), -3, max(total_steps // 2, 2))
    lr = np.logspace(np.log10(3e-4), np.log10(2e-5), max(total_steps // 10, 2))
    kw = np.linspace(0, 0.2, max(total_steps, 2)) ** 2
    return t, lr, kw


def token_dataset(n: int, max_tokens: int, seed: int = 0) -> np.ndarray:
    from serann_b200.generator import sample_population
    from serann_b200.vocabulary import encode_tokens
    # sources of up to 350 tokens as the reference's data set holds them; a shorter max_tokens truncates (ribosomal_autoencoder.py:50)
    return np.stack([encode_tokens(s, max_tokens) for s, _ in sample_population(n, seed=seed, max_tokens=max(350, max_tokens))])


def export_weights(model: GeneticAutoencoder) -> dict:
    """Decoder in the layouts srn_decoder_create takes (Keras), encoder under enc_* for completeness."""
    g = lambda t: t.detach().cpu().numpy().astype(np.float32)
    w = {'conv_kernel': g(model.d_k1), 'conv_bias': g(model.d_b1), 'dense_kernel': g(model.d_kd), 'dense_bias': g(model.d_bd)}
    for name, bn in (('bn1', model.d_bn1), ('bn2', model.d_bn2)):
        w.update({f'{name}_gamma': g(bn.gamma), f'{name}_beta': g(bn.beta), f'{name}_mean': g(bn.mean), f'{name}_var': g(bn.var)})
    for k, v in model.state_dict().items():
        if k.startswith(('emb', 'e_')):
            w['enc_' + k.replace('.', '_')] = g(v)
    return w


def train(model, data: np.ndarray, steps: int, batch: int, device, log_every=100, seed=0):
    temps, lrs, kws = schedules(steps)
    opt = torch.optim.Adam(model.parameters(), lr=float(lrs[0]), eps=1e-7)
    x = torch.as_tensor(data, device=device)
    gen = torch.Generator(device='cpu').manual_seed(seed)
    hist = []
    for step in range(1, steps + 1):
        idx = torch.randint(0, len(x), (batch,), generator=gen).to(device)
        for grp in opt.param_groups:
            grp['lr'] = float(lrs[min(step, len(lrs) - 1)])
        m = concrete_loss(model, x[idx], float(temps[min(step, len(temps) - 1)]), float(kws[min(step, len(kws) - 1)]))
        opt.zero_grad(set_to_none=True)
        m['loss'].backward()
        opt.step()
        hist.append(float(m['nll'].detach()))
        if log_every and step % log_every == 0:
            with torch.no_grad():
                rec = (model.decode(model.encode(x[idx[:64]])) == x[idx[:64]]).float().mean()
            print(f'step {step}: loss {float(m["loss"]):.3f} nll {float(m["nll"]):.3f} kld {float(m["kld"]):.3f} token accuracy {float(rec):.3f}', flush=True)
    return hist


if __name__ == '__main__':
    ap = argparse.ArgumentParser()
    ap.add_argument('--steps', type=int, default=2000)
    ap.add_argument('--batch', type=int, default=512)
    ap.add_argument('--sources', type=int, default=20000)
    ap.add_argument('--max-tokens', type=int, default=350)
    ap.add_argument('--embedding', type=int, default=50)
    ap.add_argument('--device', default='cuda' if torch.cuda.is_available() else 'cpu')
    ap.add_argument('--out', default='ae_trained.npz')
    a = ap.parse_args()
    torch.manual_seed(534213)                  # ribosomal_autoencoder/config.py:12
    dev = torch.device(a.device)
    data = token_dataset(a.sources, a.max_tokens)
    model = GeneticAutoencoder(max_tokens=a.max_tokens, embedding=a.embedding).to(dev)
    train(model, data, a.steps, a.batch, dev)
    np.savez_compressed(a.out, **export_weights(model))
    print('wrote', a.out)
