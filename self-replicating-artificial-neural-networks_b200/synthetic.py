"""Synthetic stand-ins for the assets that are not available offline.

The figshare encodings, the trained ribosomal autoencoder and the MNIST download cannot be fetched
here, so benchmarks and tests use arrays of the same shapes and dtypes drawn from fixed seeds:
genotypes/encodings seed 1, images seed 2, labels seed 3, autoencoder weights seed 4.
"""
from __future__ import annotations

from typing import Dict

import numpy as np

from .vocabulary import VOCABULARY_SIZE


def synthetic_dataset(n_train=60000, n_test=10000, image_dims=(28, 28), genotype_size=100, n_classes=10):
    """Arrays shaped like ``_build_dataset`` output (reference ``experiment_worker.py:197-220``)."""
    n = n_train + n_test
    enc = np.random.Generator(np.random.PCG64(1)).integers(0, 2, (n, genotype_size)).astype(np.uint8)
    x = np.random.Generator(np.random.PCG64(2)).random((n,) + tuple(image_dims), dtype=np.float32)
    y = np.random.Generator(np.random.PCG64(3)).integers(0, n_classes, n).astype(np.uint8)
    return {'x_train': x[:n_train], 'y_train': y[:n_train], 'g_train': enc[:n_train],
            'x_test': x[n_train:], 'y_test': y[n_train:], 'g_test': enc[n_train:], 'encodings': enc}


def synthetic_decoder_weights(genotype_size=100, max_tokens=350, vocab=VOCABULARY_SIZE, seed=4) -> Dict[str, np.ndarray]:
    """Random-init ``generative_net`` weights in Keras layouts (reference ``ribosomal_autoencoder/model.py:35-48``).

    Glorot-uniform kernels; the BatchNormalization statistics are perturbed by N(0, 0.1) so that the
    second, per-token normalisation actually changes the argmax.
    """
    rng = np.random.Generator(np.random.PCG64(seed))
    L = genotype_size - 4

    def glorot(shape, fan_in, fan_out):
        lim = np.sqrt(6.0 / (fan_in + fan_out))
        return rng.uniform(-lim, lim, shape).astype(np.float32)

    def bn(c):
        return {'gamma': (1 + 0.1 * rng.standard_normal(c)).astype(np.float32),
                'beta': (0.1 * rng.standard_normal(c)).astype(np.float32),
                'mean': (0.1 * rng.standard_normal(c)).astype(np.float32),
                'var': np.abs(1 + 0.1 * rng.standard_normal(c)).astype(np.float32)}

    w = {'conv_kernel': glorot((5, 2, 32), 5 * 2, 5 * 32), 'conv_bias': np.zeros(32, np.float32),
         'dense_kernel': glorot((L * 32, max_tokens * vocab), L * 32, max_tokens * vocab),
         'dense_bias': np.zeros(max_tokens * vocab, np.float32)}
    for name, c in (('bn1', 32), ('bn2', vocab)):
        for k, v in bn(c).items():
            w[f'{name}_{k}'] = v
    return w
