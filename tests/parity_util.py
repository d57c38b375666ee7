"""Shared helpers for the GPU parity tests: small populations, seeded data, layer-by-layer comparison."""
import os
import pathlib

import numpy as np
import torch

from oracle import serann_oracle as O
from serann_b200 import frontend, generator

NET_A = """X_layer = Conv2D(filters=8, kernel_size=3, strides=1)(X_layer)
X_layer = MaxPool2D(pool_size=2)(X_layer)
X_layer = BatchNormalization()(X_layer)
X_layer = Conv2D(filters=16, kernel_size=5, strides=2)(X_layer)
X_layer = Dense(units=24, activation='sigmoid')(X_layer)

g_layer = Conv1D(filters=4, kernel_size=7, strides=2)(g_layer)
g_layer = BatchNormalization()(g_layer)
g_layer = Dense(units=19, activation='relu')(g_layer)

con = concatenate([Reshape((1, -1))(X_layer), Reshape((1, -1))(g_layer)])

con = Dense(units=71, activation='relu')(con)
con = BatchNormalization()(con)

loss_balance = 0.7"""

NET_B = """con = concatenate([Reshape((1, -1))(X_layer), Reshape((1, -1))(g_layer)])

con = Dense(units=33, activation='relu')(con)

loss_balance = 0.4123"""

# K = 1 first Dense, overlapping pooling windows, a tensor consumed twice (gradient accumulation)
NET_C = """X_layer = Dense(units=9, activation='relu')(X_layer)
X_layer = MaxPool2D(pool_size=3, strides=2)(X_layer)
g_layer = Dense(units=12, activation='sigmoid')(g_layer)
con = concatenate([Reshape((1, -1))(X_layer), Reshape((1, -1))(g_layer), Reshape((1, -1))(X_layer)])
con = Dense(units=40, activation='sigmoid')(con)

loss_balance = 0.25"""

# wide filters (64-column tiles), kernel tuple, stride 2 both branches, conv activation
NET_D = """X_layer = Conv2D(filters=64, kernel_size=(3, 5), strides=2, activation='relu')(X_layer)
X_layer = Conv2D(filters=40, kernel_size=3, strides=1)(X_layer)
g_layer = Conv1D(filters=64, kernel_size=9, strides=1)(g_layer)
g_layer = Conv1D(filters=20, kernel_size=3, strides=2)(g_layer)
con = concatenate([Reshape((1, -1))(X_layer), Reshape((1, -1))(g_layer)])
con = Dense(units=130, activation='relu')(con)
con = Dense(units=32, activation='sigmoid')(con)

loss_balance = 0.9"""

# tensor-core shapes: 64-wide and 256-wide N tiles, stride-2 convolution from 64 channels, Conv1D from 48 channels
NET_E = """X_layer = Conv2D(filters=16, kernel_size=3, strides=1)(X_layer)
X_layer = MaxPool2D(pool_size=2)(X_layer)
X_layer = Conv2D(filters=64, kernel_size=3, strides=1)(X_layer)
X_layer = Conv2D(filters=8, kernel_size=3, strides=2)(X_layer)
g_layer = Dense(units=16, activation='relu')(g_layer)
g_layer = Dense(units=48, activation='sigmoid')(g_layer)
g_layer = Conv1D(filters=24, kernel_size=5, strides=1)(g_layer)
con = concatenate([Reshape((1, -1))(X_layer), Reshape((1, -1))(g_layer)])
con = Dense(units=200, activation='relu')(con)
con = Dense(units=256, activation='sigmoid')(con)

loss_balance = 0.5"""

HANDCRAFTED = [NET_A, NET_B, NET_C, NET_D, NET_E]


def small_population(n_generated=3, seed=11, max_flops=2.0e6):
    irs = [frontend.compile_source(s) for s in HANDCRAFTED]
    srcs = list(HANDCRAFTED)
    for s, ir in generator.sample_population(n_generated, seed=seed, max_flops_per_sample=max_flops):
        srcs.append(s)
        irs.append(ir)
    return srcs, irs


def small_dataset(n_train=160, n_test=64, seed=0, G=100):
    r = np.random.default_rng(seed)
    return {'x_train': r.random((n_train, 28, 28), dtype=np.float32), 'y_train': r.integers(0, 10, n_train).astype(np.uint8),
            'g_train': r.integers(0, 2, (n_train, G)).astype(np.uint8),
            'x_test': r.random((n_test, 28, 28), dtype=np.float32), 'y_test': r.integers(0, 10, n_test).astype(np.uint8),
            'g_test': r.integers(0, 2, (n_test, G)).astype(np.uint8)}


def relerr(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    both_nan = np.isnan(a) & np.isnan(b)
    d = np.where(both_nan, 0.0, np.abs(a - b))
    if np.isnan(d).any():
        return float('inf')
    return float(d.max() / (np.abs(np.where(np.isnan(b), 0, b)).max() + 1e-12)) if d.size else 0.0


def weights_close(x, y):
    """fp16 Adam divides by eps + sqrt(v) with v often flushed to zero, so a one-ulp gradient difference can move one
    element by a whole step (<= 1e-3 per step).  Judge the tensor by how many elements are off, not by the worst one:
    returns (fraction of elements further than 1e-2 of the tensor's scale, mean error relative to the scale)."""
    x = np.asarray(x, np.float32)
    y = np.asarray(y, np.float32)
    both_nan = np.isnan(x) & np.isnan(y)
    d = np.where(both_nan, 0.0, np.abs(x - y))
    if np.isnan(d).any():
        return 1.0, float('inf')
    scale = float(np.abs(np.where(np.isnan(y), 0, y)).max()) + 1e-6
    return float((d > 1e-2 * scale).mean()), float(d.mean() / scale)


def report(lines):
    path = os.environ.get('SRN_REPORT')
    if path:
        pathlib.Path(path).parent.mkdir(parents=True, exist_ok=True)
        with open(path, 'a') as f:
            f.write('\n'.join(lines) + '\n')


def make_engine_and_oracle(irs, data, max_batch=64, weight_seed=5):
    from serann_b200.engine import Engine
    eng = Engine(max_batch=max_batch)
    eng.set_dataset(data['x_train'], data['y_train'], data['g_train'], data['x_test'], data['y_test'], data['g_test'])
    eng.build(irs, seed=1)
    rng = np.random.default_rng(weight_seed)
    inds = []
    for i, ir in enumerate(irs):
        w = O.init_weights(ir, rng)
        # non-trivial BN parameters / statistics so that every term of the BN formulas is exercised
        for l, ws in zip(ir.layers, w):
            if l.kind == frontend.KIND_BN:
                c = len(ws[0])
                ws[0] = (1 + 0.2 * rng.standard_normal(c)).astype(np.float16)
                ws[1] = (0.1 * rng.standard_normal(c)).astype(np.float16)
                ws[2] = (0.1 * rng.standard_normal(c)).astype(np.float16)
                ws[3] = np.abs(1 + 0.2 * rng.standard_normal(c)).astype(np.float16)
            elif l.kind in (frontend.KIND_DENSE, frontend.KIND_CONV):
                ws[1] = (0.05 * rng.standard_normal(len(ws[1]))).astype(np.float16)
        eng.set_weights(i, w)
        inds.append(O.Individual(ir, w))
    return eng, inds

# pooling / BatchNormalization applied directly to the one-channel model inputs, overlapping and uneven pooling windows
NET_F = """X_layer = MaxPool2D(pool_size=2)(X_layer)
X_layer = Conv2D(filters=8, kernel_size=3, strides=1)(X_layer)
X_layer = MaxPool2D(pool_size=5)(X_layer)
g_layer = BatchNormalization()(g_layer)
g_layer = Conv1D(filters=8, kernel_size=5, strides=2)(g_layer)
con = concatenate([Reshape((1, -1))(X_layer), Reshape((1, -1))(g_layer)])
con = Dense(units=50, activation='relu')(con)

loss_balance = 0.35"""
HANDCRAFTED.append(NET_F)

# genotype first in the concatenation (the image branch lands at an unaligned offset and is compacted: 12 channels),
# 5x5 and 7x7 first layers (reductions of 32 and 56 over the shared im2col matrix), stride-2 convolution whose input
# needs a gradient (residue classes with 3, 2 and 2, 1 taps), a long Dense reduction (split along k), per-pixel Dense
NET_G = """X_layer = Conv2D(filters=12, kernel_size=5, strides=1)(X_layer)
X_layer = Conv2D(filters=12, kernel_size=(5, 4), strides=2, activation='relu')(X_layer)
g_layer = Dense(units=6, activation='relu')(g_layer)
con = concatenate([Reshape((1, -1))(g_layer), Reshape((1, -1))(X_layer)])
con = Dense(units=90, activation='relu')(con)
con = Dense(units=20)(con)

loss_balance = 0.6"""
HANDCRAFTED.append(NET_G)

NET_H = """X_layer = Conv2D(filters=3, kernel_size=7, strides=1)(X_layer)
X_layer = BatchNormalization()(X_layer)
X_layer = Conv2D(filters=72, kernel_size=3, strides=2)(X_layer)
X_layer = Conv2D(filters=16, kernel_size=3, strides=1, activation='relu')(X_layer)
g_layer = Conv1D(filters=16, kernel_size=3, strides=1)(g_layer)
g_layer = Dense(units=10)(g_layer)
con = concatenate([Reshape((1, -1))(X_layer), Reshape((1, -1))(g_layer)])
con = Dense(units=64, activation='sigmoid')(con)

loss_balance = 0.15"""
HANDCRAFTED.append(NET_H)
