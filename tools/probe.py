"""Micro-benchmark of one architecture replicated N times (clonal shard): python tools/probe.py [N] [steps]
Prints per-category device time per training step.  Used with ncu to study one kernel at a time."""
import sys, pathlib
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import numpy as np
from serann_b200 import frontend
from serann_b200.engine import Engine
from serann_b200.synthetic import synthetic_dataset

SRC = {
    'dense': """X_layer = Conv2D(filters=32, kernel_size=3, strides=1)(X_layer)
con = concatenate([Reshape((1, -1))(X_layer), Reshape((1, -1))(g_layer)])
con = Dense(units=128, activation='relu')(con)
loss_balance = 0.5""",
    'conv': """X_layer = Conv2D(filters=32, kernel_size=3, strides=1)(X_layer)
X_layer = Conv2D(filters=32, kernel_size=3, strides=1)(X_layer)
X_layer = MaxPool2D(pool_size=2)(X_layer)
X_layer = Conv2D(filters=64, kernel_size=3, strides=1)(X_layer)
X_layer = MaxPool2D(pool_size=2)(X_layer)
con = concatenate([Reshape((1, -1))(X_layer), Reshape((1, -1))(g_layer)])
con = Dense(units=64, activation='relu')(con)
loss_balance = 0.5""",
    'wide': """X_layer = Conv2D(filters=128, kernel_size=3, strides=1)(X_layer)
X_layer = Conv2D(filters=64, kernel_size=3, strides=1)(X_layer)
X_layer = MaxPool2D(pool_size=4)(X_layer)
con = concatenate([Reshape((1, -1))(X_layer), Reshape((1, -1))(g_layer)])
con = Dense(units=64, activation='relu')(con)
loss_balance = 0.5""",
    'mid': """X_layer = Conv2D(filters=32, kernel_size=3, strides=1)(X_layer)
X_layer = Conv2D(filters=24, kernel_size=3, strides=1)(X_layer)
X_layer = MaxPool2D(pool_size=4)(X_layer)
con = concatenate([Reshape((1, -1))(X_layer), Reshape((1, -1))(g_layer)])
con = Dense(units=64, activation='relu')(con)
loss_balance = 0.5""",
}

if __name__ == '__main__':
    which = sys.argv[1] if len(sys.argv) > 1 else 'dense'
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 16
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 4
    ir = frontend.compile_source(SRC[which])
    data = synthetic_dataset(n_train=6000, n_test=100)
    eng = Engine(max_batch=750)
    eng.set_dataset(data['x_train'], data['y_train'], data['g_train'], data['x_test'], data['y_test'], data['g_test'])
    eng.build([ir] * n, seed=1)
    eng.fit(1, 750, 0, shuffle_seed=1, max_steps=3)
    eng.profile(True)
    res = eng.fit(1, 750, 0, shuffle_seed=2, max_steps=steps)
    prof = eng.profile_read()
    eng.profile(False)
    gflop = ir.train_flops_per_sample() * 750 * n / 1e9
    print(f'{which} x{n}: {res["device_ms"] / steps:.3f} ms/step, {gflop:.1f} GFLOP/step -> {gflop / (res["device_ms"] / steps):.1f} TFLOP/s')
    for k, (ms, c) in prof.items():
        if c:
            print(f'  {k:12s} {ms / steps:8.3f} ms/step  {c // steps:3d} launches')
    eng.profile(True)
    eng.fit(1, 750, 0, shuffle_seed=2, max_steps=1)
    ms, info, work = eng.profile_detail(750)
    eng.profile(False)
    for t, i, w in zip(ms, info, work):
        if 1 <= i[0] <= 3:
            print(f'    {Engine.CATEGORIES[i[0]]:11s} bucket {i[1]:2d} tc {i[3]} BNf {i[4]} BNc {i[5]} probs {i[6]}: {t:.4f} ms {w[0] / 1e9 / t:7.1f} TFLOP/s')
