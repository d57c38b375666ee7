"""Writes tests/golden/oracle_small.npz: the oracle's outputs on a small seeded case.

The reference cannot be imported here (TensorFlow 2.6 is not installable), so these are NOT reference
outputs; they pin the oracle restatement against accidental drift and give the GPU tests a file-based
target that travels to the GPU box.  Run from the repo root:  python tests/golden/make_golden.py
"""
import pathlib
import sys

import numpy as np
import torch

ROOT = pathlib.Path(__file__).resolve().parent.parent.parent
for p in (str(ROOT), str(ROOT / 'tests')):
    if p not in sys.path:
        sys.path.insert(0, p)


def build_case():
    import parity_util as U
    from oracle import serann_oracle as O
    from serann_b200 import frontend
    torch.set_num_threads(1)
    irs = [frontend.compile_source(s) for s in (U.NET_A, U.NET_C)]
    data = U.small_dataset(n_train=96, n_test=16, seed=42)
    rng = np.random.default_rng(7)
    weights = [O.init_weights(ir, rng) for ir in irs]
    inds = [O.Individual(ir, w) for ir, w in zip(irs, weights)]
    rows = np.arange(32)
    xi, gi = O.prepare_inputs(data['x_train'][rows], data['g_train'][rows])
    labels = torch.tensor(data['y_train'][rows].astype(np.int64))
    bits = torch.tensor(data['g_train'][rows].astype(np.float32))
    logits0, losses = [], []
    for ind in inds:
        per = []
        for t in range(3):
            r = O.train_step(ind, xi, gi, labels, bits, t + 1)
            if t == 0:
                logits0.append(r['logits'].numpy().astype(np.float16))
            per.append([r['ce'], r['mse']])
        losses.append(per)
    off = O.replicate(inds, data['g_test'][:2], data['x_test'], 4)
    return {'logits_step0': np.stack(logits0), 'losses': np.asarray(losses, dtype=np.float32),
            'offspring': np.stack(off), 'weight_seed': 7, 'data_seed': 42}


if __name__ == '__main__':
    out = pathlib.Path(__file__).resolve().parent / 'oracle_small.npz'
    np.savez_compressed(out, **build_case())
    print('wrote', out)
