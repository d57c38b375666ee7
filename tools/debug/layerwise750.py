"""Layer-by-layer comparison of one 750-row training step (no optimiser) of chosen bench-population nets against the oracle:
python tools/debug/layerwise750.py 4 [1 ...]   (indices into sample_population(100, seed=0)[::12])"""
import sys, pathlib
ROOT = pathlib.Path(__file__).resolve().parent.parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / 'tests'))
import numpy as np, torch
from oracle import serann_oracle as O
from serann_b200 import frontend, generator
import parity_util as U

args = sys.argv[1:]
pop = generator.sample_population(100, seed=0, distinct=True)[::12]
show = None
if args and args[0] in ('all', 'allA'):          # the population of tests/test_gpu_bench_shape.py, report only the listed nets
    which = list(range(len(pop)))
    show = [int(a) for a in args[1:]]
else:
    which = [int(a) for a in args] or [4]
irs = [pop[i][1] for i in which]
if args and args[0] == 'allA':
    irs.append(frontend.compile_source(U.NET_A))
B = 750
data = U.small_dataset(n_train=3 * B + 300, n_test=64, seed=21)
eng, inds = U.make_engine_and_oracle(irs, data, max_batch=B)
rows = np.arange(B)
eng.debug_step(rows, apply_adam=False)
xi, gi = O.prepare_inputs(data['x_train'][rows], data['g_train'][rows])
labels = torch.tensor(data['y_train'][rows].astype(np.int64))
bits = torch.tensor(data['g_train'][rows].astype(np.float32))
for i, ind in enumerate(inds):
    if show is not None and i not in show:
        continue
    tape = {}
    O.train_step(ind, xi, gi, labels, bits, 1, tape=tape, apply_adam=False)
    for tid, t in tape.items():
        l = irs[i].layers[tid - 2]
        if l.kind == frontend.KIND_RESHAPE:
            continue
        e = U.relerr(eng.debug_tensor(i, tid, False, B), t.detach().numpy())
        line = f'net {i} tensor {tid} kind {l.kind} shape {l.out_shape} act {e:.3e}'
        if t.grad is not None and l.kind != frontend.KIND_CONCAT:
            ref_g = t.grad.numpy(); y = t.detach().numpy()
            if l.kind in (frontend.KIND_DENSE, frontend.KIND_CONV) and l.act == frontend.ACT_RELU:
                ref_g = ref_g * (y > 0)
            elif l.kind in (frontend.KIND_DENSE, frontend.KIND_CONV) and l.act == frontend.ACT_SIGMOID:
                h = lambda a: a.astype(np.float16).astype(np.float32)
                ref_g = h(h(ref_g * y) * h(1 - y))
            got = eng.debug_tensor(i, tid, True, B).astype(np.float32)
            eg = U.relerr(got, ref_g)
            # per-sample error profile: which rows are off
            d = np.abs(got - ref_g).reshape(B, -1).max(axis=1) / (np.abs(ref_g).max() + 1e-20)
            badrows = np.where(d > 3e-2)[0]
            line += f' | grad {eg:.3e} bad rows {len(badrows)} {badrows[:12].tolist()}'
            dd = np.abs(got - ref_g).reshape(B, -1)
            r, c = np.unravel_index(np.argmax(dd), dd.shape)
            line += f' worst at row {r} col {c}: got {got.reshape(B, -1)[r, c]:.3e} ref {ref_g.reshape(B, -1)[r, c]:.3e} y {t.detach().numpy().reshape(B, -1)[r, c]:.4e}'
        print(line)
    for li, pi, p in ind.trainable():
        got_g = eng.debug_grad(i, li, pi)
        ref_w = p.grad.numpy().reshape(got_g.shape)
        print(f'net {i} layer {li} kind {irs[i].layers[li].kind} param {pi} wgrad relerr {U.relerr(got_g, ref_w):.3e}')
eng.close()
