"""Per-launch-bucket device time of one training step of the bench population (serialised launches):
python tools/buckets.py [pop] [out.csv].  One row per (category, bucket) with FLOPs, activation bytes and ms."""
import sys, pathlib, collections
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import numpy as np
from serann_b200 import generator
from serann_b200.engine import Engine
PROFILE_CATEGORIES = Engine.CATEGORIES
from serann_b200.synthetic import synthetic_dataset

KIND = {0: 'dense', 1: 'conv', 2: 'pool', 3: 'bn', 4: 'concat', 5: 'reshape', -1: '-'}
BN = [32, 64, 128, 256]

if __name__ == '__main__':
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 100
    out = sys.argv[2] if len(sys.argv) > 2 else None
    pop = generator.sample_population(n, seed=0, distinct=True)
    data = synthetic_dataset()
    eng = Engine(max_batch=750)
    eng.set_dataset(data['x_train'], data['y_train'], data['g_train'], data['x_test'], data['y_test'], data['g_test'])
    eng.build([p[1] for p in pop], seed=7)
    eng.fit(1, 750, 3000, shuffle_seed=1, max_steps=3)
    steps = 2
    eng.profile(True)
    eng.fit(1, 750, 3000, shuffle_seed=2, max_steps=steps)
    ms, info, work = eng.profile_detail(750)
    eng.profile(False)
    agg = collections.OrderedDict()
    for t, i, w in zip(ms, info, work):
        key = (int(i[0]), int(i[1]))
        a = agg.setdefault(key, [0.0, 0, i, w])
        a[0] += t / steps
        a[1] += 1
    rows = []
    for (cat, grp), (t, c, i, w) in agg.items():
        name = PROFILE_CATEGORIES[cat]
        bn = BN[i[5]] if name == 'gemm_dgrad' and i[3] > 0 else (BN[i[4]] if i[3] > 0 and i[4] >= 0 else -1)
        rows.append((name, grp, KIND.get(int(i[2]), '?'), int(i[3]), bn, int(i[6]), int(i[7]), w[0] / 1e9, (w[1] + w[2]) / 1e6, t,
                     w[0] / 1e9 / t if t > 0 else 0.0, (w[1] + w[2]) / 1e6 / t if t > 0 else 0.0))
    rows.sort(key=lambda r: -r[9])
    hdr = 'category,bucket,kind,tc,BN,problems,depth,GFLOP,act_MB,ms,TFLOPs,GBs'
    lines = [hdr] + [f'{r[0]},{r[1]},{r[2]},{r[3]},{r[4]},{r[5]},{r[6]},{r[7]:.2f},{r[8]:.1f},{r[9]:.4f},{r[10]:.1f},{r[11]:.0f}' for r in rows]
    total = sum(r[9] for r in rows)
    print(f'total serialised kernel time {total:.3f} ms/step over {len(rows)} (category, bucket) rows')
    print('\n'.join(lines[:70]))
    for g in sys.argv[3:]:
        sel = [(float(t) / steps, PROFILE_CATEGORIES[i[0]]) for t, i in zip(ms, info) if int(i[1]) == int(g)]
        print('bucket', g, 'launches in issue order (ms, summed over', steps, 'steps / steps):', [f'{c}:{t:.4f}' for t, c in sel])
        print('bucket', g, '{H, W, C, Cs, Ho, Wo, F, Fs, kh, kw, sh, sw, act, din_acc, individual, has din, coff, n_el, Cj, in_sel}')
        np.set_printoptions(linewidth=200, threshold=100000)
        print(eng.group_probs(int(g)))
    if out:
        pathlib.Path(out).write_text('\n'.join(lines) + '\n')
