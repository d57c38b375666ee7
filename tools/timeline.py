"""Timeline of one training step of the bench population with the side streams on (overlapped schedule):
python tools/timeline.py [pop].  One row per depth level and direction: wall span, summed kernel time, slowest launch."""
import sys, pathlib
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import numpy as np
from serann_b200 import generator
from serann_b200.engine import Engine
from serann_b200.synthetic import synthetic_dataset

if __name__ == '__main__':
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 100
    pop = generator.sample_population(n, seed=0, distinct=True)
    data = synthetic_dataset()
    eng = Engine(max_batch=750)
    eng.set_dataset(data['x_train'], data['y_train'], data['g_train'], data['x_test'], data['y_test'], data['g_test'])
    eng.build([p[1] for p in pop], seed=7)
    eng.fit(1, 750, 3000, shuffle_seed=1, max_steps=3)
    eng.profile(2)
    res = eng.fit(1, 750, 3000, shuffle_seed=2, max_steps=1)
    ms, info, work = eng.profile_detail(750)
    start = eng.profile_start_ms[:len(ms)]
    eng.profile(False)
    cats = Engine.CATEGORIES
    print(f'step {res["device_ms"]:.3f} ms, {len(ms)} launches, summed kernel time {ms.sum():.3f} ms')
    levels = []          # (label, [indices])
    backward = False
    for i in range(len(ms)):
        c = cats[info[i, 0]]
        if c in ('gather', 'loss', 'adam') or info[i, 1] < 0:
            if c == 'loss':
                backward = True
            levels.append((c, [i]))
            continue
        label = f'{"bwd" if backward else "fwd"} depth {info[i, 7]}'
        if levels and levels[-1][0] == label:
            levels[-1][1].append(i)
        else:
            levels.append((label, [i]))
    print('level, launches, t_start, span_ms, sum_kernel_ms, slowest (category bucket ms GFLOP)')
    for label, idx in levels:
        idx = np.array(idx)
        t0 = start[idx].min(); t1 = (start[idx] + ms[idx]).max()
        k = idx[np.argmax(ms[idx])]
        print(f'{label:14s} {len(idx):4d} {t0:8.3f} {t1 - t0:7.3f} {ms[idx].sum():7.3f}   {cats[info[k, 0]]} {info[k, 1]} {ms[k]:.3f} {work[k, 0] / 1e9:.2f}')
