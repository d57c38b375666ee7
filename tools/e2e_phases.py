"""Wall-clock phases of one population-100 generation through the worker API: python tools/e2e_phases.py"""
import sys, pathlib, time
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import numpy as np, pandas as pd
import bench
from serann_b200 import engine as E, frontend
from serann_b200.synthetic import synthetic_dataset
from serann_b200.worker import B200PopulationWorker

t = time.perf_counter
pop = bench.population(100, 'diverse')
data = synthetic_dataset()
ids = [f's{i}' for i in range(len(pop))]
info = pd.DataFrame({'genotype': [data['encodings'][i] for i in range(len(pop))], 'source_code': [p[0] for p in pop]}, index=ids)
for rep in range(2):
    t0 = t(); w = B200PopulationWorker(dict(bench.PARAMS), dataset=data, device=0, shuffle_seed=4); t1 = t()
    irs, _ = w._compile(info); t2 = t()
    eng = w._engine
    eng.build([irs[i] for i in ids], seed=1); t3 = t()
    res = eng.fit(5, 750, 3000, shuffle_seed=2); t4 = t()
    eng.evaluate(1, 0, 10000, 32); t5 = t()
    parents = np.stack([np.asarray(data['encodings'][i]) for i in range(len(pop))]).astype(np.uint8)
    rows = np.arange(len(pop) * 10)
    eng.replicate(parents, 10, image_index=rows // 1); t6 = t()
    print(f'rep {rep}: worker init (H2D data set) {t1 - t0:.3f}  compile sources {t2 - t1:.3f}  build {t3 - t2:.3f}  fit wall {t4 - t3:.3f} '
          f'(device {res["device_ms"] / 1000:.3f})  test evaluate {t5 - t4:.3f}  replicate {t6 - t5:.3f}  total {t6 - t0:.3f}')
    w.close()
