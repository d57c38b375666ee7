"""Host cost of the launch plans: the first fit of a freshly built population against the second (plans cached)."""
import sys, pathlib, time
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent.parent))
from serann_b200 import generator
from serann_b200.engine import Engine
from serann_b200.synthetic import synthetic_dataset

pop = generator.sample_population(100, seed=0, distinct=True)
data = synthetic_dataset()
eng = Engine(max_batch=750)
eng.set_dataset(data['x_train'], data['y_train'], data['g_train'], data['x_test'], data['y_test'], data['g_test'])
for rep in range(3):
    t0 = time.perf_counter()
    eng.build([p[1] for p in pop], seed=7)
    t1 = time.perf_counter()
    r1 = eng.fit(1, 750, 3000, shuffle_seed=1, max_steps=5)
    t2 = time.perf_counter()
    r2 = eng.fit(1, 750, 3000, shuffle_seed=2, max_steps=5)
    t3 = time.perf_counter()
    print(f'build {1e3 * (t1 - t0):.1f} ms; first fit {1e3 * (t2 - t1):.1f} ms wall ({r1["device_ms"]:.1f} device); second fit {1e3 * (t3 - t2):.1f} ms wall ({r2["device_ms"]:.1f} device)')
