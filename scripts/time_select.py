import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, libpga_b200 as pg
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
g = pg.PGA(seed=1)
p = g.create_population(n, 8)
g.set_objective_function("sum_obj")
rng = np.random.default_rng(0)
p.set_scores(rng.integers(480, 545, size=n).astype(np.float32))
k = n // 100
for worst in (False, True):
    g.select_top(p, k, worst=worst)
t0 = time.perf_counter()
for i in range(50):
    g.select_top(p, k, worst=bool(i & 1))
print("n", n, "k", k, "pgax_select_top wall-clock per call: %.1f us" % ((time.perf_counter() - t0) / 50 * 1e6))
