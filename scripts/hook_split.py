"""Where does the time of the test3 hook workload go?  Times 5 generations at 2^20 x 100
with each combination of the demo hooks."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import libpga_b200 as pg

with open("tests/golden/tsp100.txt") as f:
    tok = f.read().split()
m = np.array(tok[1:10001], dtype=np.float32).reshape(100, 100)
for obj, xov in (("sum_obj", None), ("tsp_obj", None), ("sum_obj", "tsp_crossover"), ("tsp_obj", "tsp_crossover")):
    g = pg.PGA(seed=1)
    p = g.create_population(1 << 20, 100)
    g.tsp_load(m)
    g.set_objective_function(obj)
    g.set_crossover_function(xov)
    g.run(3)
    g.run(5)
    print(f"objective={obj:8s} crossover={xov}: {g.last_run_ms / 5 * 1000:9.1f} us/generation at 2^20")
    g.deinit()
