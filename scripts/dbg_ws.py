import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sys, numpy as np, libpga_b200 as pg
from oracle import pyoracle as po
n, cap, gens = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
g = pg.PGA(seed=1); o = po.Oracle(1, threads=16)
p = g.create_population(n, 1024, encoding=pg.GENES_BITS); q = o.create_population(n, 1024, po.GENES_BITS)
g.set_builtin_objective(pg.OBJ_ONEMAX); o.set_builtin_objective(po.OBJ_ONEMAX)
if cap: p.capture_selection(True)
g.run(gens); o.run(gens)
print("n", n, "cap", cap, "gens", gens, "genes", np.array_equal(p.genes(), q.genes()), "scores", np.array_equal(p.scores(), q.scores()), flush=True)
