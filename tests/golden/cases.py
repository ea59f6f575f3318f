"""The golden trajectories: small cases of every BASELINE config, run on any engine
that mirrors pga.h (the CPU oracle or the CUDA library).  make_digests.py runs them on
the oracle and commits the digests; the tests re-run them on both engines."""
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

F32, BITS = 0, 1
OBJ_SUM, OBJ_RASTRIGIN, OBJ_SPHERE, OBJ_ONEMAX = 1, 2, 3, 4
RING, RANDOM = 0, 1

CASES = {
    # name: dict(seed, islands, size, length, enc, objective (hook name or builtin id), xover hook, run spec)
    "c1_test_cu_small": dict(seed=42, islands=1, size=4096, length=100, enc=F32, hook="sum_obj", gens=20),
    "c1_test_cu_full": dict(seed=42, islands=1, size=40000, length=100, enc=F32, hook="sum_obj", gens=100),
    "test2_knapsack": dict(seed=7, islands=1, size=100, length=6, enc=F32, hook="knapsack_obj", gens=5),
    "test3_tsp": dict(seed=3, islands=1, size=1000, length=100, enc=F32, hook="tsp_obj", xover="tsp_crossover", gens=60),
    "c2_onemax_small": dict(seed=42, islands=1, size=1 << 14, length=1024, enc=BITS, builtin=OBJ_ONEMAX, gens=30),
    "c3_rastrigin_small": dict(seed=42, islands=1, size=1 << 12, length=256, enc=F32, builtin=OBJ_RASTRIGIN, gens=30),
    "c5_islands_ring": dict(seed=17, islands=4, size=2048, length=512, enc=BITS, builtin=OBJ_ONEMAX, gens=13,
                            migrate_every=3, pct=5.0, topology=RING, delay=1),
    "c5_islands_random_f32": dict(seed=17, islands=3, size=1500, length=40, enc=F32, builtin=OBJ_RASTRIGIN, gens=11,
                                  migrate_every=2, pct=3.0, topology=RANDOM, delay=0),
}


def tsp_matrix():
    with open(os.path.join(HERE, "tsp100.txt")) as f:
        tok = f.read().split()
    n = int(tok[0])
    return np.array(tok[1:1 + n * n], dtype=np.float32).reshape(n, n)


def fnv1a(a):
    """FNV-1a 64 over the bytes of an array (vectorised per 8-byte lane would change the
    value; this is the plain byte-serial definition, in numpy for speed on small inputs)."""
    data = np.ascontiguousarray(a).view(np.uint8).ravel()
    h = 0xcbf29ce484222325
    # chunked python loop is too slow for megabytes: fold with the oracle when available
    try:
        from oracle import pyoracle
        return pyoracle.digest(np.ascontiguousarray(a))
    except Exception:
        for b in data.tolist():
            h = ((h ^ b) * 0x100000001b3) & 0xFFFFFFFFFFFFFFFF
        return h


def run_case(engine, spec, make_population, tsp_load=None):
    """engine: object with the pga.h-mirror methods (oracle.pyoracle.Oracle or libpga_b200.PGA).
    Returns {"genes": [...], "scores": [...], "best": [...]} digests per island."""
    pops = [make_population(spec["size"], spec["length"], spec["enc"]) for _ in range(spec["islands"])]
    if "builtin" in spec:
        engine.set_builtin_objective(spec["builtin"])
    if "hook" in spec:
        (engine.set_objective_hook if hasattr(engine, "set_objective_hook") else engine.set_objective_function)(spec["hook"])
    if "xover" in spec:
        (engine.set_crossover_hook if hasattr(engine, "set_crossover_hook") else engine.set_crossover_function)(spec["xover"])
    if spec.get("hook") == "tsp_obj":
        engine.tsp_load(tsp_matrix())
    if spec["islands"] == 1:
        engine.run(spec["gens"])
    else:
        engine.set_topology(spec["topology"])
        engine.set_migration_delay(spec["delay"])
        engine.run_islands(spec["gens"], spec["migrate_every"], spec["pct"])
    out = {"genes": [], "scores": [], "best": []}
    for p in pops:
        out["genes"].append("%016x" % fnv1a(p.genes()))
        out["scores"].append("%016x" % fnv1a(p.scores()))
        idx, s = engine.best_index(p) if hasattr(engine, "best_index") else engine.best(p)
        out["best"].append([int(idx), float(s)])
    return out
