"""The reference's OWN kernels against the oracle and against this library, bit for bit.

oracle/_ref/pin/pinN (recipe: oracle/build_ref.sh, harness: oracle/ref_pin.cu) is the
unmodified /root/reference/src/pga.cu plus the reference's unmodified demo hooks,
with cuRAND replaced at link time by oracle/curand_shim.c, which serves the oracle's
Philox rand rows through the reference's four cuRAND calls (src/pga.cu:100, :153-154,
:167).  The reference thereby becomes deterministic and draws what the oracle draws.
The harness drives the reference's stepwise API (pga_fill_random_values /
pga_evaluate / pga_crossover / pga_mutate / pga_swap_generations, src/pga.cu:103-105,
:264-270, :319-325, :349-354, :362-366) and pga_run (:376-391), reads current_gen /
next_gen / score back through a mirror of struct population (:37-46) and compares
every word with oracle/pga_oracle.c after every step; it prints a digest per
generation.  Here the same workload is then replayed on this repository's CUDA
library through the C-ABI and must produce the same digests: reference == oracle ==
product, on the three workloads the reference ships (test/, test2/, test3/).

The binaries are built in the development container (the reference checkout does
not exist on the GPU box) and travel in oracle/_ref/.
"""
import os
import re
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PIN = os.path.join(ROOT, "oracle", "_ref", "pin")
TSP = os.path.join(ROOT, "tests", "golden", "tsp100.txt")


def run_pin(case, seed, *extra):
    path = os.path.join(PIN, f"pin{case}")
    if not os.path.exists(path):
        pytest.skip(f"{path} not built (run oracle/build_ref.sh where /root/reference exists)")
    cmd = [path, "--seed", str(seed), *extra]
    if case == 3:
        cmd += ["--instance", TSP]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-4000:] + r.stderr[-2000:]
    assert "PIN OK" in r.stdout and "MISMATCH" not in r.stdout
    return r.stdout


def parse(out):
    gens = [(int(m.group(1)), int(m.group(2), 16), int(m.group(3), 16))
            for m in re.finditer(r"^GEN (\d+) genes ([0-9a-f]+) scores_before ([0-9a-f]+)$", out, re.M)]
    m = re.search(r"^RUN fills (\d+) genes ([0-9a-f]+) scores ([0-9a-f]+)$", out, re.M)
    run = (int(m.group(1)), int(m.group(2), 16), int(m.group(3), 16))
    ref_best = out.split("REF_BEST_BEGIN\n")[1].split("REF_BEST_END")[0]
    oracle_best = re.search(r"^ORACLE_BEST (.*)$", out, re.M).group(1) + "\n"
    return gens, run, ref_best, oracle_best


def replay_on_product(pga, oracle, seed, size, length, gens, run_gens, obj, xover=None, tsp=None):
    """The harness's two phases on this library: digests per generation, then pga_run."""
    g = pga(seed)
    p = g.create_population(size, length)
    g.set_objective_function(obj)
    if xover:
        g.set_crossover_function(xover)
    if tsp is not None:
        g.tsp_load(tsp)
    digests = []
    for gen in range(1, gens + 1):
        g.fill_random_values(p)
        g.evaluate(p)
        scores_before = oracle.digest(p.scores())
        g.crossover(p)
        g.mutate(p)
        g.swap_generations(p)
        digests.append((gen, oracle.digest(p.genes()), scores_before))
    g2 = pga(seed)
    p2 = g2.create_population(size, length)
    g2.set_objective_function(obj)
    if xover:
        g2.set_crossover_function(xover)
    g2.run(run_gens)
    _, best = g2.best(p2)
    return digests, (oracle.digest(p2.genes()), oracle.digest(p2.scores())), "%f\n" % best


def check(out, digests, run, best_line, fills_per_run):
    gens, (fills, run_genes, run_scores), ref_best, oracle_best = parse(out)
    # the reference's own printed line (src/pga.cu:230) == the oracle's == the product's
    assert ref_best == oracle_best == best_line
    # one curandGenerateUniform per pga_fill_random_values + the one in pga_create_population (src/pga.cu:116)
    assert fills == fills_per_run
    assert gens == digests, "per-generation digests of the reference and of this library differ"
    assert (run_genes, run_scores) == run


def test_reference_test1_lockstep(pga, oracle):
    # test/test.cu:37-45: 40000 x 100, sum objective, 100 generations
    out = run_pin(1, 42)
    d, run, best = replay_on_product(pga, oracle, 42, 40000, 100, 100, 100, "sum_obj")
    check(out, d, run, best, 101)


def test_reference_test2_lockstep(pga, oracle):
    # test2/test.cu:43-58: knapsack, 100 x 6, 5 generations
    out = run_pin(2, 7)
    d, run, best = replay_on_product(pga, oracle, 7, 100, 6, 5, 5, "knapsack_obj")
    check(out, d, run, best, 6)


def test_reference_test3_lockstep(pga, oracle, tsp_matrix):
    # test3/test.cu:69-110: 1000 x 100, the reference's crossover and O(L^2) objective, 1000 generations
    out = run_pin(3, 3)
    d, run, best = replay_on_product(pga, oracle, 3, 1000, 100, 60, 1000, "tsp_obj", "tsp_crossover", tsp_matrix)
    check(out, d, run, best, 1001)


@pytest.mark.parametrize("size,seed", [(999, 5), (513, 11)])
def test_reference_ragged_sizes(pga, oracle, size, seed):
    # sizes that are not a multiple of the reference's 512-thread launch stride (src/pga.cu:62-70)
    out = run_pin(1, seed, "--size", str(size), "--gens", "12")
    d, run, best = replay_on_product(pga, oracle, seed, size, 100, 12, 12, "sum_obj")
    check(out, d, run, best, 13)
