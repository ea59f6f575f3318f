"""Parity of the CUDA engine against the CPU oracle, through the C-ABI.

Integer work (genomes, selection indices, packed bits, emigrant sets) must be
bit-exact.  Float fitness: the acceptance tolerance is 1e-5 relative
(BASELINE.json north_star); the engine is built to be bit-exact there too, and
the tests say which bar they apply.
"""
import numpy as np
import pytest

import libpga_b200 as pg

pytestmark = pytest.mark.gpu

REL_TOL = 1e-5  # north_star: "float fitness must agree within 1e-5 relative"


def assert_scores_close(got, want):
    np.testing.assert_allclose(got, want, rtol=REL_TOL, atol=0)


def same_bits(a, b):
    return np.array_equal(np.ascontiguousarray(a).view(np.uint32), np.ascontiguousarray(b).view(np.uint32))


def setup_pair(pga, oracle, seed, size, length, *, encoding=0, obj=None, builtin=None, xover=None, threads=8):
    g = pga(seed)
    o = oracle.Oracle(seed, threads=threads)
    gp = g.create_population(size, length, encoding=encoding)
    op = o.create_population(size, length, encoding)
    if builtin is not None:
        g.set_builtin_objective(builtin)
        o.set_builtin_objective(builtin)
    if obj is not None:
        g.set_objective_function(obj)
        o.set_objective_hook(obj)
    if xover is not None:
        g.set_crossover_function(xover)
        o.set_crossover_hook(xover)
    return g, o, gp, op


# ---- config 1: the reference's test/test.cu workload (hook objective) ------------

def test_initial_population_matches(pga, oracle):
    g, o, gp, op = setup_pair(pga, oracle, 42, 1000, 100, obj="sum_obj")
    assert same_bits(gp.genes(), op.genes())
    o.close()


@pytest.mark.parametrize("size,length,gens", [(4096, 100, 20), (1000, 37, 10), (33, 4, 5), (1, 8, 3)])
def test_c1_trajectory_every_generation(pga, oracle, size, length, gens):
    g, o, gp, op = setup_pair(pga, oracle, 7, size, length, obj="sum_obj")
    gp.capture_selection(True)
    for gen in range(gens):
        g.run(1)
        o.run(1)
        assert np.array_equal(gp.selection(), op.selection()), f"selection differs in generation {gen}"
        assert same_bits(gp.genes(), op.genes()), f"genomes differ in generation {gen}"
        assert same_bits(gp.scores(), op.scores()), f"scores differ in generation {gen}"
    assert g.best(gp) == o.best_index(op)
    o.close()


def test_c1_full_size_100_generations(pga, oracle):
    # test/test.cu:37,43: 40000 x 100, 100 generations
    g, o, gp, op = setup_pair(pga, oracle, 42, 40000, 100, obj="sum_obj")
    g.run(100)
    o.run(100)
    assert same_bits(gp.genes(), op.genes())
    assert same_bits(gp.scores(), op.scores())
    gi, gs = g.best(gp)
    oi, os_ = o.best_index(op)
    assert (gi, gs) == (oi, os_)
    assert 90.0 < gs <= 100.0  # optimum 100.0
    o.close()


def test_stepwise_equals_fused(pga, oracle):
    # pga_run == the loop of src/pga.cu:381-387 + the trailing evaluate
    a = pga(3)
    b = pga(3)
    pa = a.create_population(3000, 52)
    pb = b.create_population(3000, 52)
    a.set_objective_function("sum_obj")
    b.set_objective_function("sum_obj")
    a.run(6)
    for _ in range(6):
        b.fill_random_values(pb)
        b.evaluate(pb)
        b.crossover(pb)
        b.mutate(pb)
        b.swap_generations(pb)
    b.evaluate(pb)
    assert same_bits(pa.genes(), pb.genes())
    assert same_bits(pa.scores(), pb.scores())


def test_stepwise_matches_oracle_steps(pga, oracle):
    g, o, gp, op = setup_pair(pga, oracle, 11, 2000, 24, obj="sum_obj")
    gp.capture_selection(True)
    g.fill_random_values(gp); o.fill_random_values(op)
    g.evaluate(gp); o.evaluate(op)
    assert same_bits(gp.scores(), op.scores())
    g.crossover(gp); o.crossover(op)
    assert np.array_equal(gp.selection(), op.selection())
    assert same_bits(gp.next_genes(), op.next_genes())
    g.mutate(gp); o.mutate(op)
    assert same_bits(gp.next_genes(), op.next_genes())
    # draws only change on fill_random_values: crossover again gives the same children
    g.crossover(gp)
    g.mutate(gp)
    assert same_bits(gp.next_genes(), op.next_genes())
    o.close()


# ---- test2: knapsack (hook objective, L = 6: padded rows) ------------------------

def test_knapsack_demo(pga, oracle):
    g, o, gp, op = setup_pair(pga, oracle, 5, 100, 6, obj="knapsack_obj")
    g.run(5)
    o.run(5)
    assert same_bits(gp.genes(), op.genes())
    assert same_bits(gp.scores(), op.scores())
    assert g.best(gp) == o.best_index(op)
    o.close()


def test_knapsack_converges(pga):
    g = pga(1)
    p = g.create_population(4096, 6)
    g.set_objective_function("knapsack_obj")
    g.run(30)
    _, s = g.best(p)
    assert s == 285.0  # items 2+3: the optimum SURVEY.md section 4 derives


# ---- test3: TSP-like (hook objective + hook crossover that reads the rand row) ----

def test_tsp_demo(pga, oracle, tsp_matrix):
    g, o, gp, op = setup_pair(pga, oracle, 9, 1000, 100, obj="tsp_obj", xover="tsp_crossover")
    g.tsp_load(tsp_matrix)
    o.tsp_load(tsp_matrix)
    gp.capture_selection(True)
    for gen in range(0, 60, 20):
        g.run(20)
        o.run(20)
        assert np.array_equal(gp.selection(), op.selection()), gen
        assert same_bits(gp.genes(), op.genes()), gen
        assert same_bits(gp.scores(), op.scores()), gen
    assert g.best(gp) == o.best_index(op)
    o.close()


def test_tsp_hooks_lockstep_at_scale(pga, oracle, tsp_matrix):
    # config 4 (test3 hooks) at 2^17 individuals: one lock-step generation from a GPU state
    n = 1 << 17
    g, o, gp, op = setup_pair(pga, oracle, 21, n, 100, obj="tsp_obj", xover="tsp_crossover")
    g.tsp_load(tsp_matrix)
    o.tsp_load(tsp_matrix)
    g.run(3)
    op.set_genes(gp.genes())
    op.set_epoch(gp.epoch)
    gp.capture_selection(True)
    g.run(1)
    o.run(1)
    assert np.array_equal(gp.selection(), op.selection())
    assert same_bits(gp.genes(), op.genes())
    assert same_bits(gp.scores(), op.scores())
    o.close()


# ---- config 2: OneMax on packed bits -----------------------------------------------

@pytest.mark.parametrize("size,bits,gens", [(1 << 14, 1024, 100), (1000, 1000, 10), (77, 33, 6), (50, 4, 4), (300, 5000, 4)])
def test_onemax_trajectory(pga, oracle, size, bits, gens):
    g, o, gp, op = setup_pair(pga, oracle, 42, size, bits, encoding=pg.GENES_BITS, builtin=pg.OBJ_ONEMAX)
    assert np.array_equal(gp.genes(), op.genes())
    gp.capture_selection(True)
    g.run(gens)
    o.run(gens)
    assert np.array_equal(gp.selection(), op.selection())
    assert np.array_equal(gp.genes(), op.genes())
    assert np.array_equal(gp.scores(), op.scores())
    assert g.best(gp) == o.best_index(op)
    o.close()


# every compile-time tiling of the fused kernel (lanes per child x blocks per lane), ragged sizes
BIT_SHAPES = [33, 200, 400, 1000, 2000, 4096, 5000, 12000, 16000, 20000]
F32_SHAPES = [4, 7, 16, 40, 100, 128, 256, 300, 500, 1000]


@pytest.mark.parametrize("bits", BIT_SHAPES)
@pytest.mark.parametrize("generic", [False, True])
def test_onemax_shapes(pga, oracle, monkeypatch, bits, generic):
    if generic:
        monkeypatch.setenv("PGA_FORCE_GENERIC", "1")
    g, o, gp, op = setup_pair(pga, oracle, 13, 211, bits, encoding=pg.GENES_BITS, builtin=pg.OBJ_ONEMAX)
    gp.capture_selection(True)
    g.run(4); o.run(4)
    assert np.array_equal(gp.selection(), op.selection())
    assert np.array_equal(gp.genes(), op.genes())
    assert np.array_equal(gp.scores(), op.scores())
    o.close()


@pytest.mark.parametrize("length", F32_SHAPES)
@pytest.mark.parametrize("generic", [False, True])
def test_rastrigin_shapes(pga, oracle, monkeypatch, length, generic):
    if generic:
        monkeypatch.setenv("PGA_FORCE_GENERIC", "1")
    g, o, gp, op = setup_pair(pga, oracle, 13, 211, length, builtin=pg.OBJ_RASTRIGIN)
    g.run(4); o.run(4)
    assert_scores_close(gp.scores(), op.scores())
    assert same_bits(gp.genes(), op.genes())
    assert same_bits(gp.scores(), op.scores())
    o.close()


def test_onemax_full_size_lockstep(pga, oracle):
    # config 2 at full size: 2^20 x 1024 bits, three lock-step generations
    n, bits = 1 << 20, 1024
    g, o, gp, op = setup_pair(pga, oracle, 42, n, bits, encoding=pg.GENES_BITS, builtin=pg.OBJ_ONEMAX)
    g.run(20)  # get away from the initial population on the GPU alone
    op.set_genes(gp.genes())
    op.set_epoch(gp.epoch)
    for _ in range(3):
        g.run(1)
        o.run(1)
        assert np.array_equal(gp.genes(), op.genes())
        assert np.array_equal(gp.scores(), op.scores())
    # size-independent property: the score is the popcount of the row
    rows = gp.genes()
    pop = np.unpackbits(rows.view(np.uint8), axis=1).sum(axis=1)
    assert np.array_equal(pop.astype(np.float32), gp.scores())
    o.close()


# ---- config 3: Rastrigin, float32 -----------------------------------------------------

@pytest.mark.parametrize("which", [pg.OBJ_RASTRIGIN, pg.OBJ_SUM, pg.OBJ_SPHERE])
@pytest.mark.parametrize("size,length,gens", [(1 << 12, 256, 30), (999, 100, 8), (64, 7, 5), (500, 1000, 3)])
def test_builtin_f32_trajectory(pga, oracle, which, size, length, gens):
    g, o, gp, op = setup_pair(pga, oracle, 42, size, length, builtin=which)
    gp.capture_selection(True)
    g.run(gens)
    o.run(gens)
    assert_scores_close(gp.scores(), op.scores())
    # built to be bit-exact (same expression tree, explicit fma): check the stronger bar too
    assert np.array_equal(gp.selection(), op.selection())
    assert same_bits(gp.genes(), op.genes())
    assert same_bits(gp.scores(), op.scores())
    o.close()


def test_rastrigin_full_size_lockstep(pga, oracle):
    n, length = 1 << 20, 256
    g, o, gp, op = setup_pair(pga, oracle, 42, n, length, builtin=pg.OBJ_RASTRIGIN)
    g.run(5)
    op.set_genes(gp.genes())
    op.set_epoch(gp.epoch)
    g.run(1)
    o.run(1)
    assert_scores_close(gp.scores(), op.scores())
    assert same_bits(gp.genes(), op.genes())
    # against float64 numpy Rastrigin on a sample (the objective means what it says)
    rows = gp.genes()[:2048].astype(np.float64)
    x = (rows - 0.5) * 10.24
    ref = -(10.0 * length + np.sum(x * x - 10.0 * np.cos(2 * np.pi * x), axis=1))
    np.testing.assert_allclose(gp.scores()[:2048], ref, rtol=REL_TOL)
    o.close()


def test_builtin_objective_with_hook_crossover(pga, oracle, tsp_matrix):
    # hook crossover + built-in objective: the hook kernel scores with the canonical reduction
    g, o, gp, op = setup_pair(pga, oracle, 4, 600, 100, builtin=pg.OBJ_RASTRIGIN, xover="tsp_crossover")
    g.tsp_load(tsp_matrix); o.tsp_load(tsp_matrix)
    g.run(5); o.run(5)
    assert same_bits(gp.genes(), op.genes())
    assert same_bits(gp.scores(), op.scores())
    o.close()


@pytest.mark.parametrize("size,length", [(2000, 100), (333, 37), (5000, 10)])
def test_crossover_hook_that_leaves_genes_unwritten(pga, oracle, size, length):
    # src/pga.cu:302: the child row handed to the crossover hook starts as the row next_gen
    # holds at that index; a hook that writes only the even genes leaves the odd ones as they
    # were two generations ago (zero in the first generation: the buffer starts zeroed here
    # and in the oracle, where the reference leaves cudaMalloc's contents)
    g, o, gp, op = setup_pair(pga, oracle, 6, size, length, obj="sum_obj", xover="even_crossover")
    for gen in range(6):
        g.run(1); o.run(1)
        assert same_bits(gp.genes(), op.genes()), gen
        assert same_bits(gp.scores(), op.scores()), gen
    # stepwise too: crossover alone, then mutate alone
    g.fill_random_values(gp); o.fill_random_values(op)
    g.crossover(gp); o.crossover(op)
    assert same_bits(gp.next_genes(), op.next_genes())
    g.mutate(gp); o.mutate(op)
    assert same_bits(gp.next_genes(), op.next_genes())
    o.close()


@pytest.mark.parametrize("xover", [None, "tsp_crossover"])
def test_mutate_hook_reads_its_rand_row(pga, oracle, tsp_matrix, xover):
    # include/pga.h:47, :71-77: a user mutation gets the genome and that individual's rand row
    g, o, gp, op = setup_pair(pga, oracle, 12, 1500, 100, obj="sum_obj", xover=xover)
    g.tsp_load(tsp_matrix); o.tsp_load(tsp_matrix)
    g.set_mutate_function("blend_mutate"); o.set_mutate_hook("blend_mutate")
    for gen in range(5):
        g.run(1); o.run(1)
        assert same_bits(gp.genes(), op.genes()), gen
        assert same_bits(gp.scores(), op.scores()), gen
    g.set_mutate_function(None); o.set_mutate_hook(None)   # NULL restores the default (include/pga.h:76)
    g.run(2); o.run(2)
    assert same_bits(gp.genes(), op.genes())
    o.close()


# ---- API behaviour -------------------------------------------------------------------

def test_create_population_limits(pga):
    g = pga(1)
    assert g.create_population(10, 3) is None           # src/pga.cu:184
    for _ in range(pg.MAX_POPULATIONS):
        assert g.create_population(16, 8) is not None
    assert g.create_population(16, 8) is None           # src/pga.cu:180


def test_get_best_prints_and_returns_row(pga, oracle, capfd):
    g = pga(42, quiet=False)
    p = g.create_population(5000, 20)
    g.set_objective_function("sum_obj")
    g.run(3)
    capfd.readouterr()
    row = g.get_best(p)
    out = capfd.readouterr().out
    idx, s = g.best(p)
    assert out == "%f\n" % s                             # src/pga.cu:230
    assert same_bits(row, p.genes()[idx])
    assert np.float32(row.sum(dtype=np.float32)) == pytest.approx(s, rel=1e-6)


def test_get_best_first_maximum_wins_and_nan_rule(pga):
    g = pga(1)
    p = g.create_population(1000, 8)
    g.set_objective_function("sum_obj")
    s = np.zeros(1000, dtype=np.float32)
    s[[10, 500, 900]] = 5.0
    p.set_scores(s)
    assert g.best(p) == (10, 5.0)
    s[0] = np.nan                                        # src/pga.cu:224: a NaN at index 0 is never displaced
    p.set_scores(s)
    i, v = g.best(p)
    assert i == 0 and np.isnan(v)
    s[0] = 1.0
    s[10] = np.nan
    p.set_scores(s)
    assert g.best(p) == (500, 5.0)


def test_upload_download_roundtrip(pga):
    g = pga(1)
    for length, enc in [(6, 0), (100, 0), (33, 1), (1024, 1)]:
        p = g.create_population(257, length, encoding=enc)
        rng = np.random.default_rng(0)
        if enc:
            rows = rng.integers(0, 2 ** 32, size=(257, p.row_words), dtype=np.uint32)
        else:
            rows = rng.random((257, length), dtype=np.float32)
        p.set_genes(rows)
        assert np.array_equal(p.genes(), rows)


def test_seed_changes_everything_and_same_seed_repeats(pga):
    res = []
    for seed in (1, 1, 2):
        g = pga(seed)
        p = g.create_population(512, 16)
        g.set_objective_function("sum_obj")
        g.run(4)
        res.append(p.genes())
    assert same_bits(res[0], res[1])
    assert not same_bits(res[0], res[2])


def test_null_hook_restores_default_operator(pga, tsp_matrix):
    # include/pga.h:76,83: "if NULL, default random is used"
    a = pga(8); b = pga(8)
    pa = a.create_population(700, 100); pb = b.create_population(700, 100)
    a.tsp_load(tsp_matrix)
    for g in (a, b):
        g.set_objective_function("sum_obj")
    a.set_crossover_function("tsp_crossover")
    a.set_crossover_function(None)
    a.run(4); b.run(4)
    assert same_bits(pa.genes(), pb.genes())


def test_run_zero_generations_only_scores(pga, oracle):
    g, o, gp, op = setup_pair(pga, oracle, 2, 500, 12, obj="sum_obj")
    before = gp.genes()
    g.run(0); o.run(0)                        # src/pga.cu:390: the trailing evaluate still happens
    assert same_bits(gp.genes(), before)
    assert same_bits(gp.scores(), op.scores())
    o.close()


def test_empty_population_is_harmless(pga):
    g = pga(1)
    p = g.create_population(0, 8)
    assert p is not None and p.size == 0
    g.set_objective_function("sum_obj")
    g.run(3)
    assert g.get_best(p) is None


def test_get_best_top_longer_than_population(pga, oracle):
    g, o, gp, op = setup_pair(pga, oracle, 2, 5, 8, obj="sum_obj")
    g.run(1); o.run(1)
    rows = g.get_best_top(gp, 9)
    want = o.get_best_top(op, 9)
    assert len(rows) == 5 and all(same_bits(x, y) for x, y in zip(rows, want))
    o.close()
