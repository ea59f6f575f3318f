"""What include/pga.h promises in comments and the reference never built (SURVEY section
8(f) ranks 3 and 4): selection schemes behind enum crossover_selection_type
(include/pga.h:36-42, :103), the early stop of pga_run / pga_run_islands (:141, :147), and
population dump / restore.  The reference has no code for any of them, so the oracle
function is the specification (parity unpinned by the reference); everything is integer /
bit work or bit-exact float work, compared exactly.
"""
import numpy as np
import pytest

import libpga_b200 as pg

pytestmark = pytest.mark.gpu


def same_bits(a, b):
    return np.array_equal(np.ascontiguousarray(a).view(np.uint32), np.ascontiguousarray(b).view(np.uint32))


def pair(pga, oracle, seed, size, length, *, bits=False, obj=None, builtin=None, count=1):
    g = pga(seed)
    o = oracle.Oracle(seed, threads=8)
    enc = pg.GENES_BITS if bits else pg.GENES_F32
    gps = [g.create_population(size, length, encoding=enc) for _ in range(count)]
    ops = [o.create_population(size, length, enc) for _ in range(count)]
    if builtin is not None:
        g.set_builtin_objective(builtin); o.set_builtin_objective(builtin)
    if obj is not None:
        g.set_objective_function(obj); o.set_objective_hook(obj)
    return g, o, gps, ops


# ---- selection schemes --------------------------------------------------------------

@pytest.mark.parametrize("k", [1, 2, 3, 4, 7, 16])
@pytest.mark.parametrize("kind", ["hook", "bits", "f32"])
def test_tournament_of_k_matches_oracle(pga, oracle, k, kind):
    if kind == "hook":
        g, o, (gp,), (op,) = pair(pga, oracle, 31, 1500, 40, obj="sum_obj")
    elif kind == "bits":
        g, o, (gp,), (op,) = pair(pga, oracle, 31, 1500, 1024, bits=True, builtin=pg.OBJ_ONEMAX)
    else:
        g, o, (gp,), (op,) = pair(pga, oracle, 31, 1500, 256, builtin=pg.OBJ_RASTRIGIN)
    g.set_selection(pg.tournament_of(k)); o.set_tournament(k)
    gp.capture_selection(True)
    for gen in range(5):
        g.run(1); o.run(1)
        assert np.array_equal(gp.selection(), op.selection()), gen
        assert same_bits(gp.genes(), op.genes()), gen
        assert same_bits(gp.scores(), op.scores()), gen
    o.close()


def test_selection_pressure_grows_with_tournament_size(pga):
    # size-independent property: after the same number of generations a larger tournament
    # has a better mean fitness (k = 1 is a random walk)
    means = []
    for k in (1, 2, 4):
        g = pga(5)
        p = g.create_population(1 << 14, 512, encoding=pg.GENES_BITS)
        g.set_builtin_objective(pg.OBJ_ONEMAX)
        g.set_selection(pg.tournament_of(k))
        g.run(15)
        means.append(float(p.scores().mean()))
    assert means[0] < means[1] < means[2]
    assert abs(means[0] - 256.0) < 2.0


def test_pga_crossover_honours_its_type_argument(pga, oracle):
    # src/pga.cu:319 drops `type`; here the call's own argument selects the scheme
    g, o, (gp,), (op,) = pair(pga, oracle, 8, 900, 24, obj="sum_obj")
    gp.capture_selection(True)
    g.fill_random_values(gp); o.fill_random_values(op)
    g.evaluate(gp); o.evaluate(op)
    g.crossover(gp, pg.tournament_of(5)); o.set_tournament(5); o.crossover(op)
    assert np.array_equal(gp.selection(), op.selection())
    assert same_bits(gp.next_genes(), op.next_genes())
    g.crossover(gp, pg.TOURNAMENT); o.set_tournament(2); o.crossover(op)   # the reference's scheme
    assert np.array_equal(gp.selection(), op.selection())
    g.crossover_all(pg.tournament_of(3)); o.set_tournament(3); o.crossover_all()
    assert np.array_equal(gp.selection(), op.selection())
    assert g.L.pgax_set_selection(g.h, 1) == -1          # MAX_SELECTION_TYPE is not a scheme
    assert g.L.pgax_set_selection(g.h, 0x100 + 17) == -1  # beyond PGAX_MAX_TOURNAMENT
    o.close()


# ---- early stop ------------------------------------------------------------------------

@pytest.mark.parametrize("every", [1, 3, 10])
def test_run_stops_at_the_target(pga, oracle, every):
    g, o, (gp,), (op,) = pair(pga, oracle, 42, 4096, 256, bits=True, builtin=pg.OBJ_ONEMAX)
    target = 160.0   # reached after a dozen generations or so
    g.set_target(target, every); o.set_target(target, every)
    g.run(200); o.run(200)
    assert g.generations_run == o.generations_run
    assert 0 < g.generations_run < 200 and g.generations_run % every == 0
    assert np.array_equal(gp.genes(), op.genes())
    assert np.array_equal(gp.scores(), op.scores())
    assert g.best(gp) == o.best_index(op)
    assert g.best(gp)[1] >= target
    assert gp.epoch == op.epoch == g.generations_run
    # the solver is still usable: disarm and go on in lock-step
    g.set_target(0.0, 0); o.set_target(0.0, 0)
    g.run(3); o.run(3)
    assert g.generations_run == 3
    assert np.array_equal(gp.genes(), op.genes())
    o.close()


def test_target_never_reached_runs_all_generations(pga, oracle):
    g, o, (gp,), (op,) = pair(pga, oracle, 1, 1000, 64, obj="sum_obj")
    g.set_target(1e9, 2); o.set_target(1e9, 2)
    g.run(9); o.run(9)
    assert g.generations_run == o.generations_run == 9
    assert same_bits(gp.genes(), op.genes())
    assert same_bits(gp.scores(), op.scores())
    o.close()


@pytest.mark.parametrize("delay", [0, 1])
def test_run_islands_stops_at_the_target(pga, oracle, delay):
    g, o, gps, ops = pair(pga, oracle, 9, 2048, 256, bits=True, builtin=pg.OBJ_ONEMAX, count=3)
    for x in (g, o):
        x.set_topology(pg.TOPO_RING)
        x.set_migration_delay(delay)
        x.set_target(158.0, 2)
    g.run_islands(300, 4, 3.0); o.run_islands(300, 4, 3.0)
    assert g.generations_run == o.generations_run
    assert 0 < g.generations_run < 300
    for gp, op in zip(gps, ops):
        assert np.array_equal(gp.genes(), op.genes())
        assert np.array_equal(gp.scores(), op.scores())
        assert g.best(gp) == o.best_index(op)
    assert max(g.best(gp)[1] for gp in gps) >= 158.0
    o.close()


# ---- dump / restore ----------------------------------------------------------------------

@pytest.mark.parametrize("bits", [True, False])
def test_save_and_restore_resume_the_same_trajectory(pga, oracle, tmp_path, bits):
    length, builtin = (1000, pg.OBJ_ONEMAX) if bits else (37, pg.OBJ_RASTRIGIN)
    g, o, gps, ops = pair(pga, oracle, 77, 3000, length, bits=bits, builtin=builtin, count=2)
    for x in (g, o):
        x.set_topology(pg.TOPO_RANDOM)
    g.run_islands(7, 3, 2.0); o.run_islands(7, 3, 2.0)
    path = tmp_path / "islands.pga"
    g.save(path)
    g.run_islands(6, 2, 2.0); o.run_islands(6, 2, 2.0)

    # a fresh, empty solver: the file creates the populations
    h = pga(1)
    pops = h.restore(path)
    assert len(pops) == 2 and pops[0].size == 3000 and pops[0].genome_len == length
    assert h.seed == 77 and pops[0].epoch == 7
    h.run_islands(6, 2, 2.0)
    for hp, gp, op in zip(pops, gps, ops):
        assert same_bits(hp.genes(), gp.genes())
        assert same_bits(hp.scores(), gp.scores())
        assert same_bits(hp.genes(), op.genes())      # == the oracle's uninterrupted run
        assert h.best(hp) == o.best_index(op)
    o.close()


def test_restore_into_existing_populations_and_rejects_other_shapes(pga, tmp_path):
    a = pga(3)
    pa = a.create_population(500, 20)
    a.set_objective_function("sum_obj")
    a.run(4)
    path = tmp_path / "one.pga"
    a.save(path)
    b = pga(9)
    pb = b.create_population(500, 20)
    b.set_objective_function("sum_obj")
    b.restore(path)
    a.run(3); b.run(3)
    assert same_bits(pa.genes(), pb.genes())
    c = pga(9)
    c.create_population(501, 20)
    with pytest.raises(OSError):
        c.restore(path)
    with pytest.raises(OSError):
        c.restore(tmp_path / "missing.pga")
