"""CPU-only tests: the oracle against published known answers and the committed
golden digests, the host-side logic that runs without a GPU, and the C-ABI surface
(the shared library loads and exports every declared symbol; no compute is called)."""
import ctypes as C
import json
import os
import re
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import cases  # noqa: E402

# ---- Philox4x32-10 known-answer vectors (Random123 kat_vectors) ----
KATS = [
    ([0, 0, 0, 0], [0, 0], [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]),
    ([0xffffffff] * 4, [0xffffffff] * 2, [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]),
    ([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0],
     [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]),
]


@pytest.mark.parametrize("ctr,key,want", KATS)
def test_philox_known_answers(oracle, ctr, key, want):
    assert [int(x) for x in oracle.philox(ctr, key)] == want


def philox_numpy(ctr, key):
    """Independent restatement of Philox4x32-10 (python ints) to pin the draw layout."""
    c = [int(x) for x in ctr]
    k = [int(x) for x in key]
    for _ in range(10):
        p0, p1 = 0xD2511F53 * c[0], 0xCD9E8D57 * c[2]
        c = [(p1 >> 32) ^ c[1] ^ k[0], p1 & 0xFFFFFFFF, (p0 >> 32) ^ c[3] ^ k[1], p0 & 0xFFFFFFFF]
        k = [(k[0] + 0x9E3779B9) & 0xFFFFFFFF, (k[1] + 0xBB67AE85) & 0xFFFFFFFF]
    return c


def test_draw_layout(oracle):
    seed, island, epoch, ind = 0x123456789ABCDEF0, 5, 9, 1234
    row = oracle.rand_row(seed, island, epoch, ind, 10)
    for j in range(10):
        w = philox_numpy([j // 4, ind, epoch, (0 << 24) | island], [seed & 0xFFFFFFFF, seed >> 32])[j % 4]
        assert row[j] == np.float32((w >> 8) * 2.0 ** -24)
        assert 0.0 <= row[j] < 1.0
    L = oracle.lib()
    w = L.opga_draw_word(seed, island, epoch, ind, oracle.STREAM_MASK, 6)
    assert w == philox_numpy([1, ind, epoch, (1 << 24) | island], [seed & 0xFFFFFFFF, seed >> 32])[2]


def test_index_draw_never_reaches_size(oracle):
    # (int)(u * size) < size for the largest uniform and every size the demos use
    u = np.float32(1.0) - np.float32(2.0 ** -24)
    for size in (4, 6, 100, 1000, 40000, 1 << 20, (1 << 22) + 1, (1 << 24) - 1):
        assert int(np.float32(u * np.float32(size))) < size


def test_cos2pi_accuracy(oracle):
    L = oracle.lib()
    xs = np.linspace(-5.12, 5.12, 20001).astype(np.float32)
    err = max(abs(L.opga_cos2pi(float(x)) - np.cos(2 * np.pi * float(x))) for x in xs)
    assert err < 3e-7


def test_rastrigin_objective_meaning(oracle):
    rng = np.random.default_rng(0)
    for length in (4, 7, 100, 256, 1000):
        g = rng.random(length, dtype=np.float32)
        x = (g.astype(np.float64) - 0.5) * 10.24
        want = -(10.0 * length + np.sum(x * x - 10.0 * np.cos(2 * np.pi * x)))
        got = oracle.builtin_objective_f32(oracle.OBJ_RASTRIGIN, g)
        assert got == pytest.approx(want, rel=1e-5)
        assert oracle.builtin_objective_f32(oracle.OBJ_SUM, g) == pytest.approx(float(g.astype(np.float64).sum()), rel=1e-5)


# ---- golden trajectories -----------------------------------------------------------------

with open(os.path.join(ROOT, "tests", "golden", "oracle_digests.json")) as f:
    GOLDEN = json.load(f)


@pytest.mark.parametrize("name", sorted(cases.CASES))
def test_oracle_reproduces_golden_digests(oracle, name):
    spec = cases.CASES[name]
    o = oracle.Oracle(spec["seed"], threads=os.cpu_count() or 1)
    got = cases.run_case(o, spec, lambda n, l, e: o.create_population(n, l, e))
    o.close()
    assert got == GOLDEN[name]


def test_serial_and_openmp_oracle_agree(oracle):
    res = []
    for threads in (1, 4):
        o = oracle.Oracle(11, threads=threads)
        p = o.create_population(3000, 64, oracle.GENES_BITS)
        o.set_builtin_objective(oracle.OBJ_ONEMAX)
        o.run(6)
        res.append((p.genes(), p.scores()))
        o.close()
    assert np.array_equal(res[0][0], res[1][0]) and np.array_equal(res[0][1], res[1][1])


def test_demo_optima_are_reachable(oracle, tsp_matrix):
    # derivable optima 100.0 / 285 / -990
    o = oracle.Oracle(1, threads=4)
    p = o.create_population(4096, 6)
    o.set_objective_hook("knapsack_obj")
    o.run(30)
    assert o.best_index(p)[1] == 285.0
    o.close()
    L = oracle.lib()
    o = oracle.Oracle(1)
    o.tsp_load(tsp_matrix)
    tour = ((np.arange(100, dtype=np.float32) + 0.5) / 100).astype(np.float32)  # 0 -> 1 -> ... -> 99
    obj = C.CFUNCTYPE(C.c_float, C.c_void_p, C.c_uint)(L.opga_demo_hook(b"tsp_obj"))
    # rows 91..99 of the device matrix stay zero (test3/test.cu:79 copies a 10000-float prefix of a 110-wide array)
    assert obj(tour.ctypes.data_as(C.c_void_p), 100) == -(91 * 10.0)
    o.close()


def test_oracle_edge_cases(oracle):
    o = oracle.Oracle(1)
    assert o.create_population(10, 3) is None          # src/pga.cu:184
    for _ in range(10):
        assert o.create_population(4, 4) is not None
    assert o.create_population(4, 4) is None           # src/pga.cu:180
    o.close()
    # get_best tie and NaN rules (src/pga.cu:221-229)
    o = oracle.Oracle(1)
    p = o.create_population(8, 4)
    s = np.array([1, 5, 5, 2, 5, 0, 0, 0], dtype=np.float32)
    p.set_scores(s)
    assert o.best_index(p) == (1, 5.0)
    s[0] = np.nan
    p.set_scores(s)
    assert o.best_index(p)[0] == 0
    o.close()


def test_tournament_sizes(oracle):
    # src/pga.cu:280-292 with TOURNAMENT_POPULATION = K: candidates from draws 0..K-1 and K..2K-1
    seed = 17
    for k in (1, 2, 3, 5):
        o = oracle.Oracle(seed)
        p = o.create_population(400, 16)
        o.set_objective_hook("sum_obj")
        o.set_tournament(k)
        o.fill_random_values(p)
        o.evaluate(p)
        o.crossover(p)
        s, sel = p.scores(), p.selection()
        for i in (0, 7, 399):
            u = oracle.rand_row(seed, 0, p.epoch, i, 2 * k)
            cand = [min(int(np.float32(x) * np.float32(400)), 399) for x in u]
            for which, group in enumerate((cand[:k], cand[k:])):
                best = group[0]
                for c in group[1:]:
                    if s[best] < s[c]:
                        best = c
                assert sel[i, which] == best
        o.close()


def test_target_stop(oracle):
    # include/pga.h:141: "until n-generations or obj_func(best_genome) == value"
    a = oracle.Oracle(3)
    pa = a.create_population(512, 128, oracle.GENES_BITS)
    a.set_builtin_objective(oracle.OBJ_ONEMAX)
    a.set_target(84.0, 2)
    a.run(100)
    g = a.generations_run
    assert 0 < g < 100 and g % 2 == 0
    assert a.best_index(pa)[1] >= 84.0
    b = oracle.Oracle(3)
    pb = b.create_population(512, 128, oracle.GENES_BITS)
    b.set_builtin_objective(oracle.OBJ_ONEMAX)
    b.run(g)                               # the stopped run is the plain run of g generations
    assert np.array_equal(pa.genes(), pb.genes())
    b2 = oracle.Oracle(3)
    p2 = b2.create_population(512, 128, oracle.GENES_BITS)
    b2.set_builtin_objective(oracle.OBJ_ONEMAX)
    b2.run(g - 2)                          # and two generations earlier the target was not met
    assert b2.best_index(p2)[1] < 84.0
    for x in (a, b, b2):
        x.close()


def test_select_top_order_and_ties(oracle):
    s = np.array([3, 7, 7, 1, np.nan, 7, 0, -0.0], dtype=np.float32)
    assert list(oracle.select_top(s, 2)) == [1, 2]              # lower index wins ties
    assert list(oracle.select_top(s, 4)) == [0, 1, 2, 5]
    assert list(oracle.select_top(s, 1, worst=True)) == [4]     # NaN ranks last
    assert list(oracle.select_top(s, 3, worst=True)) == [4, 6, 7]  # 0 == -0: higher index is worse
    assert oracle.lib().opga_migration_count(1 << 22, 1.0) == 41943


# ---- host-side logic of the product that needs no GPU ------------------------------------------

def test_shared_library_exports_every_declared_symbol():
    from libpga_b200 import _lib
    lib = _lib.load()  # raises if the extension is not built
    declared = set()
    for header in ("pga.h", "pga_b200.h"):
        text = open(os.path.join(ROOT, "include", header)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        declared |= set(re.findall(r"\b(pgax?_[a-z0-9_]+)\s*\(", text))
    declared -= {"pgax_transfer"}
    assert len(declared) >= 22 + 30
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} is declared in include/ but not exported"
        assert name in _lib.SIGNATURES, f"{name} has no ctypes signature"
    assert lib.pgax_version().startswith(b"libpga-b200")


def test_static_archive_has_all_pga_h_symbols():
    import subprocess
    path = os.path.join(ROOT, "libpga.a")
    assert os.path.exists(path), "libpga.a not built"
    out = subprocess.run(["nm", "-g", "--defined-only", path], capture_output=True, text=True).stdout
    text = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "pga.h")).read(), flags=re.S)
    for name in set(re.findall(r"\b(pga_[a-z0-9_]+)\s*\(", text)):
        assert re.search(r"\bT %s\b" % name, out), name


def test_plan_exchange_ring_and_random(oracle):
    from libpga_b200 import islands
    for world, k in ((1, 4), (2, 2), (4, 2), (8, 1)):
        total = world * k
        plan = islands.plan_exchange(42, 0, 0, world, k)
        assert [(t["src_island"], t["dst_island"]) for t in plan] == [(g, (g + 1) % total) for g in range(total)]
        for t in plan:
            assert t["src_rank"] == t["src_island"] % world and t["src_local"] == t["src_island"] // world
            assert t["dst_rank"] == t["dst_island"] % world and t["dst_local"] == t["dst_island"] // world
        for rnd in range(5):
            perm = oracle.random_permutation(42, rnd, total)
            plan = islands.plan_exchange(42, rnd, 1, world, k)
            assert [(t["src_island"], t["dst_island"]) for t in plan] == [(g, int(perm[g])) for g in range(total) if perm[g] != g]
            assert sorted(perm) == list(range(total))


def test_migration_count_matches_oracle(oracle):
    import libpga_b200 as pg
    for size, pct in ((1 << 22, 1.0), (1000, 2.5), (100, 0.0), (100, -1.0), (100, 150.0), (7, 33.3)):
        assert pg.migration_count(size, pct) == oracle.lib().opga_migration_count(size, pct)


def test_missing_extension_fails_loudly(monkeypatch):
    # no CPU fallback: without the built library the product refuses to load
    from libpga_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "SO_PATH", os.path.join(ROOT, "libpga_b200", "does_not_exist.so"))
    with pytest.raises(RuntimeError, match="not built"):
        _lib.load()


def test_product_never_touches_the_oracle():
    # the oracle is test infrastructure: nothing under libpga_b200/ may import, load or name it
    import glob
    for path in glob.glob(os.path.join(ROOT, "libpga_b200", "**", "*"), recursive=True):
        if os.path.isfile(path) and path.endswith((".py", ".c", ".cu", ".cuh", ".h")):
            text = open(path).read()
            assert "pyoracle" not in text and "pga_oracle" not in text and "opga_" not in text, path
