"""ctypes binding of the CPU oracle (oracle/pga_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  The product package
(libpga_b200/) never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libpga_oracle.so")

GENES_F32, GENES_BITS = 0, 1
OBJ_HOOK, OBJ_SUM, OBJ_RASTRIGIN, OBJ_SPHERE, OBJ_ONEMAX = 0, 1, 2, 3, 4
TOPO_RING, TOPO_RANDOM = 0, 1
STREAM_ROW, STREAM_MASK, STREAM_INITBITS, STREAM_MIGRATE = 0, 1, 2, 3


def build(force=False):
    src = [os.path.join(_HERE, f) for f in ("pga_oracle.c", "pga_oracle.h")]
    if (not force and os.path.exists(_SO)
            and all(os.path.getmtime(_SO) >= os.path.getmtime(s) for s in src)):
        return _SO
    subprocess.check_call(["make", "-C", _HERE, "_build/libpga_oracle.so"],
                          stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    L = C.CDLL(build())
    vp, u64, u32, i32, f32 = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int, C.c_float
    ul, ui = C.c_ulong, C.c_uint
    sig = {
        "opga_philox4x32_10": (None, [vp, vp, vp]),
        "opga_draw_word": (u32, [u64, u32, u32, u32, u32, u32]),
        "opga_draw_uniform": (f32, [u64, u32, u32, u32, u32]),
        "opga_rand_row": (None, [u64, u32, u32, u32, u32, vp]),
        "opga_cos2pi": (f32, [f32]),
        "opga_builtin_objective_f32": (f32, [i32, vp, ui]),
        "opga_builtin_objective_bits": (f32, [i32, vp, ui]),
        "opga_init": (vp, [u64]),
        "opga_deinit": (None, [vp]),
        "opga_set_threads": (None, [vp, i32]),
        "opga_set_island_base": (None, [vp, u32]),
        "opga_set_quiet": (None, [vp, i32]),
        "opga_create_population": (vp, [vp, ul, ui, i32]),
        "opga_create_population_on_island": (vp, [vp, ul, ui, i32, u32]),
        "opga_set_objective_function": (None, [vp, vp]),
        "opga_set_builtin_objective": (None, [vp, i32]),
        "opga_set_mutate_function": (None, [vp, vp]),
        "opga_set_crossover_function": (None, [vp, vp]),
        "opga_get_best": (vp, [vp, vp]),
        "opga_best_index": (C.c_long, [vp, vp, vp]),
        "opga_get_best_top": (vp, [vp, vp, ui]),
        "opga_get_best_all": (vp, [vp]),
        "opga_get_best_top_all": (vp, [vp, ui]),
        "opga_evaluate": (None, [vp, vp]),
        "opga_evaluate_all": (None, [vp]),
        "opga_crossover": (None, [vp, vp]),
        "opga_crossover_all": (None, [vp]),
        "opga_mutate": (None, [vp, vp]),
        "opga_mutate_all": (None, [vp]),
        "opga_swap_generations": (None, [vp, vp]),
        "opga_fill_random_values": (None, [vp, vp]),
        "opga_run": (None, [vp, ui]),
        "opga_set_tournament": (None, [vp, ui]),
        "opga_set_target": (None, [vp, f32, ui]),
        "opga_generations_run": (ui, [vp]),
        "opga_migrate": (None, [vp, f32]),
        "opga_migrate_between": (None, [vp, vp, vp, f32]),
        "opga_set_topology": (None, [vp, i32]),
        "opga_set_migration_delay": (None, [vp, ui]),
        "opga_run_islands": (None, [vp, ui, ui, f32]),
        "opga_pop_size": (ul, [vp]),
        "opga_pop_genome_len": (ui, [vp]),
        "opga_pop_row_words": (ui, [vp]),
        "opga_pop_epoch": (u32, [vp]),
        "opga_pop_current": (vp, [vp]),
        "opga_pop_next": (vp, [vp]),
        "opga_pop_scores": (vp, [vp]),
        "opga_pop_selection": (vp, [vp]),
        "opga_pop_set_current": (None, [vp, vp]),
        "opga_pop_set_scores": (None, [vp, vp]),
        "opga_pop_set_epoch": (None, [vp, u32]),
        "opga_select_top": (None, [vp, ul, ul, i32, vp]),
        "opga_migration_count": (ul, [ul, f32]),
        "opga_random_permutation": (None, [u64, u32, ui, vp]),
        "opga_demo_tsp_load": (None, [vp, i32]),
        "opga_demo_hook": (vp, [C.c_char_p]),
        "opga_digest": (u64, [vp, C.c_size_t]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype, fn.argtypes = res, args
    L.free = C.CDLL(None).free
    L.free.argtypes = [vp]
    _lib = L
    return L


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def philox(ctr, key):
    c = np.asarray(ctr, dtype=np.uint32)
    k = np.asarray(key, dtype=np.uint32)
    out = np.zeros(4, dtype=np.uint32)
    lib().opga_philox4x32_10(_ptr(c), _ptr(k), _ptr(out))
    return out


def rand_row(seed, island, epoch, individual, length):
    out = np.zeros(length, dtype=np.float32)
    lib().opga_rand_row(seed, island, epoch, individual, length, _ptr(out))
    return out


def select_top(scores, k, worst=False):
    s = np.ascontiguousarray(scores, dtype=np.float32)
    k = min(int(k), s.size)
    out = np.zeros(max(k, 1), dtype=np.uint32)
    lib().opga_select_top(_ptr(s), s.size, k, int(bool(worst)), _ptr(out))
    return out[:k]


def random_permutation(seed, rnd, count):
    out = np.zeros(count, dtype=np.uint32)
    lib().opga_random_permutation(seed, rnd, count, _ptr(out))
    return out


def builtin_objective_f32(which, genes):
    g = np.ascontiguousarray(genes, dtype=np.float32)
    return float(lib().opga_builtin_objective_f32(which, _ptr(g), g.size))


def digest(a):
    a = np.ascontiguousarray(a)
    return int(lib().opga_digest(_ptr(a), a.nbytes))


class Population:
    def __init__(self, owner, handle):
        self.owner, self.h = owner, handle
        L = lib()
        self.size = L.opga_pop_size(handle)
        self.genome_len = L.opga_pop_genome_len(handle)
        self.row_words = L.opga_pop_row_words(handle)

    def _view(self, addr, dtype, shape):
        n = int(np.prod(shape))
        if n == 0:
            return np.zeros(shape, dtype=dtype)
        buf = (C.c_uint32 * n).from_address(addr)
        return np.frombuffer(buf, dtype=dtype).reshape(shape)

    def genes(self, dtype=None):
        """Current generation, [size][row_words] (float32 genes or packed uint32 bits)."""
        dt = dtype or (np.uint32 if self._enc == GENES_BITS else np.float32)
        return self._view(lib().opga_pop_current(self.h), dt, (self.size, self.row_words)).copy()

    def next_genes(self, dtype=None):
        dt = dtype or (np.uint32 if self._enc == GENES_BITS else np.float32)
        return self._view(lib().opga_pop_next(self.h), dt, (self.size, self.row_words)).copy()

    def scores(self):
        return self._view(lib().opga_pop_scores(self.h), np.float32, (self.size,)).copy()

    def selection(self):
        return self._view(lib().opga_pop_selection(self.h), np.int32, (self.size, 2)).copy()

    @property
    def epoch(self):
        return lib().opga_pop_epoch(self.h)

    def set_genes(self, rows):
        dt = np.uint32 if self._enc == GENES_BITS else np.float32
        r = np.ascontiguousarray(rows, dtype=dt)
        assert r.size == self.size * self.row_words
        lib().opga_pop_set_current(self.h, _ptr(r))

    def set_scores(self, s):
        s = np.ascontiguousarray(s, dtype=np.float32)
        assert s.size == self.size
        lib().opga_pop_set_scores(self.h, _ptr(s))

    def set_epoch(self, e):
        lib().opga_pop_set_epoch(self.h, e)


class Oracle:
    """Mirror of pga.h on the CPU oracle; method names follow the pga_* functions."""

    def __init__(self, seed, threads=1, island_base=0, quiet=True):
        self.L = lib()
        self.h = self.L.opga_init(seed)
        self.L.opga_set_threads(self.h, threads)
        self.L.opga_set_island_base(self.h, island_base)
        self.L.opga_set_quiet(self.h, int(quiet))
        self.pops = []

    def close(self):
        if self.h:
            self.L.opga_deinit(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def create_population(self, size, genome_len, encoding=GENES_F32, island=None):
        if island is None:
            h = self.L.opga_create_population(self.h, size, genome_len, encoding)
        else:
            h = self.L.opga_create_population_on_island(self.h, size, genome_len, encoding, island)
        if not h:
            return None
        p = Population(self, h)
        p._enc = encoding
        self.pops.append(p)
        return p

    def set_objective_hook(self, name):
        self.L.opga_set_objective_function(self.h, self.L.opga_demo_hook(name.encode()))

    def set_builtin_objective(self, which):
        self.L.opga_set_builtin_objective(self.h, which)

    def set_crossover_hook(self, name):
        self.L.opga_set_crossover_function(self.h, self.L.opga_demo_hook(name.encode()) if name else None)

    def set_mutate_hook(self, name):
        self.L.opga_set_mutate_function(self.h, self.L.opga_demo_hook(name.encode()) if name else None)

    def tsp_load(self, matrix):
        m = np.ascontiguousarray(matrix, dtype=np.float32)
        n = int(round(m.size ** 0.5))
        self.L.opga_demo_tsp_load(_ptr(m), n)

    def evaluate(self, pop): self.L.opga_evaluate(self.h, pop.h)
    def evaluate_all(self): self.L.opga_evaluate_all(self.h)
    def crossover(self, pop): self.L.opga_crossover(self.h, pop.h)
    def crossover_all(self): self.L.opga_crossover_all(self.h)
    def mutate(self, pop): self.L.opga_mutate(self.h, pop.h)
    def mutate_all(self): self.L.opga_mutate_all(self.h)
    def swap_generations(self, pop): self.L.opga_swap_generations(self.h, pop.h)
    def fill_random_values(self, pop): self.L.opga_fill_random_values(self.h, pop.h)
    def run(self, n): self.L.opga_run(self.h, n)
    def set_tournament(self, k): self.L.opga_set_tournament(self.h, k)
    def set_target(self, value, check_every=1): self.L.opga_set_target(self.h, value, check_every)

    @property
    def generations_run(self): return int(self.L.opga_generations_run(self.h))

    def migrate(self, pct): self.L.opga_migrate(self.h, pct)
    def migrate_between(self, a, b, pct): self.L.opga_migrate_between(self.h, a.h, b.h, pct)
    def set_topology(self, t): self.L.opga_set_topology(self.h, t)
    def set_migration_delay(self, d): self.L.opga_set_migration_delay(self.h, d)
    def run_islands(self, n, m, pct): self.L.opga_run_islands(self.h, n, m, pct)

    def best_index(self, pop):
        s = C.c_float(0)
        i = self.L.opga_best_index(self.h, pop.h, C.byref(s))
        return int(i), float(s.value)

    def _take_row(self, addr, length):
        row = np.frombuffer((C.c_float * length).from_address(addr), dtype=np.float32).copy()
        self.L.free(addr)
        return row

    def get_best(self, pop):
        addr = self.L.opga_get_best(self.h, pop.h)
        return self._take_row(addr, pop.genome_len)

    def get_best_top(self, pop, length):
        length = min(length, pop.size)
        arr = self.L.opga_get_best_top(self.h, pop.h, length)
        ptrs = (C.c_void_p * max(length, 1)).from_address(arr)
        rows = [self._take_row(ptrs[i], pop.genome_len) for i in range(length)]
        self.L.free(arr)
        return rows

    def get_best_all(self):
        addr = self.L.opga_get_best_all(self.h)
        return self._take_row(addr, self.pops[0].genome_len)

    def get_best_top_all(self, length):
        length = min(length, sum(p.size for p in self.pops))
        arr = self.L.opga_get_best_top_all(self.h, length)
        ptrs = (C.c_void_p * max(length, 1)).from_address(arr)
        rows = [self._take_row(ptrs[i], self.pops[0].genome_len) for i in range(length)]
        self.L.free(arr)
        return rows
