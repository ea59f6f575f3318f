"""Python mirror of libpga's C API (include/pga.h) over the C-ABI shared library.

Method names, argument order and behaviour follow the pga_* functions
(reference: include/pga.h:53-150), so a test written against the reference reads
the same here.  All compute happens in the CUDA library; this module only
marshals pointers.  numpy arrays are host buffers: every upload/download is a
real H2D/D2H copy inside the call.
"""
import ctypes as C

import numpy as np

from . import _lib

RANDOM_POPULATION = 0   # enum population_type (include/pga.h:31-34)
TOURNAMENT = 0          # enum crossover_selection_type (include/pga.h:39-42)
MAX_POPULATIONS = 10    # include/pga.h:44



def tournament_of(k):
    """PGAX_TOURNAMENT_OF(k): tournament among k candidates as a crossover_selection_type."""
    return 0x100 + int(k)


GENES_F32, GENES_BITS = 0, 1
OBJ_HOOK, OBJ_SUM, OBJ_RASTRIGIN, OBJ_SPHERE, OBJ_ONEMAX = 0, 1, 2, 3, 4
TOPO_RING, TOPO_RANDOM = 0, 1


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class Population:
    """Borrowed handle to one population (population_t*); owned by its PGA."""

    def __init__(self, owner, handle, encoding):
        self.owner, self.h, self.encoding = owner, handle, encoding
        L = owner.L
        self.size = L.pgax_population_size(handle)
        self.genome_len = L.pgax_genome_len(handle)
        self.row_words = L.pgax_row_words(handle)

    @property
    def dtype(self):
        return np.uint32 if self.encoding == GENES_BITS else np.float32

    @property
    def epoch(self):
        return self.owner.L.pgax_epoch(self.h)

    @property
    def island(self):
        return self.owner.L.pgax_island_id(self.h)

    def set_epoch(self, e):
        self.owner.L.pgax_set_epoch(self.h, e)

    def genes(self, out=None):
        a = out if out is not None else np.empty((self.size, self.row_words), dtype=self.dtype)
        rc = self.owner.L.pgax_download_genomes(self.owner.h, self.h, _ptr(a), a.nbytes)
        assert rc == 0, rc
        return a

    def next_genes(self):
        a = np.empty((self.size, self.row_words), dtype=self.dtype)
        rc = self.owner.L.pgax_download_next_genomes(self.owner.h, self.h, _ptr(a), a.nbytes)
        assert rc == 0, rc
        return a

    def set_genes(self, rows):
        a = np.ascontiguousarray(rows, dtype=self.dtype)
        rc = self.owner.L.pgax_upload_genomes(self.owner.h, self.h, _ptr(a), a.nbytes)
        assert rc == 0, rc

    def scores(self):
        a = np.empty(self.size, dtype=np.float32)
        rc = self.owner.L.pgax_download_scores(self.owner.h, self.h, _ptr(a), a.size)
        assert rc == 0, rc
        return a

    def set_scores(self, s):
        a = np.ascontiguousarray(s, dtype=np.float32)
        rc = self.owner.L.pgax_upload_scores(self.owner.h, self.h, _ptr(a), a.size)
        assert rc == 0, rc

    def capture_selection(self, enable=True):
        self.owner.L.pgax_capture_selection(self.owner.h, self.h, int(enable))

    def selection(self):
        a = np.empty((self.size, 2), dtype=np.int32)
        rc = self.owner.L.pgax_download_selection(self.owner.h, self.h, _ptr(a), self.size)
        assert rc == 0, rc
        return a


class PGA:
    """pga_t: one solver on the current CUDA device."""

    def __init__(self, seed=None, quiet=True):
        self.L = _lib.load()
        self.h = self.L.pga_init()
        if not self.h:
            raise MemoryError("pga_init failed")
        if seed is not None:
            self.L.pgax_set_seed(self.h, seed)
        self.L.pgax_set_quiet(self.h, int(quiet))
        self.pops = []

    # -- lifetime
    def deinit(self):
        if self.h:
            self.L.pga_deinit(self.h)
            self.h = None

    close = deinit

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.deinit()

    # -- set-up
    @property
    def seed(self):
        return self.L.pgax_get_seed(self.h)

    def create_population(self, size, genome_len, type=RANDOM_POPULATION, encoding=GENES_F32):
        if encoding == GENES_F32:
            h = self.L.pga_create_population(self.h, size, genome_len, type)
        else:
            h = self.L.pgax_create_population(self.h, size, genome_len, encoding)
        if not h:
            return None
        p = Population(self, h, encoding)
        self.pops.append(p)
        return p

    def demo_hook(self, name):
        return self.L.pgax_demo_hook(name.encode())

    def _hook(self, f):
        if f is None:
            return None
        return self.demo_hook(f) if isinstance(f, str) else f

    def set_objective_function(self, f):
        self.L.pga_set_objective_function(self.h, self._hook(f))

    def set_mutate_function(self, f):
        self.L.pga_set_mutate_function(self.h, self._hook(f))

    def set_crossover_function(self, f):
        self.L.pga_set_crossover_function(self.h, self._hook(f))

    def set_builtin_objective(self, which):
        rc = self.L.pgax_set_builtin_objective(self.h, which)
        assert rc == 0, rc

    def set_selection(self, type):
        rc = self.L.pgax_set_selection(self.h, type)
        assert rc == 0, rc

    def set_target(self, value, check_every=1):
        self.L.pgax_set_target(self.h, value, check_every)

    @property
    def generations_run(self):
        return int(self.L.pgax_generations_run(self.h))

    def save(self, path):
        rc = self.L.pgax_save(self.h, str(path).encode())
        if rc != 0:
            raise OSError(f"pgax_save failed: {rc}")

    def restore(self, path):
        rc = self.L.pgax_restore(self.h, str(path).encode())
        if rc != 0:
            raise OSError(f"pgax_restore failed: {rc}")
        if not self.pops:  # populations created by the restore: wrap them
            i = 0
            while True:
                h = self.L.pgax_population(self.h, i)
                if not h:
                    break
                enc = self.L.pgax_encoding(h)
                self.pops.append(Population(self, h, enc))
                i += 1
        return self.pops

    def set_stream(self, cuda_stream):
        self.L.pgax_set_stream(self.h, cuda_stream)

    def tsp_load(self, matrix):
        m = np.ascontiguousarray(matrix, dtype=np.float32)
        n = int(round(m.size ** 0.5))
        rc = self.L.pgax_demo_tsp_load(_ptr(m), n)
        assert rc == 0, rc

    # -- steps (include/pga.h:95-134)
    def evaluate(self, pop): self.L.pga_evaluate(self.h, pop.h)
    def evaluate_all(self): self.L.pga_evaluate_all(self.h)
    def crossover(self, pop, type=TOURNAMENT): self.L.pga_crossover(self.h, pop.h, type)
    def crossover_all(self, type=TOURNAMENT): self.L.pga_crossover_all(self.h, type)
    def mutate(self, pop): self.L.pga_mutate(self.h, pop.h)
    def mutate_all(self): self.L.pga_mutate_all(self.h)
    def swap_generations(self, pop): self.L.pga_swap_generations(self.h, pop.h)
    def fill_random_values(self, pop): self.L.pga_fill_random_values(self.h, pop.h)
    def run(self, n): self.L.pga_run(self.h, n)
    def run_islands(self, n, m, pct): self.L.pga_run_islands(self.h, n, m, pct)
    def migrate(self, pct): self.L.pga_migrate(self.h, pct)
    def migrate_between(self, a, b, pct): self.L.pga_migrate_between(self.h, a.h, b.h, pct)
    def set_topology(self, t): self.L.pgax_set_topology(self.h, t)
    def set_migration_delay(self, d): self.L.pgax_set_migration_delay(self.h, d)

    # -- results
    def _take_row(self, addr, length):
        row = np.frombuffer((C.c_float * length).from_address(addr), dtype=np.float32).copy()
        self.L._free(addr)
        return row

    def get_best(self, pop):
        addr = self.L.pga_get_best(self.h, pop.h)
        return None if not addr else self._take_row(addr, pop.genome_len)

    def _take_rows(self, arr, length, genome_len):
        if not arr:
            return None
        ptrs = (C.c_void_p * max(length, 1)).from_address(arr)
        rows = [self._take_row(ptrs[i], genome_len) for i in range(length) if ptrs[i]]
        self.L._free(arr)
        return rows

    def get_best_top(self, pop, length):
        return self._take_rows(self.L.pga_get_best_top(self.h, pop.h, length), length, pop.genome_len)

    def get_best_all(self):
        addr = self.L.pga_get_best_all(self.h)
        return None if not addr else self._take_row(addr, self.pops[0].genome_len)

    def get_best_top_all(self, length):
        return self._take_rows(self.L.pga_get_best_top_all(self.h, length), length, self.pops[0].genome_len)

    def best(self, pop):
        idx, s = C.c_ulong(0), C.c_float(0)
        rc = self.L.pgax_best(self.h, pop.h, C.byref(idx), C.byref(s))
        assert rc == 0, rc
        return int(idx.value), float(s.value)

    def best_words(self, pop):
        a = np.empty(pop.row_words, dtype=np.uint32)
        rc = self.L.pgax_get_best_words(self.h, pop.h, _ptr(a))
        assert rc == 0, rc
        return a

    def select_top(self, pop, k, worst=False):
        k = min(int(k), pop.size)
        a = np.empty(max(k, 1), dtype=np.uint32)
        rc = self.L.pgax_select_top(self.h, pop.h, k, int(bool(worst)), _ptr(a))
        assert rc == 0, rc
        return a[:k]

    # -- multi-process islands
    def dist_set_layout(self, rank, world):
        rc = self.L.pgax_dist_set_layout(self.h, rank, world)
        assert rc == 0, rc

    def dist_init(self, rank, world, unique_id_bytes):
        buf = (C.c_ubyte * 128).from_buffer_copy(bytes(unique_id_bytes))
        rc = self.L.pgax_dist_init(self.h, rank, world, buf)
        if rc != 0:
            raise RuntimeError(f"pgax_dist_init failed: {rc}")

    # -- timing
    @property
    def last_run_ms(self):
        return float(self.L.pgax_last_run_ms(self.h))

    @property
    def kernel_launches(self):
        return int(self.L.pgax_kernel_launches(self.h))

    @property
    def bytes_to_host(self):
        return int(self.L.pgax_bytes_to_host(self.h))


def dist_unique_id():
    buf = (C.c_ubyte * 128)()
    rc = _lib.load().pgax_dist_unique_id(buf)
    if rc != 0:
        raise RuntimeError(f"pgax_dist_unique_id failed: {rc} (NCCL not loadable?)")
    return bytes(buf)


def migration_count(size, pct):
    return int(_lib.load().pgax_migration_count(size, pct))
