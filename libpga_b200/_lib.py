"""Loader for the C-ABI shared library (libpga_b200.so).

There is no Python or CPU fallback: if the CUDA extension has not been built the
import of anything that needs it fails loudly.  Build with `make` at the repo root
(or `python -c "import __graft_entry__ as g; g.build()"`).
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# $PGA_B200_SO selects another build of the same library (the asserting debug build, `make debug`)
SO_PATH = os.environ.get("PGA_B200_SO") or os.path.join(_HERE, "libpga_b200.so")

_vp, _u64, _u32, _i32, _f32 = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int, C.c_float
_ul, _ui, _sz = C.c_ulong, C.c_uint, C.c_size_t

# every symbol include/pga.h and include/pga_b200.h declare: name -> (restype, argtypes)
SIGNATURES = {
    # include/pga.h
    "pga_init": (_vp, []),
    "pga_deinit": (None, [_vp]),
    "pga_create_population": (_vp, [_vp, _ul, _ui, _i32]),
    "pga_set_objective_function": (None, [_vp, _vp]),
    "pga_set_mutate_function": (None, [_vp, _vp]),
    "pga_set_crossover_function": (None, [_vp, _vp]),
    "pga_get_best": (_vp, [_vp, _vp]),
    "pga_get_best_top": (_vp, [_vp, _vp, _ui]),
    "pga_get_best_all": (_vp, [_vp]),
    "pga_get_best_top_all": (_vp, [_vp, _ui]),
    "pga_evaluate": (None, [_vp, _vp]),
    "pga_evaluate_all": (None, [_vp]),
    "pga_crossover": (None, [_vp, _vp, _i32]),
    "pga_crossover_all": (None, [_vp, _i32]),
    "pga_migrate": (None, [_vp, _f32]),
    "pga_migrate_between": (None, [_vp, _vp, _vp, _f32]),
    "pga_mutate": (None, [_vp, _vp]),
    "pga_mutate_all": (None, [_vp]),
    "pga_swap_generations": (None, [_vp, _vp]),
    "pga_fill_random_values": (None, [_vp, _vp]),
    "pga_run": (None, [_vp, _ui]),
    "pga_run_islands": (None, [_vp, _ui, _ui, _f32]),
    # include/pga_b200.h
    "pgax_set_seed": (_i32, [_vp, _u64]),
    "pgax_get_seed": (_u64, [_vp]),
    "pgax_create_population": (_vp, [_vp, _ul, _ui, _i32]),
    "pgax_set_builtin_objective": (_i32, [_vp, _i32]),
    "pgax_set_quiet": (_i32, [_vp, _i32]),
    "pgax_set_selection": (_i32, [_vp, _i32]),
    "pgax_set_target": (_i32, [_vp, _f32, _ui]),
    "pgax_generations_run": (_ui, [_vp]),
    "pgax_save": (_i32, [_vp, C.c_char_p]),
    "pgax_restore": (_i32, [_vp, C.c_char_p]),
    "pgax_set_stream": (_i32, [_vp, _vp]),
    "pgax_capture_selection": (_i32, [_vp, _vp, _i32]),
    "pgax_population": (_vp, [_vp, _ui]),
    "pgax_encoding": (_i32, [_vp]),
    "pgax_row_words": (_ui, [_vp]),
    "pgax_population_size": (_ul, [_vp]),
    "pgax_genome_len": (_ui, [_vp]),
    "pgax_epoch": (_u32, [_vp]),
    "pgax_set_epoch": (_i32, [_vp, _u32]),
    "pgax_island_id": (_u32, [_vp]),
    "pgax_upload_genomes": (_i32, [_vp, _vp, _vp, _sz]),
    "pgax_download_genomes": (_i32, [_vp, _vp, _vp, _sz]),
    "pgax_download_next_genomes": (_i32, [_vp, _vp, _vp, _sz]),
    "pgax_upload_scores": (_i32, [_vp, _vp, _vp, _sz]),
    "pgax_download_scores": (_i32, [_vp, _vp, _vp, _sz]),
    "pgax_download_selection": (_i32, [_vp, _vp, _vp, _sz]),
    "pgax_best": (_i32, [_vp, _vp, _vp, _vp]),
    "pgax_get_best_words": (_i32, [_vp, _vp, _vp]),
    "pgax_select_top": (_i32, [_vp, _vp, _ul, _i32, _vp]),
    "pgax_select_top_noalloc": (_i32, [_vp, _ul, _ul, _i32, _vp, _vp]),
    "pgax_set_topology": (_i32, [_vp, _i32]),
    "pgax_set_migration_delay": (_i32, [_vp, _ui]),
    "pgax_migration_count": (_ul, [_ul, _f32]),
    "pgax_dist_unique_id": (_i32, [_vp]),
    "pgax_dist_init": (_i32, [_vp, _i32, _i32, _vp]),
    "pgax_dist_set_layout": (_i32, [_vp, _i32, _i32]),
    "pgax_dist_rank": (_i32, [_vp]),
    "pgax_dist_world": (_i32, [_vp]),
    "pgax_plan_exchange": (_i32, [_u64, _u32, _i32, _i32, _ui, _vp, _ui]),
    "pgax_last_run_ms": (_f32, [_vp]),
    "pgax_kernel_launches": (C.c_ulonglong, [_vp]),
    "pgax_bytes_to_host": (C.c_ulonglong, [_vp]),
    "pgax_demo_hook": (_vp, [C.c_char_p]),
    "pgax_demo_tsp_load": (_i32, [_vp, _i32]),
    "pgax_version": (C.c_char_p, []),
}

_lib = None


def load():
    """dlopen libpga_b200.so and type every entry point; raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise RuntimeError(
            f"libpga_b200: CUDA extension not built ({SO_PATH} missing). "
            "Run `make` in the repository root. There is no CPU fallback.")
    lib = C.CDLL(SO_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError = the library lacks a declared symbol
        fn.restype, fn.argtypes = res, args
    libc = C.CDLL(None)
    libc.free.argtypes = [_vp]
    lib._free = libc.free
    _lib = lib
    return lib
