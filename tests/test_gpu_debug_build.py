"""Parity cases through the asserting build of the library (`make debug`: -DPGA_DEBUG).

compute-sanitizer is closed on this GPU pool, so memory safety is checked from inside: the
debug build asserts, on the device, every index that addresses a population row, a score or
an output slot (tournament candidates, parent rows, child rows, mutation positions, the slots
of the top-k compaction, the rows packed and unpacked by a migration).  A failed assert traps
the kernel, the next CUDA call fails and the library exits with "GPUassert: ...", so a clean
run of these cases at ragged sizes is the evidence.  The cases run in a child process with
$PGA_B200_SO pointing at the debug library (the loader honours it), because a process can hold
only one build of the library.
"""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DBG = os.path.join(ROOT, "libpga_b200", "libpga_b200_dbg.so")

CASES = ("test_c1_trajectory_every_generation or test_onemax_trajectory or test_onemax_shapes or test_rastrigin_shapes "
         "or test_tsp_demo or test_crossover_hook_that_leaves_genes_unwritten or test_mutate_hook_reads_its_rand_row "
         "or test_islands_of_unequal_size or test_run_islands_matches_oracle or test_select_top_matches_oracle "
         "or test_tournament_of_k_matches_oracle or test_run_islands_stops_at_the_target")


def test_parity_cases_pass_with_device_asserts_enabled():
    if not os.path.exists(DBG):
        pytest.skip("libpga_b200_dbg.so not built (make debug)")
    env = dict(os.environ, PGA_B200_SO=DBG)
    cmd = [sys.executable, "-m", "pytest", "-x", "-q", "-m", "gpu", "-k", CASES,
           os.path.join(ROOT, "tests", "test_gpu_parity.py"), os.path.join(ROOT, "tests", "test_gpu_islands.py"),
           os.path.join(ROOT, "tests", "test_gpu_extensions.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=1500, cwd=ROOT)
    tail = r.stdout[-1500:] + r.stderr[-1500:]
    assert r.returncode == 0, tail
    assert " passed" in r.stdout and "failed" not in r.stdout, tail


def test_the_debug_build_does_assert():
    # the asserts are really compiled in: the debug library carries the device-side assert
    # machinery (__assertfail), the release library does not
    if not os.path.exists(DBG):
        pytest.skip("libpga_b200_dbg.so not built (make debug)")
    rel = os.path.join(ROOT, "libpga_b200", "libpga_b200.so")
    assert b"__assertfail" in open(DBG, "rb").read()
    assert b"__assertfail" not in open(rel, "rb").read()
