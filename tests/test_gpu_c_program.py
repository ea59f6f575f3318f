"""A plain C program (gcc, no CUDA in the caller) against include/pga.h + pga_b200.h and
libpga_b200.so: the "host side is plain C over a thin C-ABI" claim, exercised end to end."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_c_example_builds_and_matches_oracle(tmp_path, oracle):
    exe = str(tmp_path / "onemax")
    lib = os.path.join(ROOT, "libpga_b200")
    cc = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else "gcc"
    subprocess.check_call([cc, "-O2", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "examples", "onemax.c"),
                           "-o", exe, "-L" + lib, "-lpga_b200", "-Wl,-rpath," + lib])
    r = subprocess.run([exe, "5000", "333", "12", "77"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    o = oracle.Oracle(77, threads=4)
    p = o.create_population(5000, 333, oracle.GENES_BITS)
    o.set_builtin_objective(oracle.OBJ_ONEMAX)
    o.run(12)
    idx, best = o.best_index(p)
    o.close()
    lines = r.stdout.strip().split("\n")
    assert lines[0] == "%f" % best                      # pga_get_best's print, src/pga.cu:230
    assert lines[1].startswith(f"best individual {idx}: {int(best)} of 333 bits set")
