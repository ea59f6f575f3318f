"""The CUDA engine against the COMMITTED golden digests (tests/golden/oracle_digests.json,
made by tests/golden/make_digests.py from the oracle): one small case per BASELINE config,
full trajectories, bit-exact genomes and scores, same best individual."""
import json
import os
import sys

import pytest

import libpga_b200 as pg

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import cases  # noqa: E402

with open(os.path.join(ROOT, "tests", "golden", "oracle_digests.json")) as f:
    GOLDEN = json.load(f)


@pytest.mark.parametrize("name", sorted(cases.CASES))
def test_cuda_reproduces_golden_digests(pga, name):
    spec = cases.CASES[name]
    g = pga(spec["seed"])
    got = cases.run_case(g, spec, lambda n, l, e: g.create_population(n, l, encoding=e))
    assert got == GOLDEN[name]
