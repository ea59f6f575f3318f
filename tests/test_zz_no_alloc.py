"""pgax_select_top_noalloc: ranking for callers that must not let the library
allocate -- programs whose GPU memory is wholly owned by their own pool, and loops
that cannot absorb the device-wide stall of cudaMalloc/cudaFree.  Its working
memory is a fixed array inside the library, so the interesting failure modes are
stale contents from an earlier call and sizes beyond the array.

The file name sorts last on purpose: with `pytest -x` every other test has already
run when this one starts.  It holds no state of its own, so it also passes alone in
a fresh process.
"""
import ctypes as C

import numpy as np
import pytest

import libpga_b200 as pg
from libpga_b200 import _lib

pytestmark = pytest.mark.gpu

MAX_N = 1 << 24  # PGAX_NOALLOC_MAX_INDIVIDUALS


def existing_entry_point(pga, scores, k, worst):
    """pgax_select_top, the allocating entry point, on a population holding `scores`."""
    g = pga(1)
    p = g.create_population(scores.size, 8)
    g.set_objective_function("sum_obj")
    p.set_scores(scores)
    return g.select_top(p, k, worst=worst)


def noalloc_indices(scores, k, worst, check_no_allocation=False):
    import torch
    L = _lib.load()
    dev_scores = torch.from_numpy(scores).cuda()
    out = torch.full((max(min(k, scores.size), 1),), -1, dtype=torch.int32, device="cuda")
    stream = torch.cuda.current_stream()
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]  # cudaMemGetInfo
    rc = L.pgax_select_top_noalloc(C.c_void_p(dev_scores.data_ptr()), scores.size, k, int(worst),
                                   C.c_void_p(out.data_ptr()), C.c_void_p(stream.cuda_stream))
    assert rc == 0, rc
    stream.synchronize()
    if check_no_allocation:
        assert torch.cuda.mem_get_info()[0] == free_before, "device memory was allocated or freed during the call"
    return out.cpu().numpy().view(np.uint32)[:min(k, scores.size)]


def test_two_calls_in_a_row_match_the_oracle_and_select_top(pga, oracle):
    # call 1: many ties, best-k; call 2: a different size, distinct values with NaNs, worst-k.
    # The second call runs on whatever the first left in the shared array.  Both are checked
    # against the CPU oracle's ranking and against the existing (allocating) entry point.
    rng = np.random.default_rng(7)
    first = rng.integers(500, 520, size=200_000).astype(np.float32)
    second = ((rng.random(70_001, dtype=np.float32) - 0.5) * 100).astype(np.float32)
    second[::97] = np.nan
    got1 = noalloc_indices(first, 2000, worst=False)
    got2 = noalloc_indices(second, 701, worst=True, check_no_allocation=True)
    assert np.array_equal(got1, oracle.select_top(first, 2000, worst=False))   # integer indices: exact
    assert np.array_equal(got2, oracle.select_top(second, 701, worst=True))
    assert np.array_equal(got1, existing_entry_point(pga, first, 2000, False))
    assert np.array_equal(got2, existing_entry_point(pga, second, 701, True))
    # and once more with the first input: a larger problem after a smaller one
    assert np.array_equal(noalloc_indices(first, 2000, worst=False, check_no_allocation=True), got1)


def test_largest_accepted_size(oracle):
    # n = 2^24 is the limit: 8192 per-block counts, i.e. all eight trips of the scan's
    # 1024-wide loop; OneMax-like scores so that the k-th key has thousands of ties
    rng = np.random.default_rng(24)
    scores = rng.integers(480, 545, size=MAX_N).astype(np.float32)
    k = 167_772  # top 1 %
    for worst in (False, True):
        got = noalloc_indices(scores, k, worst=worst, check_no_allocation=True)
        assert np.array_equal(got, oracle.select_top(scores, k, worst=worst))


def test_sizes_beyond_the_fixed_array_are_refused(capfd):
    import torch
    L = _lib.load()
    small = torch.zeros(16, dtype=torch.float32, device="cuda")
    out = torch.zeros(16, dtype=torch.int32, device="cuda")
    rc = L.pgax_select_top_noalloc(C.c_void_p(small.data_ptr()), MAX_N + 1, 4, 0, C.c_void_p(out.data_ptr()), None)
    assert rc == -2  # PGAX_ERANGE
    assert "exceed" in capfd.readouterr().err
    assert L.pgax_select_top_noalloc(None, 16, 4, 0, C.c_void_p(out.data_ptr()), None) == -1  # PGAX_EINVAL
