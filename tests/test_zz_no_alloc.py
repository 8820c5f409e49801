"""The allocation-free entry point (no counterpart in the reference):
smcphys_resample_systematic_noalloc - systematic resampling on device buffers with all of its scratch in one static
device array of the library - against the existing entry smcphys_resample_systematic, through the C ABI.

The bar is the one the resampling tests use everywhere: indices bit exact.  Two calls in a row share the scratch array
(it keeps the last call's contents), so the second call uses other inputs and a smaller n.  Collected last on purpose;
passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def sp():
    import __graft_entry__ as entry
    entry.build()
    import smc_physics_b200 as m
    return m


def _noalloc(sp, torch, lw, u0):
    d_lw = torch.from_numpy(np.ascontiguousarray(lw, dtype=np.float64)).cuda()
    d_out = torch.full((lw.shape[0],), -1, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream().cuda_stream
    rc = sp.lib().smcphys_resample_systematic_noalloc(0, C.c_void_p(d_lw.data_ptr()), lw.shape[0], float(u0),
                                                      C.c_void_p(d_out.data_ptr()), C.c_void_p(stream))
    assert rc == 0, sp.lib().smcphys_last_error()
    torch.cuda.synchronize()
    return d_out.cpu().numpy()


def test_noalloc_resampling_matches_existing_entry_twice_in_a_row(sp):
    import torch
    rng = np.random.default_rng(2024)
    # first call: the largest size the static scratch takes; second call right after it: other weights, smaller n
    # (stale CDF / tile sums of the first call lie beyond it); third: tiny
    for n, spread, u0 in ((1 << 20, 3.0, 0.37), (100003, 30.0, 1.0 - 2.0 ** -32), (33, 0.1, 0.0)):
        lw = rng.normal(0, spread, n) - 1234.5
        got = _noalloc(sp, torch, lw, u0)
        want = sp.resample_systematic(lw, u0)          # the existing entry point (allocates per call)
        assert np.array_equal(got, want), (n, np.flatnonzero(got != want)[:5])
    # the same input again gives the same answer (nothing left over from the calls in between)
    assert np.array_equal(_noalloc(sp, torch, lw, 0.0), want)


def test_noalloc_rejects_sizes_beyond_the_static_scratch(sp):
    import torch
    L = sp.lib()
    n = (1 << 20) + 1
    d = torch.zeros(8, dtype=torch.float64, device="cuda")
    rc = L.smcphys_resample_systematic_noalloc(0, C.c_void_p(d.data_ptr()), n, 0.5, C.c_void_p(d.data_ptr()), None)
    assert rc == -1 and b"SMCPHYS_NOALLOC_MAX_N" in L.smcphys_last_error()
    assert L.smcphys_resample_systematic_noalloc(0, None, 0, 0.5, None, None) == 0          # empty input: nothing to do
    assert L.smcphys_resample_systematic_noalloc(0, None, 5, 0.5, None, None) == -1         # NULL buffers
