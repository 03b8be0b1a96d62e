"""fb200_tracer_extrema_noalloc — the allocation-free variant of the extrema scan (static device scratch, no cudaMalloc) —
against fb200_tracer_extrema through the C ABI, on two calls in a row.  min/max are exact operations, so the two entry
points must agree bit for bit (the tolerance of the tracer tests, 1e-12, is asserted as well).  Collected last: the file
name sorts after every other test file."""
import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu
TOL = 1e-12


def test_noalloc_extrema_matches_allocating_entry_point_twice_in_a_row():
    from fesom2_b200 import api
    hm = util.build_model("band")
    eng = api.TracerEngine(hm)
    eng.push_step_state()
    eng.push_tracers()
    wet = util.wet_mask(hm)
    for call in range(2):                                   # second call: the shared scratch holds the first call's partials
        eng.solve()                                         # a new tracer state for every call
        tr = np.zeros_like(hm["tr_arr"])
        tro = np.zeros_like(hm["tr_arr_old"])
        eng.pull_tracers(tr, tro)
        for k in range(tr.shape[0]):
            ref = eng.extrema(k + 1)
            for nblocks in (0, 1, 7, 2048):
                got = eng.extrema_noalloc(k + 1, nblocks)
                assert got == ref, (call, k, nblocks, got, ref)
                assert abs(got[0] - ref[0]) <= TOL * abs(ref[0]) and abs(got[1] - ref[1]) <= TOL * abs(ref[1])
            assert ref == (tr[k][wet].min(), tr[k][wet].max())
    # sizes beyond the static scratch are refused with a clear error, and the library stays usable
    with pytest.raises(api.FesomB200Error, match="exceeds the static scratch"):
        eng.extrema_noalloc(1, 2049)
    assert eng.extrema_noalloc(1, 0) == eng.extrema(1)
    eng.close()
