"""GPU parity of the output-mean accumulation of the resident tracers (fb200_mean_accumulate / fb200_mean_fetch;
io_meandata.F90:768-794, 894-900) against the oracle's stream restatement fed with the oracle's tracers.  Own file, sorted
after the other GPU tests: written after the round's GPU budget was spent, not yet run on a B200."""
import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("accuracy", [8, 4])
def test_device_means_match_stream_restatement(accuracy):
    from fesom2_b200 import api
    from oracle import binding as ob
    A, B = util.build_model("pi"), util.build_model("pi")
    n, nl = A.ipar("myDim_nod2D"), A.ipar("nl")
    o = ob.TracerOracle(A)
    eng = api.TracerEngine(B)
    eng.push_step_state()
    eng.push_tracers()
    out = np.zeros_like(B["tr_arr"])
    streams = [ob.MeanStream((n, nl - 1), accuracy) for _ in range(2)]
    for interval in range(2):                           # two averaging intervals: the accumulator must come back cleared
        for s in range(3):
            o.step()
            eng.solve_tracers_ale_resident(out)
            for k in range(2):
                streams[k].update(A["tr_arr"][k][:n])
                eng.mean_accumulate(k + 1, accuracy)
        for k in range(2):
            want = streams[k].output()
            got = eng.mean_fetch(k + 1, accuracy)
            assert got.dtype == want.dtype
            assert np.array_equal(got, want) or util.rel_err(got, want) < (1e-12 if accuracy == 8 else 1e-6), (interval, k)
    eng.close()
