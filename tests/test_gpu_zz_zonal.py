"""GPU parity of relax_zonal_temp (toy_channel_soufflet.F90:78-103; fb200_set_zonal_relax, k_relax_zonal) against the
oracle on the soufflet channel, host-tracer and resident stepping.  Own file, sorted after the other GPU tests: written
after the round's GPU budget was spent, not yet run on a B200."""
import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu
TOL = 1e-12


@pytest.mark.parametrize("resident", [False, True])
def test_relax_zonal_temp_matches_oracle(resident):
    from fesom2_b200 import api
    from oracle import binding as ob
    A, B = util.build_model("soufflet"), util.build_model("soufflet")
    ztem, Tclim, lat0, dy, tau_inv = util.soufflet_zonal_setup(A)
    Tclim = Tclim + 0.3
    o = ob.TracerOracle(A)
    o.set_zonal_relax(ztem, np.ascontiguousarray(A["coord_nod2D"]), Tclim, lat0, dy, tau_inv)
    eng = api.TracerEngine(B)
    eng.push_step_state()
    eng.set_zonal_relax(ztem, Tclim, lat0, dy, tau_inv)
    g_tr, g_tro = B["tr_arr"].copy(), B["tr_arr_old"].copy()
    if resident:
        eng.push_tracers(g_tr, g_tro)
    for s in range(3):
        o.step()
        if resident:
            eng.solve_tracers_ale_resident(g_tr)
        else:
            eng.solve_tracers_ale(g_tr, g_tro)
        for k in range(2):
            assert util.rel_err(g_tr[k], A["tr_arr"][k]) < TOL, (s, k)
    # and the restoring really acted: without it the temperature differs
    C = util.build_model("soufflet")
    oc = ob.TracerOracle(C)
    for s in range(3):
        oc.step()
    assert np.abs(C["tr_arr"][0] - A["tr_arr"][0]).max() > 1e-4
    eng.close()
