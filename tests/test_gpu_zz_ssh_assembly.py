"""GPU parity of the device-side SSH assembly (fb200_update_stiff_mat, fb200_compute_ssh_rhs,
fb200_update_vel) against the host-side mirror of update_stiff_mat_ale / compute_ssh_rhs_ale and a numpy restatement of
update_vel, inside the coupled step order of oce_timestep_ale (oce_ale.F90:2694-2760) with the device-side ALE producers.
Own file, sorted after the other GPU tests: written after the round's GPU budget was spent, not yet run on a B200."""
import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu
TOL = 1e-12


@pytest.mark.parametrize("name,kw", [("pi", {"which_ale": "zstar"}), ("band", {"which_ale": "zstar", "water_flux": 2e-7}),
                                     ("band", {"which_ale": "zlevel", "water_flux": 2e-7}), ("soufflet", {})])
def test_device_ssh_assembly_matches_host_mirror(name, kw):
    from fesom2_b200 import api
    from oracle import binding as ob
    kw = dict({"umax": 0.05}, **kw)
    A, B = util.build_model(name, **kw), util.build_model(name, **kw)
    linfs = A.ipar("which_ale") == 0
    n, myE, nl = A.ipar("myDim_nod2D"), A.ipar("myDim_elem2D"), A.ipar("nl")
    o = ob.TracerOracle(A)
    eng = api.TracerEngine(B)
    eng.push_step_state()
    eng.push_tracers()
    eng.ale_init()
    eng.ssh_assembly_init()
    out = np.zeros_like(B["tr_arr"])
    wet = util.wet_mask(A)
    rng = np.random.default_rng(7)
    lon = A["geo_coord_nod2D"][:, 0]
    # step 0: the state fill_state produced
    o.step()
    eng.solve_tracers_ale_resident(out)
    A.update_thickness()
    eng.ale_step_end()
    for s in range(1, 4):
        if not linfs:                                           # oce_ale.F90:2686
            A.stiff_mat_update()
            eng.update_stiff_mat()
            assert np.array_equal(A["ssh_values"], B["ssh_values"]), s
        r = 1e-3 * rng.standard_normal(A["UV_rhs"].shape)      # the momentum step's increment (outside the path)
        A["UV_rhs"][:] = r
        B["UV_rhs"][:] = r
        A.compute_ssh_rhs()
        eng.compute_ssh_rhs()
        assert np.array_equal(A["ssh_rhs"][:n], B["ssh_rhs"][:n]), s
        d = 1e-2 * np.sin(3 * lon + s)                          # stands in for the SSH solution
        A["d_eta"][:] = d
        B["d_eta"][:] = d
        util.update_vel_host(A)
        eng.update_vel()
        assert np.array_equal(A["UV"], B["UV"]), s
        A.compute_hbar()
        A.vert_vel()
        eng.ale_step_begin(uv_resident=True)
        for nm in ("hbar", "ssh_rhs_old", "dhe"):
            assert np.array_equal(A[nm], B[nm]), (s, nm)
        assert np.array_equal(eng.debug("Wvel"), A["Wvel"]), s
        o.step()
        eng.solve_tracers_ale_resident(out)
        for k in range(2):
            assert util.rel_err(out[k][wet], A["tr_arr"][k][wet]) < TOL, (s, k)
        A.update_thickness()
        eng.ale_step_end()
    eng.close()
