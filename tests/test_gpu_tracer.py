"""GPU parity tests of the tracer path: CUDA kernels (through the C ABI) vs the oracle on the same inputs.
Tolerance: BASELINE.json north_star asks for 1e-12 relative per field per step; the kernels follow the
reference's operation order with FMA contraction off, so most fields are expected to agree bit for bit —
the assertion is the stated 1e-12 (max|diff| / max|field|), bit-equality is reported by test_bitwise_report."""
import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu
TOL = 1e-12


def run_pair(name, opts=None, model_kw=None, steps=1):
    from fesom2_b200 import api
    from oracle import binding as ob
    opts = opts or {}
    hm = util.build_model(name, **(model_kw or {}))
    tr0, tro0 = hm["tr_arr"].copy(), hm["tr_arr_old"].copy()
    eng = api.TracerEngine(hm, **opts)
    eng.push_step_state()
    g_tr, g_tro = tr0.copy(), tro0.copy()
    for _ in range(steps):
        eng.solve_tracers_ale(g_tr, g_tro)
    o = ob.TracerOracle(hm, **opts)
    for _ in range(steps):
        o.step()
    return hm, eng, o, g_tr, g_tro


@pytest.mark.parametrize("name,model_kw,opts", [
    ("pi", {}, {}),
    ("soufflet", {}, {}),
    ("band", {}, {}),
    ("pi", {"w_split": True, "w_max_cfl": 0.2}, {}),
    ("pi", {}, {"tra_adv_hor": "MUSCL"}),
    ("pi", {}, {"tra_adv_hor": "UPW1", "tra_adv_ver": "UPW1"}),
    ("band", {}, {"tra_adv_ver": "CDIFF"}),
    ("band", {}, {"tra_adv_ver": "PPM"}),
    ("pi", {}, {"tra_adv_ver": "PPM"}),
    ("soufflet", {"w_split": True, "w_max_cfl": 0.2}, {"tra_adv_ver": "PPM"}),
    ("pi", {}, {"i_vert_diff": False}),
    ("band", {"water_flux": 1e-7, "heat_flux": 50.0}, {"i_vert_diff": False, "tra_adv_lim": "NONE", "tra_adv_hor": "MUSCL"}),
    ("pi", {}, {"tra_adv_lim": "NONE", "tra_adv_hor": "MUSCL"}),
    ("band", {}, {"tra_adv_lim": "NONE", "tra_adv_hor": "MUSCL", "tra_adv_ver": "PPM"}),
    ("pi", {"w_split": True, "w_max_cfl": 0.2}, {"tra_adv_lim": "NONE", "tra_adv_hor": "UPW1", "tra_adv_ver": "CDIFF"}),
    ("soufflet", {"w_split": True, "w_max_cfl": 0.2}, {"tra_adv_lim": "NONE", "tra_adv_hor": "MUSCL"}),
    ("band", {"water_flux": 1e-7, "heat_flux": 50.0}, {}),
    ("band", {"num_tracers": 3}, {"tracer_ids": (0, 1, 301)}),
    ("farclike", {}, {}),                         # BASELINE.json config 4 stand-in: multi-resolution mesh
    ("farclike", {"shuffle": True}, {}),          # the same mesh with a random node/element numbering (no locality)
    ("core2like", {}, {}),                        # BASELINE.json config 3 stand-in: 127 k surface nodes x 47 layers
    ("band", {"nlon": 60, "nlat": 36, "nl": 63}, {}),                                   # up to 59 wet layers: k_fct_bounds<4>, three level chunks
    ("band", {"nlon": 60, "nlat": 36, "nl": 56, "w_split": True, "w_max_cfl": 0.2}, {"tra_adv_ver": "PPM"}),
])
def test_tracer_step_matches_oracle(name, model_kw, opts):
    hm, eng, o, g_tr, g_tro = run_pair(name, opts, model_kw)
    wet = util.wet_mask(hm)
    c_tr, c_tro = hm["tr_arr"], hm["tr_arr_old"]
    for k in range(c_tr.shape[0]):
        assert util.rel_err(g_tr[k][wet], c_tr[k][wet]) < TOL, "tracer %d" % k
        assert util.rel_err(g_tro[k], c_tro[k]) < TOL


_fill_redi_sw = util.fill_redi_sw


@pytest.mark.parametrize("name,opts", [
    ("band", {"Redi": True}),
    ("pi", {"Redi": True, "use_sw_pene": True}),
    ("soufflet", {"use_sw_pene": True, "tra_adv_lim": "NONE", "tra_adv_hor": "MUSCL"}),
    ("cavity", {"Redi": True, "fct_cavity_neutral": True}),
    ("pi", {"Redi": True, "i_vert_diff": False}),                   # diff_ver_part_expl_ale and diff_ver_part_redi_expl in one step
    ("pi", {"Redi": True, "do_hor_diff": False}),                   # the tr_xy(T) the term needs is still computed
    ("pi", {"smooth_bh_tra": True}),                                # diff_part_bh with the namelist's gamma0_tra/gamma1_tra
    ("band", {"smooth_bh_tra": True, "gamma_tra": (0.0002, 0.05, 3.0), "Redi": True}),
    ("cavity", {"smooth_bh_tra": True, "fct_cavity_neutral": True}),
])
def test_redi_and_shortwave_match_oracle(name, opts):
    from fesom2_b200 import api
    from oracle import binding as ob
    hm = util.build_model(name)
    _fill_redi_sw(hm)
    g_tr, g_tro = hm["tr_arr"].copy(), hm["tr_arr_old"].copy()
    eng = api.TracerEngine(hm, **{k: v for k, v in opts.items() if k != "fct_cavity_neutral"})
    eng.push_step_state()
    eng.solve_tracers_ale(g_tr, g_tro)
    ob.TracerOracle(hm, **opts).step()
    wet = util.wet_mask(hm)
    for k in range(g_tr.shape[0]):
        assert util.rel_err(g_tr[k][wet], hm["tr_arr"][k][wet]) < TOL, "tracer %d" % k
        assert np.array_equal(g_tr[k][wet], hm["tr_arr"][k][wet]), "tracer %d not bit-identical" % k
    eng.close()


def test_intermediate_fields_match_oracle():
    """last tracer's work arrays: fct_LO, antidiffusive fluxes (before limiting on the GPU side they are
    unlimited; compare the limiter factors and LO instead), fct_plus/minus, gradients"""
    hm, eng, o, g_tr, g_tro = run_pair("pi")
    nl, N, Ed = hm.ipar("nl"), hm.ipar("myDim_nod2D"), hm.ipar("myDim_edge2D")
    wet = util.wet_mask(hm)
    for nm in ("fct_LO", "fct_plus", "fct_minus"):
        ref = o.work(nm, (nl - 1, N))
        got = eng.debug(nm)
        assert util.rel_err(got[wet], ref[wet]) < TOL, nm


def test_five_steps_stay_within_tolerance():
    hm, eng, o, g_tr, g_tro = run_pair("band", steps=5)
    wet = util.wet_mask(hm)
    for k in range(2):
        assert util.rel_err(g_tr[k][wet], hm["tr_arr"][k][wet]) < 1e-11


def test_bitwise_report(capsys):
    hm, eng, o, g_tr, g_tro = run_pair("pi")
    wet = util.wet_mask(hm)
    same = [(g_tr[k][wet] == hm["tr_arr"][k][wet]).mean() for k in range(2)]
    with capsys.disabled():
        print("\n[bitwise] fraction of identical values after one step: T %.6f  S %.6f" % tuple(same))
    assert min(same) > 0.5


def test_device_resident_steps_equal_push_pull_steps():
    from fesom2_b200 import api
    hm = util.build_model("band")
    eng = api.TracerEngine(hm)
    eng.push_step_state()
    a_tr, a_tro = hm["tr_arr"].copy(), hm["tr_arr_old"].copy()
    b_tr, b_tro = a_tr.copy(), a_tro.copy()
    for _ in range(3):
        eng.solve_tracers_ale(a_tr, a_tro)
    eng.push_tracers(b_tr, b_tro)
    for _ in range(3):
        eng.solve()
    eng.pull_tracers(b_tr, b_tro)
    assert np.array_equal(a_tr, b_tr) and np.array_equal(a_tro, b_tro)


def test_split_state_push_equals_single_push():
    """fb200_push_state_early (asynchronous) + psolve + fb200_push_state_late give the same step as fb200_push_step_state"""
    from fesom2_b200 import api
    hm = util.build_model("band")
    hm.compute_ssh_rhs()
    a_tr, a_tro = hm["tr_arr"].copy(), hm["tr_arr_old"].copy()
    b_tr, b_tro = a_tr.copy(), a_tro.copy()
    eng = api.TracerEngine(hm)
    eng.push_step_state()
    eng.solve_tracers_ale(a_tr, a_tro)
    solver = api.SSHSolver(hm)
    sol = np.zeros_like(hm["d_eta"])
    eng.push_state_early()
    solver.solve(new_values=1, sol=sol)
    eng.push_state_late()
    eng.solve_tracers_ale(b_tr, b_tro)
    assert np.array_equal(a_tr, b_tr) and np.array_equal(a_tro, b_tro)
    assert solver.info()[1] < 1e-10
    eng.close()


def test_div_by_matches_ieee_division():
    """the kernels divide several fluxes by the same areasvol through one reciprocal + an exact-residual correction;
    it must round exactly like '/' (bit-exact parity with the reference's divisions depends on it)"""
    from fesom2_b200 import api
    assert api.selftest_div(1 << 24) == 0


@pytest.mark.parametrize("name", ["pi", "band"])
def test_resident_time_stepping_matches_oracle(name):
    """fb200_solve_tracers_ale_resident: three steps on device-resident tracers (only tr_arr comes back) with the early
    state pushed through the NULL conventions (Ki kept resident; hnode kept = the host's unchanged hnode) == three oracle
    steps, and == the host-array path step by step."""
    from fesom2_b200 import api
    from oracle import binding as ob
    hm = util.build_model(name)
    eng = api.TracerEngine(hm)
    eng.push_step_state()
    eng.push_tracers()
    out = np.zeros_like(hm["tr_arr"])
    o = ob.TracerOracle(hm)
    wet = util.wet_mask(hm)
    for step in range(3):
        eng.push_state_early(ki_resident=True)
        eng.push_state_late()
        eng.solve_tracers_ale_resident(out)
        o.step()
        for k in range(out.shape[0]):
            assert util.rel_err(out[k][wet], hm["tr_arr"][k][wet]) < TOL, "step %d tracer %d" % (step, k)
    # hnode = NULL: the old thickness becomes the resident hnode_new.  Make the host agree (hnode := hnode_new, what
    # update_thickness_ale does) and compare one more step
    hm["hnode"][:] = hm["hnode_new"]
    eng.push_state_early(hnode_resident=True, ki_resident=True)
    eng.push_state_late()
    eng.solve_tracers_ale_resident(out)
    ob.TracerOracle(hm).step()
    for k in range(out.shape[0]):
        assert util.rel_err(out[k][wet], hm["tr_arr"][k][wet]) < TOL


@pytest.mark.parametrize("name", ["pi", "band"])
def test_100_coupled_steps_match_oracle_mass_and_extrema(name):
    """BASELINE.json north_star: fields to 1e-12 per step and 'stated bounds on global tracer mass and extrema after 100
    steps'.  The same 100-step time loop (host: compute_hbar_ale -> vert_vel_ale, update_thickness_ale; device:
    resident tracer stepping) on the GPU and in the oracle: final fields agree to 1e-10 (100 x the per-step tolerance;
    measured: bit-identical), global mass to 1e-13 relative, extrema to 1e-12, and both stay inside the initial
    [min, max] with mass conserved."""
    from fesom2_b200 import api
    from oracle import binding as ob
    kw = dict(which_ale="zstar", umax=0.02)
    ho = util.build_model(name, **kw)                       # oracle's copy of the model
    o = ob.TracerOracle(ho)
    hg = util.build_model(name, **kw)                       # the GPU run's copy
    wet = util.wet_mask(hg)
    m0 = [util.tracer_mass(hg, k) for k in range(2)]
    ext0 = [(hg["tr_arr"][k][wet].min(), hg["tr_arr"][k][wet].max()) for k in range(2)]
    eng = api.TracerEngine(hg)
    eng.push_tracers()
    out = hg["tr_arr"]                                      # the device's tr_arr lands in the host model's array
    for s in range(100):
        for hm in (ho, hg):
            util.advance_ale_state(hm, first=(s == 0))
        o.step()
        eng.push_step_state()
        eng.solve_tracers_ale_resident(out)
        for hm in (ho, hg):
            hm.update_thickness()
    for k in range(2):
        a, b = hg["tr_arr"][k][wet], ho["tr_arr"][k][wet]
        assert util.rel_err(a, b) < 1e-10
        mg, mo = util.tracer_mass(hg, k), util.tracer_mass(ho, k)
        assert abs(mg - mo) <= 1e-13 * abs(mo) and abs(mg - m0[k]) <= 1e-13 * abs(m0[k])
        assert abs(a.min() - b.min()) <= 1e-12 * abs(b.min()) and abs(a.max() - b.max()) <= 1e-12 * abs(b.max())
        assert a.min() >= ext0[k][0] - 1e-12 and a.max() <= ext0[k][1] + 1e-12
    eng.close()


@pytest.mark.parametrize("name", ["pi", "band"])
def test_relax_to_clim_and_extrema_match_oracle(name):
    """relax_to_clim inside the step (oce_tracer_mod.F90:87-119, between diff_tracers_ale and exchange_nod(tr_arr)) and
    the device-side extrema scan of the resident tracers (what check_blowup / write_step_info test on the host)"""
    from fesom2_b200 import api
    from oracle import binding as ob
    hm = util.build_model(name)
    N, L = hm["tr_arr"].shape[1], hm["tr_arr"].shape[2]
    rng = np.random.default_rng(3)
    lat = hm["coord_nod2D"][:, 1]
    relax2clim = np.where(np.abs(lat) > 1.0, 1.0 / (10.0 * 86400.0), 0.0) * rng.uniform(0.5, 1.0, N)      # ~10-day restoring poleward of 57 deg
    Tclim = hm["tr_arr"][0] + 0.5 * rng.uniform(-1, 1, (N, L))
    Sclim = hm["tr_arr"][1] + 0.1 * rng.uniform(-1, 1, (N, L))
    eng = api.TracerEngine(hm)
    eng.set_clim(1.0 / (10.0 * 86400.0), relax2clim, Tclim, Sclim)
    eng.push_step_state()
    eng.push_tracers()
    o = ob.TracerOracle(hm)
    o.set_clim(1.0 / (10.0 * 86400.0), relax2clim, Tclim, Sclim)
    ref_norelax = None
    out = np.zeros_like(hm["tr_arr"])
    wet = util.wet_mask(hm)
    for _ in range(2):
        eng.solve_tracers_ale_resident(out)
        o.step()
    for k in range(2):
        assert util.rel_err(out[k][wet], hm["tr_arr"][k][wet]) < TOL
        lo, hi = eng.extrema(k + 1)
        own = wet[:hm.ipar("myDim_nod2D")]
        vals = hm["tr_arr"][k][:hm.ipar("myDim_nod2D")][own]
        assert abs(lo - vals.min()) <= 1e-12 * abs(vals.min()) and abs(hi - vals.max()) <= 1e-12 * abs(vals.max())
    # the relaxation really acted: switching it off gives a different field
    hm2 = util.build_model(name)
    o2 = ob.TracerOracle(hm2)
    for _ in range(2):
        o2.step()
    assert util.rel_err(hm2["tr_arr"][0][wet], hm["tr_arr"][0][wet]) > 1e-8
    eng.close()


@pytest.mark.parametrize("name,local,opts,model_kw", [
    ("band", True, {}, {}),
    ("pi", False, {}, {}),
    ("soufflet", True, {"tra_adv_lim": "NONE", "tra_adv_hor": "MUSCL"}, {"w_split": True, "w_max_cfl": 0.2}),     # with the implicit Wvel_i terms
])
def test_kpp_nonlocal_transport_matches_oracle(name, local, opts, model_kw):
    """use_kpp_nonlclflx (oce_ale_tracer.F90:688-781) in the implicit vertical diffusion: ghats*blmc products above and
    below 1 (the MIN acts), non-zero heat and freshwater fluxes, ref_sss and ref_sss_local"""
    from fesom2_b200 import api
    from oracle import binding as ob
    from tests.test_oracle_tracer import _kpp_arrays
    hm = util.build_model(name, water_flux=2e-7, heat_flux=80.0, **model_kw)
    arrs = _kpp_arrays(hm)
    eng = api.TracerEngine(hm, **opts)
    eng.push_step_state()
    eng.push_kpp_nonlocal(*arrs, ref_sss=34.2, ref_sss_local=local)
    g_tr, g_tro = hm["tr_arr"].copy(), hm["tr_arr_old"].copy()
    for _ in range(2):
        eng.solve_tracers_ale(g_tr, g_tro)
    o = ob.TracerOracle(hm, **opts)
    o.set_kpp_nonlocal(*arrs, ref_sss=34.2, ref_sss_local=local)
    for _ in range(2):
        o.step()
    wet = util.wet_mask(hm)
    for k in range(2):
        assert util.rel_err(g_tr[k][wet], hm["tr_arr"][k][wet]) < TOL
    # switching it off again restores the plain step
    eng.push_kpp_nonlocal(*arrs, enable=False)
    h2 = util.build_model(name, water_flux=2e-7, heat_flux=80.0, **model_kw)
    a_tr, a_tro = h2["tr_arr"].copy(), h2["tr_arr_old"].copy()
    eng.solve_tracers_ale(a_tr, a_tro)
    ob.TracerOracle(h2, **opts).step()
    assert util.rel_err(a_tr[0][wet], h2["tr_arr"][0][wet]) < TOL
    eng.close()


def cavity_ring(hm):
    """nodes whose elements start at different levels (the edge of an ice-shelf cavity) and their edge neighbours"""
    C = hm["ulevels_nod2D"] < hm["ulevels_nod2D_max"]
    ed = hm["edges"].reshape(-1, 2) - 1
    ring = C.copy()
    ring[ed[C[ed[:, 1]], 0]] = True
    ring[ed[C[ed[:, 0]], 1]] = True
    return ring


@pytest.mark.parametrize("opts", [{}, {"tra_adv_hor": "MUSCL", "tra_adv_ver": "PPM"}, {"tra_adv_lim": "NONE", "tra_adv_hor": "MUSCL"}])
def test_cavity_patch_matches_oracle(opts):
    """ice-shelf cavity (ulevels 3/4 inside a box; boundary edges of cavity elements, all five level ranges of
    oce_adv_tra_hor.F90:129-209).  Under a cavity edge the reference's limiter reads scratch entries it never wrote
    (oracle.h: fct_cavity_neutral); the CUDA path treats them as neutral, so it is bit-compared with the oracle in that
    mode (tests/test_oracle_tracer.py shows that the oracle's two modes differ only in the ring around the cavity edge)."""
    from fesom2_b200 import api
    from oracle import binding as ob
    hm = util.build_model("cavity")
    eng = api.TracerEngine(hm, **opts)
    eng.push_step_state()
    g_tr, g_tro = hm["tr_arr"].copy(), hm["tr_arr_old"].copy()
    for _ in range(2):
        eng.solve_tracers_ale(g_tr, g_tro)
    eng.close()
    assert np.isfinite(g_tr).all()
    o1 = ob.TracerOracle(hm, fct_cavity_neutral=True, **opts)
    for _ in range(2):
        o1.step()
    wet = util.wet_mask(hm)
    for k in range(2):
        assert util.rel_err(g_tr[k][wet], hm["tr_arr"][k][wet]) < TOL
        assert util.rel_err(g_tro[k], hm["tr_arr_old"][k]) < TOL


@pytest.mark.parametrize("name,kw", [
    ("pi", {"which_ale": "zstar"}),
    ("band", {"which_ale": "zstar", "water_flux": 2e-7}),
    ("band", {"which_ale": "zlevel", "water_flux": 2e-7}),
    ("soufflet", {}),                                              # linfs
    ("pi", {"which_ale": "zstar", "w_split": True, "w_max_cfl": 0.2, "umax": 0.5}),      # Courant numbers above w_max_cfl: Wvel_i != 0
])
def test_device_ale_producers_match_host_mirror(name, kw):
    """fb200_ale_step_begin / _end (compute_hbar_ale, vert_vel_ale, update_thickness_ale on the device) against the
    host-side mirror of the same routines, bit for bit, over four coupled steps; the tracers stepped on the resident,
    device-produced state must equal the oracle stepped on the host-produced one."""
    from fesom2_b200 import api
    from oracle import binding as ob
    kw = dict({"umax": 0.05}, **kw)
    A, B = util.build_model(name, **kw), util.build_model(name, **kw)
    o = ob.TracerOracle(A)
    eng = api.TracerEngine(B)
    eng.push_step_state()
    eng.push_tracers()
    eng.ale_init()
    out = np.zeros_like(B["tr_arr"])
    wet = util.wet_mask(A)
    myE = A.ipar("myDim_elem2D")
    for s in range(4):
        if s > 0:
            util.advance_ale_state(A, first=False)
            eng.ale_step_begin()
            for nm in ("hbar", "ssh_rhs_old", "dhe"):
                assert np.array_equal(A[nm], B[nm]), (s, nm)
            for nm in ("Wvel", "hnode_new") + (("Wvel_e", "Wvel_i") if kw.get("w_split") else ()):
                assert np.array_equal(eng.debug(nm), A[nm]), (s, nm)
        o.step()
        eng.solve_tracers_ale_resident(out)
        for k in range(2):
            assert util.rel_err(out[k][wet], A["tr_arr"][k][wet]) < TOL, (s, k)
        A.update_thickness()
        eng.ale_step_end()
        for nm in ("hnode", "Z_3d_n", "zbar_3d_n"):
            assert np.array_equal(eng.debug(nm), A[nm]), (s, nm)
        assert np.array_equal(eng.debug("helem")[:myE], A["helem"][:myE]), s
    eng.close()
