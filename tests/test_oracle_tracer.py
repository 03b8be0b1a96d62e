"""CPU tests (no GPU): the oracle's restatement of the tracer step against the invariants the reference's
code implies; the reference ships no kernel-level golden vectors for these routines."""
import numpy as np
import pytest

from oracle import binding as ob
from tests import util


@pytest.mark.parametrize("name,kw", [("pi", {}), ("soufflet", {}), ("band", {}), ("pi", {"w_split": True, "w_max_cfl": 0.2}),
                                     ("cavity", {})])
def test_constant_tracer_is_preserved(name, kw):
    hm = util.build_model(name, **kw)
    wet = util.wet_mask(hm)
    tr, tro = hm["tr_arr"], hm["tr_arr_old"]
    tr[0] = np.where(wet, 7.0, 0.0)
    tro[0] = tr[0]
    o = ob.TracerOracle(hm)
    o.step()
    assert np.abs(tr[0][wet] - 7.0).max() < 5e-14
    assert not np.isnan(tr).any()
    assert np.array_equal(tro[0][wet], np.full(wet.sum(), 7.0))      # tr_arr_old = tr_arr (oce_ale_tracer.F90:274)


@pytest.mark.parametrize("name", ["pi", "soufflet", "band"])
def test_mass_conservation_without_surface_flux(name):
    """sum_n sum_nz T*h*areasvol is conserved by advection + diffusion on a closed domain (integrate_nod_3D,
    gen_support.F90:288-311, with the ALE thickness)"""
    hm = util.build_model(name, which_ale="zstar")     # linfs lets tracer leave through the fixed surface (W(1) != 0)
    wet = util.wet_mask(hm)
    nl = hm.ipar("nl")
    vol_old = hm["hnode"] * hm["areasvol"][:, :nl - 1]
    vol_new = hm["hnode_new"] * hm["areasvol"][:, :nl - 1]
    tr = hm["tr_arr"]
    m0 = [(tr[k] * vol_old)[wet].sum() for k in range(2)]
    o = ob.TracerOracle(hm, clamp_salt=False)
    o.step()
    m1 = [(tr[k] * vol_new)[wet].sum() for k in range(2)]
    for a, b in zip(m0, m1):
        assert abs(a - b) / abs(a) < 1e-12


def test_fct_solution_stays_within_low_order_bounds():
    """after advection alone (no diffusion) the FCT solution lies within the min/max of (T, LO) over the
    node's cluster and the levels above/below (oce_adv_tra_fct.F90:93-162)"""
    hm = util.build_model("pi")
    wet = util.wet_mask(hm)
    hm["Kv"][:] = 0.0
    T0 = hm["tr_arr"][0].copy()
    o = ob.TracerOracle(hm, do_hor_diff=False, clamp_salt=False)
    for ph in (1, 2, 3, 4):
        o.phase(ph, 1)
    nl = hm.ipar("nl")
    N = hm.ipar("myDim_nod2D")
    LO = o.work("fct_LO", (nl - 1, N)).copy()
    o.phase(5, 1)
    T1 = hm["tr_arr"][0]
    hi = np.maximum(T0, LO)
    lo = np.minimum(T0, LO)
    gmax, gmin = hi[wet].max(), lo[wet].min()
    assert T1[wet].max() <= gmax + 1e-9 and T1[wet].min() >= gmin - 1e-9


@pytest.mark.parametrize("hor,ver,lim", [("MFCT", "QR4C", "FCT"), ("MUSCL", "QR4C", "FCT"), ("UPW1", "UPW1", "FCT"), ("MFCT", "CDIFF", "FCT"),
                                         ("MFCT", "PPM", "FCT"), ("MUSCL", "QR4C", "NONE"), ("MUSCL", "PPM", "NONE"), ("UPW1", "CDIFF", "NONE")])
def test_all_scheme_combinations_run_and_conserve(hor, ver, lim):
    hm = util.build_model("band")
    wet = util.wet_mask(hm)
    nl = hm.ipar("nl")
    tr = hm["tr_arr"]
    m0 = (tr[0] * hm["hnode"] * hm["areasvol"][:, :nl - 1])[wet].sum()
    o = ob.TracerOracle(hm, tra_adv_hor=hor, tra_adv_ver=ver, tra_adv_lim=lim, clamp_salt=False)
    o.step()
    m1 = (tr[0] * hm["hnode_new"] * hm["areasvol"][:, :nl - 1])[wet].sum()
    assert np.isfinite(tr).all()
    assert abs(m0 - m1) / abs(m0) < 1e-12


def test_explicit_vertical_diffusion_runs():
    """i_vert_diff=.false. (diff_ver_part_expl_ale, oce_ale_tracer.F90:330-393).  As written in the reference the flux
    divergence is divided by the layer thickness before it joins the thickness-weighted del_ttf, so this legacy
    variant is not exactly conservative; the restatement follows the code, the test only bounds the drift."""
    hm = util.build_model("band")
    wet = util.wet_mask(hm)
    nl = hm.ipar("nl")
    tr = hm["tr_arr"]
    m0 = (tr[0] * hm["hnode"] * hm["areasvol"][:, :nl - 1])[wet].sum()
    o = ob.TracerOracle(hm, i_vert_diff=False, clamp_salt=False)
    o.step()
    m1 = (tr[0] * hm["hnode_new"] * hm["areasvol"][:, :nl - 1])[wet].sum()
    assert np.isfinite(tr).all()
    assert abs(m0 - m1) / abs(m0) < 1e-6


@pytest.mark.parametrize("name", ["pi", "band"])
def test_100_coupled_steps_conserve_mass_and_keep_extrema(name):
    """BASELINE.json north_star: 'stated bounds on global tracer mass and extrema after 100 steps'.  100 steps of the
    time loop (compute_hbar_ale -> vert_vel_ale -> solve_tracers_ale -> update_thickness_ale, zstar, prescribed
    velocities, no surface flux): tracer mass sum(T h areasvol) is conserved to 1e-13 relative, and FCT advection +
    diffusion create no new extrema (every value stays inside the initial [min, max])."""
    hm = util.build_model(name, which_ale="zstar", umax=0.02)
    wet = util.wet_mask(hm)
    tr = hm["tr_arr"]
    m0 = [util.tracer_mass(hm, k) for k in range(2)]
    ext0 = [(tr[k][wet].min(), tr[k][wet].max()) for k in range(2)]
    o = ob.TracerOracle(hm, clamp_salt=False)
    for s in range(100):
        util.advance_ale_state(hm, first=(s == 0))
        o.step()
        hm.update_thickness()
    assert hm["hnode"][wet].min() > 0.0
    for k in range(2):
        assert abs(util.tracer_mass(hm, k) - m0[k]) <= 1e-13 * abs(m0[k])
        assert tr[k][wet].min() >= ext0[k][0] - 1e-12 and tr[k][wet].max() <= ext0[k][1] + 1e-12


def test_relax_to_clim_pulls_towards_climatology():
    """relax_to_clim (oce_tracer_mod.F90:87-119): T_new = T + relax2clim*dt*(Tclim - T) on own wet cells after the
    diffusion step; with relax2clim*dt = 1 the tracer lands exactly on the climatology, with 0 nothing changes"""
    hm = util.build_model("band")
    wet = util.wet_mask(hm)
    N, L = hm["tr_arr"].shape[1], hm["tr_arr"].shape[2]
    dt = hm.par("dt")
    half = np.arange(N) % 2 == 0
    relax2clim = np.where(half, 1.0 / dt, 0.0)
    Tclim = np.full((N, L), 5.0)
    Sclim = np.full((N, L), 35.0)
    ref = util.build_model("band")
    ob.TracerOracle(ref).step()
    o = ob.TracerOracle(hm)
    o.set_clim(1.0, relax2clim, Tclim, Sclim)
    o.step()
    on = wet & half[:, None]
    off = wet & ~half[:, None]
    assert np.abs(hm["tr_arr"][0][on] - 5.0).max() < 1e-12 and np.abs(hm["tr_arr"][1][on] - 35.0).max() < 1e-12
    assert np.array_equal(hm["tr_arr"][0][off], ref["tr_arr"][0][off])


def _kpp_arrays(hm, seed=11):
    rng = np.random.default_rng(seed)
    N, nl = hm["Kv"].shape
    z = np.arange(nl)[None, :]
    g = 2.0 * np.exp(-z / 6.0) * rng.uniform(0.2, 1.0, (N, nl))          # "ghats": decays below the boundary layer
    gS = 1.5 * np.exp(-z / 5.0) * rng.uniform(0.2, 1.0, (N, nl))
    b2 = rng.uniform(0.0, 1.2, (N, nl))                                    # ghats*blmc exceeds 1 in places: the MIN(.., 1) acts
    b3 = rng.uniform(0.0, 1.2, (N, nl))
    return [np.ascontiguousarray(a) for a in (g, gS, b2, b3)]


def test_kpp_nonlocal_transport_redistributes_without_changing_the_column_content():
    """use_kpp_nonlclflx (oce_ale_tracer.F90:688-781): the non-local flux is a telescoping sum over the column (no flux
    through the surface or the bottom), so it changes the profile but not sum_z T h area — the step with it conserves
    exactly what the step without it conserves"""
    kw = dict(water_flux=2e-7, heat_flux=80.0)
    ha, hb = util.build_model("band", **kw), util.build_model("band", **kw)
    wet = util.wet_mask(ha)
    nl = ha.ipar("nl")
    oa, obb = ob.TracerOracle(ha, clamp_salt=False), ob.TracerOracle(hb, clamp_salt=False)
    oa.set_kpp_nonlocal(*_kpp_arrays(ha), ref_sss=34.7, ref_sss_local=True)
    oa.step()
    obb.step()
    for k in range(2):
        ca = (ha["tr_arr"][k] * ha["hnode_new"] * ha["areasvol"][:, :nl - 1] * wet).sum(axis=1)
        cb = (hb["tr_arr"][k] * hb["hnode_new"] * hb["areasvol"][:, :nl - 1] * wet).sum(axis=1)
        assert np.abs(ca - cb).max() <= 1e-11 * np.abs(cb).max()
        assert util.rel_err(ha["tr_arr"][k][wet], hb["tr_arr"][k][wet]) > 1e-9            # but the profile did change


def test_cavity_limiter_scratch_only_matters_at_the_cavity_edge():
    """oce_tra_adv_fct a3 reads element scratch above an element's cavity top that a2 never wrote (oracle.h:
    fct_cavity_neutral).  The two readings of that (zeros as allocated / neutral entries) may only differ at nodes whose
    elements start at different levels and at their edge neighbours; both keep a constant tracer constant."""
    from tests.test_gpu_tracer import cavity_ring
    ha, hb = util.build_model("cavity"), util.build_model("cavity")
    ob.TracerOracle(ha).step()
    ob.TracerOracle(hb, fct_cavity_neutral=True).step()
    wet = util.wet_mask(ha)
    ring = cavity_ring(ha)
    assert ring.sum() > 0 and (~ring).sum() > 0
    for k in range(2):
        far = wet & ~ring[:, None]
        assert np.array_equal(ha["tr_arr"][k][far], hb["tr_arr"][k][far])
    hc = util.build_model("cavity")
    hc["tr_arr"][0] = np.where(wet, 7.0, 0.0)
    hc["tr_arr_old"][0] = hc["tr_arr"][0]
    ob.TracerOracle(hc, fct_cavity_neutral=True).step()
    assert np.abs(hc["tr_arr"][0][wet] - 7.0).max() < 5e-14


def test_relax_zonal_temp_adds_the_zonal_restoring_once_per_tracer_pass():
    """relax_zonal_temp (toy_channel_soufflet.F90:78-103) replaces relax_to_clim on the soufflet channel: the increment
    dt*tau_inv*(Tclim - Tzon) does not depend on the current temperature and the tracer loop reaches the routine once per
    tracer (oce_ale_tracer.F90:150-151), so after one step T = (T_without + inc) + inc exactly and S is untouched."""
    A, B = util.build_model("soufflet"), util.build_model("soufflet")
    ztem, Tclim, lat0, dy, tau_inv = util.soufflet_zonal_setup(B)
    Tclim = Tclim + 0.3                                   # make the restoring term non-trivial
    ob.TracerOracle(A).step()
    o = ob.TracerOracle(B)
    o.set_zonal_relax(ztem, np.ascontiguousarray(B["coord_nod2D"]), Tclim, lat0, dy, tau_inv)
    o.step()
    nl, dt = B.ipar("nl"), B.par("dt")
    yy = B["coord_nod2D"][:, 1] - lat0
    nn = np.where(yy < dy / 2, 1, np.floor(yy / dy - 0.5).astype(np.int64) + 1)
    nn1 = np.where(yy < dy / 2, 1, np.where(nn + 1 > 100, nn, nn + 1))
    a = np.where(yy < dy / 2, 0.0, yy / dy + 0.5 - nn)
    Tzon = (1.0 - a)[:, None] * ztem[nn - 1] + a[:, None] * ztem[nn1 - 1]
    inc = dt * tau_inv * (Tclim - Tzon)
    lev = np.arange(1, nl)[None, :] <= (B["nlevels_nod2D"] - 1)[:, None]
    want = np.where(lev, (A["tr_arr"][0] + inc) + inc, A["tr_arr"][0])
    assert np.abs(inc[lev]).max() > 1e-4
    assert np.array_equal(B["tr_arr"][0], want)
    assert np.array_equal(B["tr_arr"][1], A["tr_arr"][1])


@pytest.mark.parametrize("accuracy", [8, 4])
def test_output_mean_stream_restatement(accuracy):
    """update_means / compute_means (io_meandata.F90:768-794, 894-900) in the stream's precision: against numpy in that precision"""
    rng = np.random.default_rng(3)
    fields = [15.0 + rng.standard_normal((50, 7)) for _ in range(5)]
    ms = ob.MeanStream((50, 7), accuracy)
    dt = np.float64 if accuracy == 8 else np.float32
    acc = np.zeros((50, 7), dtype=dt)
    for f in fields:
        ms.update(f)
        acc = acc + f.astype(dt)
    got = ms.output()
    assert got.dtype == dt and np.array_equal(got, acc / dt(5))
    assert ms.count == 0 and not ms.acc.any()


def test_redi_explicit_vertical_term_is_there_and_conserves():
    """diff_ver_part_redi_expl (oce_ale_tracer.F90:860-930, call site :288; Redi=.true. is the reference's default,
    config/namelist.oce:40).  (i) the term is not zero on pi: with Redi on, zeroing slope_tapered(1:2) — which leaves the
    Ki*slope(3)^2 terms of the implicit solve and removes only the slope*gradient products of diff_part_hor_redi and of
    diff_ver_part_redi_expl — changes the result, and so does the routine on its own (del_ttf before/after); (ii) the term is
    a flux divergence per column: it adds nothing to sum(del_ttf * areasvol) of a column; (iii) the whole step with Redi
    conserves tracer mass like the step without."""
    hm = util.build_model("pi", which_ale="zstar")
    util.fill_redi_sw(hm)
    wet = util.wet_mask(hm)
    nl, N = hm.ipar("nl"), hm.ipar("myDim_nod2D")
    o = ob.TracerOracle(hm, Redi=True, clamp_salt=False)
    for ph in (1, 2, 3, 4):
        o.phase(ph, 1)
    d = o.work("del_ttf", (nl - 1, N))
    before = d.copy()
    ob.lib().o_diff_ver_part_redi_expl.argtypes = [ob.C.c_void_p]
    ob.lib().o_diff_ver_part_redi_expl.restype = None
    ob.lib().o_diff_ver_part_redi_expl(o.o)
    term = d - before
    assert np.abs(term[wet]).max() > 0.0
    assert not np.any(term[~wet])
    col = (term * hm["areasvol"][:, :nl - 1]).sum(axis=1)
    scale = (np.abs(term) * hm["areasvol"][:, :nl - 1]).sum(axis=1)
    assert np.all(np.abs(col) <= 1e-12 * np.maximum(scale, 1e-300))
    # the full step: mass is conserved with the term in place
    hm2 = util.build_model("pi", which_ale="zstar")
    util.fill_redi_sw(hm2)
    vol_old = hm2["hnode"] * hm2["areasvol"][:, :nl - 1]
    vol_new = hm2["hnode_new"] * hm2["areasvol"][:, :nl - 1]
    tr = hm2["tr_arr"]
    m0 = [(tr[k] * vol_old)[wet].sum() for k in range(2)]
    ob.TracerOracle(hm2, Redi=True, clamp_salt=False).step()
    m1 = [(tr[k] * vol_new)[wet].sum() for k in range(2)]
    for a, b in zip(m0, m1):
        assert abs(a - b) / abs(a) < 1e-12
    # and it matters: the same step without the explicit term (Redi terms of the other two routines kept) differs
    hm3 = util.build_model("pi", which_ale="zstar")
    util.fill_redi_sw(hm3)
    o3 = ob.TracerOracle(hm3, Redi=True, clamp_salt=False)
    o3.L.oracle_set_int(o3.o, b"Redi", 1)
    for k in (1, 2):
        for ph in (1, 2, 3, 4):
            o3.phase(ph, k)
        o3.L.oracle_set_int(o3.o, b"skip_redi_expl", 1)
        o3.phase(5, k)
        o3.L.oracle_set_int(o3.o, b"skip_redi_expl", 0)
    assert np.abs(hm3["tr_arr"][0][wet] - hm2["tr_arr"][0][wet]).max() > 1e-9


def test_biharmonic_filter_conserves_and_smooths():
    """diff_part_bh (oce_ale_tracer.F90:1081-1148): both rounds add +x to one node of an edge and -x (divided by the nodes'
    areas in round 2) to the other, so sum(T * area) over a level is unchanged; a 2-grid-interval checkerboard is damped;
    boundary edges (one triangle) take no part; switched off, the step is untouched."""
    hm = util.build_model("band")
    nl, N = hm.ipar("nl"), hm.ipar("myDim_nod2D")
    wet = util.wet_mask(hm)
    o = ob.TracerOracle(hm, smooth_bh_tra=True)
    L = ob.lib()
    L.o_diff_part_bh.argtypes = [ob.C.c_void_p, ob.C.c_void_p]
    L.o_diff_part_bh.restype = None
    rng = np.random.default_rng(3)
    T = hm["tr_arr"][0]
    T[:] = np.where(wet, 10.0 + rng.uniform(-1, 1, T.shape), 0.0)
    T0 = T.copy()
    area = hm["area"][:, :nl - 1]
    # levels where every node of the level's wet set is wet in all its neighbours' sense: compare per-level sums on the
    # range each edge works on — the total over all levels and nodes is conserved exactly (up to round-off)
    L.o_diff_part_bh(o.o, ob.C.c_void_p(T.ctypes.data))
    assert np.abs(T - T0).max() > 0.0
    s0, s1 = (T0 * area).sum(), (T * area).sum()
    assert abs(s0 - s1) <= 1e-12 * abs(s0)
    # damping: the variance of the noise shrinks
    assert (T[wet] - 10.0).std() < (T0[wet] - 10.0).std()
    # off by default
    A, B = util.build_model("band"), util.build_model("band")
    ob.TracerOracle(A).step()
    ob.TracerOracle(B, smooth_bh_tra=False).step()
    assert np.array_equal(A["tr_arr"], B["tr_arr"])
    C_ = util.build_model("band")
    ob.TracerOracle(C_, smooth_bh_tra=True).step()
    assert not np.array_equal(A["tr_arr"], C_["tr_arr"])
