"""Shared builders of test cases: a HostModel (the Fortran driver's state) for each reference mesh fixture or
synthetic mesh, with a consistent state for the tracer step."""
import os

import numpy as np

from harness import synth
from harness.host import HostModel

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def mesh_fixture(name):
    z = np.load(os.path.join(GOLD, name + "_mesh.npz"))
    return {k: z[k] for k in z.files}


def build_model(name="pi", dt=None, which_ale=None, w_split=False, w_max_cfl=1.0, num_tracers=2, seed=12345,
                water_flux=0.0, heat_flux=0.0, umax=0.5, **mesh_kw):
    """name: 'pi' (test_pi: rotated global mesh, zstar), 'soufflet' (test_souf: cyclic channel, linfs),
    'band' (synthetic lon/lat band with land and ragged bathymetry), 'farclike' (band refined 3x towards the north: the
    stand-in for BASELINE.json's farc mesh) or 'core2like' (band of CORE2's size: ~127 k surface nodes x 47 layers)"""
    hm = HostModel()
    if name == "pi":
        m = mesh_fixture("pi")
        hm.set_global_mesh(m["lon"], m["lat"], m["tri"], m["zbar"], m["depth"], m["nlevels_nod"], m["nlevels_elem"],
                           m["edges"], m["edge_tri"], int(m["edge2D_in"]), force_rotation=True, cyclic_deg=360.0)
        dt = 900.0 if dt is None else dt                # setups/test_pi: step_per_day 96
        which_ale = which_ale or "zstar"
        pc = True
    elif name == "soufflet":
        m = mesh_fixture("soufflet")
        hm.set_global_mesh(m["lon"], m["lat"], m["tri"], m["zbar"], m["depth"], m["nlevels_nod"], m["nlevels_elem"],
                           m["edges"], m["edge_tri"], int(m["edge2D_in"]), force_rotation=False, cyclic_deg=4.5)
        dt = 1200.0 if dt is None else dt               # setups/test_souf: step_per_day 72
        which_ale = which_ale or "linfs"
        pc = False
    elif name in ("band", "farclike", "core2like"):
        kw = dict(nlon=72, nlat=40, nl=24)
        if name == "farclike":           # BASELINE.json config 4 stand-in: multi-resolution (3x finer north of 65 N)
            kw = dict(nlon=120, nlat=90, nl=24, refine_north=3.0)
        if name == "core2like":          # BASELINE.json config 3 stand-in: ~127 k surface nodes, 47 layers
            kw = dict(nlon=528, nlat=264, nl=48)
        kw.update(mesh_kw)
        m = synth.lonlat_band_mesh(**kw)
        hm.set_global_mesh(m["lon"], m["lat"], m["tri"], m["zbar"], m["depth"], force_rotation=False, cyclic_deg=360.0)
        dt = 1800.0 if dt is None else dt
        which_ale = which_ale or "zstar"
        pc = True
    elif name == "cavity":
        # band mesh with an ice-shelf cavity patch ("a second run with a cavity patch"): elements
        # inside a box get ulevels = 3 or 4 (neighbouring elements differ, so all five level ranges of
        # oce_adv_tra_hor.F90:129-209 occur), nodes the minimum over their elements (find_levels_cavity).  linfs: under
        # zstar the reference treats cavity nodes as linfs, which the host-side mirror of the ALE setup does not restate.
        kw = dict(nlon=72, nlat=40, nl=24)
        kw.update(mesh_kw)
        m = synth.lonlat_band_mesh(**kw)
        hm.set_global_mesh(m["lon"], m["lat"], m["tri"], m["zbar"], m["depth"], force_rotation=False, cyclic_deg=360.0)
        tri = m["tri"] - 1
        inb = (m["lat"] < -55) & (m["lon"] > 100) & (m["lon"] < 220)
        ule, uln, nle = hm["ulevels"], hm["ulevels_nod2D"], hm["nlevels"]
        cav = np.flatnonzero(inb[tri].all(axis=1) & (nle - 4 >= 4))
        ule[cav] = np.where(np.arange(cav.size) % 2 == 0, 4, 3)
        un = np.full(m["lon"].size, 10 ** 6)
        for k in range(3):
            np.minimum.at(un, tri[:, k], ule)
        uln[:] = un
        dt = 1800.0 if dt is None else dt
        which_ale = which_ale or "linfs"
        pc = True
    else:
        raise ValueError(name)
    hm.derive()
    hm.setup_ocean(dt=dt, which_ale=which_ale, use_partial_cell=pc, num_tracers=num_tracers, w_split=w_split,
                   w_max_cfl=w_max_cfl)
    synth.fill_state(hm, seed=seed, water_flux=water_flux, heat_flux=heat_flux, umax=umax)
    return hm


def fill_redi_sw(hm, seed=7):
    """neutral slopes (|s| <= 2e-3, smooth + noise) and a shortwave profile decaying with depth: the inputs
    diff_part_hor_redi / diff_ver_part_impl_ale read when Redi and use_sw_pene are on"""
    rng = np.random.default_rng(seed)
    sl = hm["slope_tapered"]
    n, L, _ = sl.shape
    x = np.linspace(0.0, 6.0, n)[:, None]
    z = np.arange(L)[None, :]
    sl[:, :, 0] = 1e-3 * np.sin(x + 0.1 * z) + 2e-4 * rng.uniform(-1, 1, (n, L))
    sl[:, :, 1] = 1e-3 * np.cos(2 * x - 0.2 * z) + 2e-4 * rng.uniform(-1, 1, (n, L))
    sl[:, :, 2] = np.sqrt(sl[:, :, 0] ** 2 + sl[:, :, 1] ** 2)
    sw = hm["sw_3d"]
    sw[:, :] = 1e-5 * (1.0 + 0.3 * np.sin(3 * x)) * np.exp(-np.arange(sw.shape[1])[None, :] / 3.0)


def wet_mask(hm):
    nl = hm.ipar("nl")
    lev = np.arange(1, nl)[None, :]
    return (lev >= hm["ulevels_nod2D"][:, None]) & (lev <= (hm["nlevels_nod2D"] - 1)[:, None])


def rel_err(a, b):
    """max |a-b| / max |b| : the 'relative 1e-12 per field' measure of BASELINE.json's north_star"""
    d = np.abs(np.asarray(a) - np.asarray(b)).max()
    s = np.abs(np.asarray(b)).max()
    return d / s if s > 0 else d


def advance_ale_state(hm, first):
    """the host side of one time step around solve_tracers_ale with a prescribed velocity field (oce_timestep_ale,
    src/oce_ale.F90:2694-2760): compute_ssh_rhs_ale -> compute_hbar_ale -> vert_vel_ale before the tracer step (skipped
    on the first step: fill_state has already run them), update_thickness_ale after it (the caller runs hm.update_thickness())"""
    if not first:
        hm.compute_ssh_rhs()
        hm.compute_hbar()
        hm.vert_vel()


def tracer_mass(hm, k):
    """sum_n sum_nz T h areasvol over wet cells with the current thickness (integrate_nod_3D, gen_support.F90:288-311)"""
    nl = hm.ipar("nl")
    wet = wet_mask(hm)
    return float((hm["tr_arr"][k] * hm["hnode"] * hm["areasvol"][:, :nl - 1])[wet].sum())


def soufflet_zonal_setup(hm, nybins=100, ysize=2000000.0, lat0=0.0):
    """the soufflet channel's zonal restoring inputs: (ztem, Tclim, lat0, dy, tau_inv).  ztem (nybins, nl-1) follows
    compute_zonal_mean_ini / compute_zonal_mean (toy_channel_soufflet.F90:110-200) on one partition; Tclim is the initial
    temperature (initial_state_soufflet sets Tclim = tr_arr(:,:,1)); tau_inv = 1/50 days (:19)"""
    r_earth = 6367500.0
    dy = ysize / nybins / r_earth
    nl = hm.ipar("nl")
    myE = hm.ipar("myDim_elem2D")
    en = hm["elem2D_nodes"][:myE] - 1
    lat = hm["coord_nod2D"][:, 1]
    ymean = ((lat[en[:, 0]] + lat[en[:, 1]]) + lat[en[:, 2]]) / 3.0
    bpos = np.floor((ymean - lat0) / dy).astype(np.int64) + 1
    assert bpos.min() >= 1 and bpos.max() <= nybins
    T = hm["tr_arr"][0]
    ztem = np.zeros((nybins, nl - 1))
    znum = np.zeros((nybins, nl - 1))
    nle = hm["nlevels"]
    for e in range(myE):
        znum[bpos[e] - 1, :] += 1.0
        k = nle[e] - 1
        ztem[bpos[e] - 1, :k] += ((T[en[e, 0], :k] + T[en[e, 1], :k]) + T[en[e, 2], :k]) / 3.0
    ztem = np.ascontiguousarray(ztem / (znum + 0.001))
    return ztem, np.ascontiguousarray(T.copy()), lat0, dy, 1.0 / 50.0 / 24.0 / 3600.0


def update_vel_host(hm, g=9.81):
    """update_vel (oce_dyn.F90:101-127) on the host model, own elements: UV += UV_rhs - g theta dt grad(d_eta); eta_n += d_eta"""
    myE, theta, dt = hm.ipar("myDim_elem2D"), hm.par("theta"), hm.par("dt")
    en = hm["elem2D_nodes"][:myE] - 1
    eta = -(g * theta * dt * hm["d_eta"][en])                          # (myE, 3)
    gs = hm["gradient_sca"][:myE]                                      # (myE, 6)
    Fx = (gs[:, 0] * eta[:, 0] + gs[:, 1] * eta[:, 1]) + gs[:, 2] * eta[:, 2]
    Fy = (gs[:, 3] * eta[:, 0] + gs[:, 4] * eta[:, 1]) + gs[:, 5] * eta[:, 2]
    nl = hm.ipar("nl")
    lev = np.arange(1, nl)[None, :]
    act = (lev >= hm["ulevels"][:myE, None]) & (lev <= (hm["nlevels"][:myE] - 1)[:, None])
    UV, UVr = hm["UV"], hm["UV_rhs"]
    UV[:myE, :, 0] = np.where(act, (UV[:myE, :, 0] + UVr[:myE, :, 0]) + Fx[:, None], UV[:myE, :, 0])
    UV[:myE, :, 1] = np.where(act, (UV[:myE, :, 1] + UVr[:myE, :, 1]) + Fy[:, None], UV[:myE, :, 1])
    hm["eta_n"][:] = hm["eta_n"] + hm["d_eta"]
