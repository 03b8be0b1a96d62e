"""Synthetic meshes and model states: inputs for parity tests and for bench.py.

Nothing here is timed.  The fields are smooth analytic profiles plus seeded noise; the vertical velocity and
the new layer thicknesses are NOT analytic — they come from the restated `compute_hbar_ale` / `vert_vel_ale`
(harness/host/ale.cpp), so that discrete continuity holds exactly as it does in the reference and a
constant tracer stays constant.
"""
import numpy as np

from .host import HostModel


def elem_centers(hm):
    """lon/lat (radians, model coordinates) of element centroids; cyclic-aware like elem_center (oce_mesh.F90:1818)"""
    c = hm["coord_nod2D"]                 # (N, 2)
    en = hm["elem2D_nodes"] - 1           # (E, 3)
    cyc = hm.par("cyclic_length")
    ax = c[en, 0]
    amin = ax.min(axis=1, keepdims=True)
    ax = np.where(ax - amin >= cyc / 2, ax - cyc, ax)
    ax = np.where(ax - amin < -cyc / 2, ax + cyc, ax)
    return ax.mean(axis=1), c[en, 1].mean(axis=1)


def fill_state(hm, seed=12345, umax=0.5, hbar_amp=0.3, noise=0.1, water_flux=0.0, heat_flux=0.0):
    """Fill tracers, velocities, mixing coefficients and surface elevation of an already set-up HostModel,
    then run the reference's producers (compute_hbar_ale, vert_vel_ale) once so that the state is the one
    `solve_tracers_ale` sees in the time loop (src/oce_ale.F90:2709-2752)."""
    rng = np.random.default_rng(seed)
    nl = hm.ipar("nl")
    N = hm.ipar("myDim_nod2D") + hm.ipar("eDim_nod2D")
    E = hm.ipar("myDim_elem2D") + hm.ipar("eDim_elem2D")
    zbar = hm["zbar"]
    Z = hm["Z"]
    lon, lat = hm["coord_nod2D"][:, 0], hm["coord_nod2D"][:, 1]
    nln, uln = hm["nlevels_nod2D"], hm["ulevels_nod2D"]
    lev = np.arange(1, nl)[None, :]
    wet = (lev >= uln[:, None]) & (lev <= (nln - 1)[:, None])          # (N, nl-1)

    # surface elevation, then ALE thicknesses consistent with it (init_thickness_ale)
    if hm.ipar("which_ale") != 0:
        hb = hbar_amp * np.sin(2 * lon) * np.cos(lat)
        hm["hbar"][:] = hb
        hm["hbar_old"][:] = hb
    hm.init_thickness()

    tr = hm["tr_arr"]          # (ntr, N, nl-1)
    tro = hm["tr_arr_old"]
    T = 2.0 + 18.0 * np.cos(lat)[:, None] ** 2 * np.exp(Z / 800.0)[None, :] + noise * rng.uniform(-1, 1, (N, nl - 1))
    S = 34.5 + 0.5 * np.sin(3 * lon)[:, None] * np.exp(Z / 500.0)[None, :] + 0.1 * noise * rng.uniform(-1, 1, (N, nl - 1))
    tr[0] = np.where(wet, T, 0.0)
    tro[0] = np.where(wet, T + 0.5 * noise * rng.uniform(-1, 1, (N, nl - 1)), 0.0)
    if tr.shape[0] > 1:
        tr[1] = np.where(wet, S, 0.0)
        tro[1] = np.where(wet, S + 0.05 * noise * rng.uniform(-1, 1, (N, nl - 1)), 0.0)
    for k in range(2, tr.shape[0]):
        P = 1.0 + 0.5 * np.cos(2 * lat)[:, None] * np.exp(Z / 300.0)[None, :]
        tr[k] = np.where(wet, P, 0.0)
        tro[k] = np.where(wet, P, 0.0)

    elon, elat = elem_centers(hm)
    nle, ule = hm["nlevels"][:E], hm["ulevels"][:E]
    wet_e = (lev >= ule[:, None]) & (lev <= (nle - 1)[:, None])
    prof = np.exp(Z / 1500.0)[None, :]
    u = umax * np.cos(elat)[:, None] * np.cos(2 * elon)[:, None] * prof + 0.02 * umax * rng.uniform(-1, 1, (E, nl - 1))
    v = 0.5 * umax * np.sin(2 * elat)[:, None] * np.sin(3 * elon)[:, None] * prof + 0.02 * umax * rng.uniform(-1, 1, (E, nl - 1))
    UV = hm["UV"]              # (E, nl-1, 2)
    UV[:, :, 0] = np.where(wet_e, u, 0.0)
    UV[:, :, 1] = np.where(wet_e, v, 0.0)

    hm["Kv"][:] = (1e-5 + 1e-2 * np.exp(zbar / 50.0))[None, :]
    hm["water_flux"][:] = water_flux * np.cos(lat)
    hm["heat_flux"][:] = heat_flux * np.cos(2 * lat)

    # producers of the tracer step's inputs, in the reference's order
    if hm.ipar("which_ale") != 0:
        hm.compute_hbar()
    hm.vert_vel()
    return hm


# ---------------------------------------------------------------------------------------------------
def lonlat_band_mesh(nlon, nlat, lat_max=80.0, nl=24, depth_max=5000.0, seed=20261104, land_frac=0.12,
                     shuffle=False, refine_north=0.0, rcm=False):
    """Triangulated lon/lat band (cyclic in longitude) with elliptical land patches and ragged bathymetry.
    Returns the raw arrays `HostModel.set_global_mesh` takes.  Structured-like but stored fully unstructured.
    refine_north = r > 1: multi-resolution variant (the stand-in for the Arctic-refined farc mesh of BASELINE.json
    config 4, which is not in the repository): the meridional spacing shrinks by the factor r between 50 N and 65 N
    and stays there up to lat_max.  shuffle: random node and element numbering (no locality at all).  rcm: the shuffled mesh
    renumbered by reverse Cuthill-McKee on the node graph, elements sorted by their smallest node — what an offline
    bandwidth-reducing renumbering makes of a badly numbered mesh."""
    rng = np.random.default_rng(seed)
    lons = np.linspace(0.0, 360.0, nlon, endpoint=False)
    lats = np.linspace(-lat_max, lat_max, nlat)
    if refine_north and refine_north > 1.0:
        fine = np.linspace(-lat_max, lat_max, 20 * nlat)
        dens = 1.0 + (refine_north - 1.0) * np.clip((fine - 50.0) / 15.0, 0.0, 1.0)      # rows per degree, relative
        cdf = np.concatenate([[0.0], np.cumsum(0.5 * (dens[1:] + dens[:-1]) * np.diff(fine))])
        lats = np.interp(np.linspace(0.0, cdf[-1], nlat), cdf, fine)
    LON, LAT = np.meshgrid(lons, lats, indexing="xy")          # (nlat, nlon)
    # land mask from a few ellipses
    land = np.zeros_like(LON, dtype=bool)
    nblob = max(1, int(round(land_frac * 25)))
    for _ in range(nblob):
        c0, c1 = rng.uniform(0, 360), rng.uniform(-0.8 * lat_max, 0.8 * lat_max)
        a0, a1 = rng.uniform(15, 40), rng.uniform(8, 25)
        dl = (LON - c0 + 180.0) % 360.0 - 180.0
        land |= (dl / a0) ** 2 + ((LAT - c1) / a1) ** 2 < 1.0
    # quads -> triangles, keep triangles with no land corner
    j, i = np.meshgrid(np.arange(nlat - 1), np.arange(nlon), indexing="ij")
    n00 = j * nlon + i
    n10 = j * nlon + (i + 1) % nlon
    n01 = (j + 1) * nlon + i
    n11 = (j + 1) * nlon + (i + 1) % nlon
    flip = ((i + j) % 2 == 0)
    t1 = np.where(flip[..., None], np.stack([n00, n10, n11], -1), np.stack([n00, n10, n01], -1))
    t2 = np.where(flip[..., None], np.stack([n00, n11, n01], -1), np.stack([n10, n11, n01], -1))
    tri = np.concatenate([t1.reshape(-1, 3), t2.reshape(-1, 3)], axis=0)
    lf = land.reshape(-1)
    tri = tri[~lf[tri].any(axis=1)]
    # drop nodes without elements and elements with < 2 neighbours (iteratively), renumber
    for _ in range(50):
        used = np.zeros(nlon * nlat, dtype=bool)
        used[tri.reshape(-1)] = True
        # element neighbours through shared edges
        e = np.sort(np.concatenate([tri[:, [0, 1]], tri[:, [1, 2]], tri[:, [2, 0]]], axis=0), axis=1)
        key = e[:, 0].astype(np.int64) * (nlon * nlat) + e[:, 1]
        _, inv, cnt = np.unique(key, return_inverse=True, return_counts=True)
        shared = (cnt[inv] == 2).reshape(3, -1).sum(axis=0)
        keep = shared >= 2
        if keep.all():
            break
        tri = tri[keep]
    used = np.zeros(nlon * nlat, dtype=bool)
    used[tri.reshape(-1)] = True
    newid = np.cumsum(used) - 1
    tri = newid[tri]
    lon = LON.reshape(-1)[used]
    lat = LAT.reshape(-1)[used]
    if shuffle or rcm:                # destroy the numbering locality (what a real mesh may look like)
        perm = rng.permutation(lon.size)
        inv = np.empty_like(perm)
        inv[perm] = np.arange(perm.size)
        lon, lat = lon[perm], lat[perm]
        tri = inv[tri]
        tri = tri[rng.permutation(tri.shape[0])]
    if rcm:
        import scipy.sparse as sp
        from scipy.sparse.csgraph import reverse_cuthill_mckee
        n = lon.size
        a = np.concatenate([tri[:, 0], tri[:, 1], tri[:, 2]])
        b = np.concatenate([tri[:, 1], tri[:, 2], tri[:, 0]])
        G = sp.coo_matrix((np.ones(a.size, np.int8), (a, b)), shape=(n, n)).tocsr()
        G = ((G + G.T) > 0).astype(np.int8).tocsr()
        perm = np.asarray(reverse_cuthill_mckee(G, symmetric_mode=True))
        inv = np.empty_like(perm)
        inv[perm] = np.arange(n)
        lon, lat = lon[perm], lat[perm]
        tri = inv[tri]
        tri = tri[np.argsort(tri.min(axis=1), kind="stable")]
    # vertical grid: stretched, 5 m at the top
    k = np.arange(nl)
    zbar = -depth_max * (np.expm1(3.0 * k / (nl - 1)) / np.expm1(3.0))
    zbar[1] = min(zbar[1], -5.0) if nl > 2 else zbar[1]
    zbar = -np.sort(-zbar)
    lr, tr_ = np.radians(lon), np.radians(lat)
    sm = np.abs(0.6 * np.sin(3 * lr) * np.cos(2 * tr_) + 0.4 * np.cos(5 * lr + 1.0) * np.sin(3 * tr_ + 0.5))
    depth = -(0.2 * depth_max + 0.8 * depth_max * sm)
    return dict(lon=lon, lat=lat, tri=(tri + 1).astype(np.int32), zbar=zbar, depth=depth)


def model_from_raw(m, force_rotation=False, cyclic_deg=360.0):
    hm = HostModel()
    hm.set_global_mesh(m["lon"], m["lat"], m["tri"], m["zbar"], m["depth"], m.get("nlevels_nod"), m.get("nlevels_elem"),
                       m.get("edges"), m.get("edge_tri"), m.get("edge2D_in", 0), force_rotation=force_rotation,
                       cyclic_deg=cyclic_deg)
    hm.derive()
    return hm
