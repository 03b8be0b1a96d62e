// Device-side declarations shared by the tracer kernel files (tracer_kernels.cu, tracer_kernels2.cu): the kernel
// parameter block, indexing of the level-blocked layout, the reference's flux formulas as inline device functions.
#pragma once
#include "common.cuh"

#define LB(n) __launch_bounds__(n)

struct KP {                       // kernel parameter block (by value)
    int nl, nlp, myN, N, myE, Eall, myEd, maxdeg, maxnie;
    int zero_edge;                 // tile base of an edge-field entry column that is never written (see SlotBatch in tracer_kernels2.cu)
    double dt, ph, pv, eps_ab, is_nonlinfs;
    int hor, ver, w_split, Redi, use_sw_pene, do_hor_diff, tracer_id, tr_num;
    const EdgeRec* edge; const EdgeHo* edgeho; const int* ne_slot; const int* nie_slot; const int* el_nb;
    const int* tile_deg;           // per node tile: largest number of incident edges among its 32 nodes
    const int4* ns_i;              // per (node, slot): tile bases (other node, element 1, element 2 or -1, edge)
    const ulonglong2* ns_m;        // level masks (m1, m2)
    const SlotRec* ns_p;           // packed record the node kernels read: one 16-byte load per (node, slot)
    const double2 *ns_d1, *ns_d2;  // (dX1,dY1), (dX2,dY2)
    const unsigned* el_lev; const double* el_g; const double* el_area; const unsigned* n_lev; const int* n_nbl; const double* n_znb;
    const double2* UV;              // (u, v) per element-level
    const double *area, *areasvol, *helem, *W, *We, *Wi, *hnode, *hnode_new, *Z3, *zbar3, *Kv, *Ki, *sl1, *sl2, *sl3, *sw3d;
    const double *heat_flux, *water_flux, *virtual_salt, *relax_salt, *real_salt_flux;
    const double *kgT, *kgS, *kb2, *kb3; double ref_sss; int ref_sss_local;      // KPP non-local transport (fb200_push_kpp_nonlocal)
    double *T, *Told, *T1;        // this tracer's fields; T1 = tracer 1 (bc_surface)
    double2 *txy, *fpm;            // (tr_xy(1), tr_xy(2)) per element-level; (fct_plus, fct_minus) per node-level
    double *trz, *LO, *adf_h, *adf_v, *tstar;
    const double* redi_v;          // diff_ver_part_redi_expl's contribution to del_ttf per (node, level), from k_redi_expl (Redi only)
    double* eflx;                  // per (edge, level) scratch: upwind flux of ttf (k_ho_edge -> k_lo), then the diffusive flux (k_diff_edge -> k_update)
};

__device__ __forceinline__ int tbase(int ent, int nlp) { return (ent >> 5) * (nlp * TILE) + (ent & 31); }
#define AT(p, base, nz) (p)[(base) + (((nz) - 1) << 5)]          // 1-based level
__device__ __forceinline__ int node_of_base_d(int base, int nlp) { return (base / (nlp * TILE)) * TILE + base % (nlp * TILE); }   // entity id of a tile base

// min/max of two non-NaN doubles as one compare + select (the values here are never NaN; fmin/fmax spend several
// more instructions on NaN propagation)
__device__ __forceinline__ double omin(double a, double b) { return a < b ? a : b; }     // MIN/MAX as the oracle evaluates them
__device__ __forceinline__ double omax(double a, double b) { return a > b ? a : b; }
__device__ __forceinline__ double dmin(double a, double b) { return (b < a) ? b : a; }
__device__ __forceinline__ double dmax(double a, double b) { return (b > a) ? b : a; }
// x / b, correctly rounded, given rb = 1.0 / b: quotient estimate plus one exact-residual correction (the last two
// steps of the IEEE division sequence; the Newton iterations for 1/b are paid once per level instead of per edge).
// tests/test_gpu_tracer.py::test_div_by_matches_ieee_division checks it against '/' on 2^24 operand pairs.
__device__ __forceinline__ double div_by(double x, double b, double rb) {
    const double q = x * rb;
    const double r = fma(-b, q, x);
    return fma(r, rb, q);
}

// Level masks of an edge: bit nz of m1 (m2) is set where the median-dual segment in element 1 (2) carries a flux at
// level nz — the five ranges (A)-(E) of oce_adv_tra_hor.F90:129-209 evaluated once on the host (context.cu), so the
// kernels test two bits instead of re-deriving the ranges for every (edge, level).
__device__ __forceinline__ int mask_segs(unsigned long long m1, unsigned long long m2, int nz) {
    return (int)((m1 >> nz) & 1ull) | ((int)((m2 >> nz) & 1ull) << 1);
}
// volume flux through the two median-dual segments of an edge, oce_adv_tra_hor.F90:139,157,176
__device__ __forceinline__ double seg_vflux(const KP& p, int segs, int be1, int be2, double dX1, double dY1, double dX2, double dY2, int nz) {
    double c1 = 0.0, c2 = 0.0;
    if (segs & 1) { const double2 uv = AT(p.UV, be1, nz); c1 = (-uv.y * dX1 + uv.x * dY1) * AT(p.helem, be1, nz); }
    if (segs & 2) { const double2 uv = AT(p.UV, be2, nz); c2 = (uv.y * dX2 - uv.x * dY2) * AT(p.helem, be2, nz); }
    if (segs == 3) return c1 + c2;
    return (segs & 1) ? c1 : c2;
}
__device__ __forceinline__ double upw1_flux(double t1, double t2, double vflux) {     // oce_adv_tra_hor.F90:143-146
    return -0.5 * (t1 * (vflux + fabs(vflux)) + t2 * (vflux - fabs(vflux)));
}


// ---------------------------------------------------------------------------------------------------
// vertical fluxes at interface nz of one column (flux_in = 0 variants; the caller subtracts)
__device__ __forceinline__ double ver_upw1(const double* __restrict__ T, const double* __restrict__ W, const double* __restrict__ area,
                                           int b, int nz, int nzmin, int nzmax) {                 // oce_adv_tra_ver.F90:255-281
    if (nz == nzmax) return 0.0;
    const double w = AT(W, b, nz);
    if (nz == nzmin) return -w * AT(T, b, nz) * AT(area, b, nz);
    return -0.5 * (AT(T, b, nz) * (w + fabs(w)) + AT(T, b, nz - 1) * (w - fabs(w))) * AT(area, b, nz);
}
__device__ __forceinline__ double ver_qr4c(const KP& p, const double* __restrict__ T, const double* __restrict__ W,
                                           int b, int nz, int nzmin, int nzmax) {                 // oce_adv_tra_ver.F90:314-356
    if (nz == nzmax) return 0.0;
    const double w = AT(W, b, nz), ar = AT(p.area, b, nz);
    if (nz == nzmin) return -AT(T, b, nz) * w * ar;
    if (nz == nzmin + 1 || nz == nzmax - 1) return -0.5 * (AT(T, b, nz - 1) + AT(T, b, nz)) * w * ar;
    const double t0 = AT(T, b, nz), tm1 = AT(T, b, nz - 1), tm2 = AT(T, b, nz - 2), tp1 = AT(T, b, nz + 1);
    const double z0 = AT(p.Z3, b, nz), zm1 = AT(p.Z3, b, nz - 1), zm2 = AT(p.Z3, b, nz - 2), zp1 = AT(p.Z3, b, nz + 1);
    const double zb = AT(p.zbar3, b, nz);
    const double qc = (tm1 - t0) / (zm1 - z0);
    const double qu = (t0 - tp1) / (z0 - zp1);
    const double qd = (tm2 - tm1) / (zm2 - zm1);
    const double Tmean1 = t0 + (2 * qc + qu) * (zb - z0) / 3.0;
    const double Tmean2 = tm1 + (2 * qc + qd) * (zb - zm1) / 3.0;
    const double Tmean = (w + fabs(w)) * Tmean1 + (w - fabs(w)) * Tmean2;
    return (-0.5 * (1.0 - p.pv) * Tmean - p.pv * (0.5 * (Tmean1 + Tmean2)) * w) * ar;
}
__device__ __forceinline__ double ver_cdiff(const KP& p, const double* __restrict__ T, const double* __restrict__ W,
                                            int b, int nz, int nzmin) {                           // oce_adv_tra_ver.F90:566-585
    const double w = AT(W, b, nz), ar = AT(p.area, b, nz);
    if (nz == nzmin) return -w * AT(T, b, nz) * ar;
    const double tv = 0.5 * (AT(T, b, nz - 1) + AT(T, b, nz));
    return -tv * w * ar;
}

// PPM (Colella & Woodward) vertical flux, adv_tra_vert_ppm, oce_adv_tra_ver.F90:361-538.  The reference walks down the
// column keeping the interface values tv(:) and writes tvert(nz) from the cell the flow leaves; here every interface
// is evaluated on its own: ppm_iface() is tv(k) with the reference's assignment order deciding which rule wins on short
// columns, ver_ppm() the flux through interface nz from its upwind cell.
static __device__ __noinline__ double ppm_iface(const double* __restrict__ T, const double* __restrict__ hn, const double* __restrict__ W,
                                         int b, int k, int nzmin, int nzmax) {
    if (k == nzmax) return AT(T, b, nzmax - 1);                                                        // :411
    if (k == nzmax - 1) {                                                                              // :408-409
        const double s = copysign(1.0, AT(W, b, nzmax - 1));
        return -AT(T, b, max(nzmax - 2, 1)) * dmin(s, 0.0) + AT(T, b, nzmax - 1) * dmax(s, 0.0);
    }
    if (k == nzmin) return AT(T, b, nzmin);                                                            // :401
    if (k == nzmin + 1) return 0.5 * (AT(T, b, nzmin) + AT(T, b, nzmin + 1));                          // :404
    const int nz = k - 1;                                                                              // :416-456
    const double tm1 = AT(T, b, nz - 1), t0 = AT(T, b, nz), tp1 = AT(T, b, nz + 1), tp2 = AT(T, b, nz + 2);
    const double dzjm1 = AT(hn, b, nz - 1), dzj = AT(hn, b, nz), dzjp1 = AT(hn, b, nz + 1), dzjp2 = AT(hn, b, nz + 2);
    double deltaj = dzj / (dzjm1 + dzj + dzjp1) *
                    ((2.0 * dzjm1 + dzj) / (dzjp1 + dzj) * (tp1 - t0) + (dzj + 2.0 * dzjp1) / (dzjm1 + dzj) * (t0 - tm1));
    double deltajp1 = dzjp1 / (dzj + dzjp1 + dzjp2) *
                      ((2.0 * dzj + dzjp1) / (dzjp2 + dzjp1) * (tp2 - tp1) + (dzjp1 + 2.0 * dzjp2) / (dzj + dzjp1) * (tp1 - t0));
    if ((tp1 - t0) * (t0 - tm1) > 0.0)
        deltaj = copysign(dmin(dmin(fabs(deltaj), 2.0 * fabs(tp1 - t0)), 2.0 * fabs(t0 - tm1)), deltaj);
    else deltaj = 0.0;
    if ((tp2 - tp1) * (tp1 - t0) > 0.0)
        deltajp1 = copysign(dmin(dmin(fabs(deltajp1), 2.0 * fabs(tp2 - tp1)), 2.0 * fabs(tp1 - t0)), deltajp1);
    else deltajp1 = 0.0;
    return t0 + dzj / (dzj + dzjp1) * (tp1 - t0) +
           1.0 / (dzjm1 + dzj + dzjp1 + dzjp2) *
               ((2.0 * dzjp1 * dzj) / (dzj + dzjp1) * ((dzjm1 + dzj) / (2.0 * dzj + dzjp1) - (dzjp2 + dzjp1) / (2.0 * dzjp1 + dzj)) * (tp1 - t0) -
                dzj * (dzjm1 + dzj) / (2.0 * dzj + dzjp1) * deltajp1 + dzjp1 * (dzjp1 + dzjp2) / (dzj + 2.0 * dzjp1) * deltaj);
}
static __device__ __noinline__ double ver_ppm(const KP& p, const double* __restrict__ T, const double* __restrict__ W,
                                       int b, int nz, int nzmin, int nzmax) {
    if (nz == nzmax) return 0.0;                                                                       // :527
    const double w = AT(W, b, nz), ar = AT(p.area, b, nz);
    if (nz == nzmin) return -ppm_iface(T, p.hnode_new, W, b, nzmin, nzmin, nzmax) * w * ar;            // :525
    if (w == 0.0) return 0.0;
    const int j = (w > 0.0) ? nz : nz - 1;                                                             // the cell the flow leaves (:461-519)
    const double tj = AT(T, b, j);
    double aL = ppm_iface(T, p.hnode_new, W, b, j, nzmin, nzmax), aR = ppm_iface(T, p.hnode_new, W, b, j + 1, nzmin, nzmax);
    if ((aR - tj) * (tj - aL) <= 0.0) { aL = tj; aR = tj; }
    if ((aR - aL) * (tj - 0.5 * (aL + aR)) > (aR - aL) * (aR - aL) / 6.0) aL = 3.0 * tj - 2.0 * aR;
    if ((aR - aL) * (tj - 0.5 * (aR + aL)) < -((aR - aL) * (aR - aL)) / 6.0) aR = 3.0 * tj - 2.0 * aL;
    const double dzj = AT(p.hnode, b, j);
    const double aj = 6.0 * (tj - 0.5 * (aL + aR));
    double f;
    if (w > 0.0) {
        const double x = dmin(w * p.dt / dzj, 1.0);
        f = (-aL - 0.5 * x * (aR - aL + (1.0 - 2.0 / 3.0 * x) * aj));
    } else {
        const double x = dmin(-w * p.dt / dzj, 1.0);
        f = (-aR + 0.5 * x * (aR - aL - (1.0 - 2.0 / 3.0 * x) * aj));
    }
    return f * ar * w;
}


// ---------------------------------------------------------------------------------------------------
// area-weighted mean of the elemental gradients around a node (fill_up_dn_grad fallback, oce_muscl_adv.F90:313-328)
static __device__ __noinline__ void node_mean_grad(const int* __restrict__ nie_slot, int maxnie, const unsigned* __restrict__ el_lev,
                                            const double* __restrict__ el_area, const double2* __restrict__ txy,
                                            int nlp, int node, int nz, double& gx, double& gy) {
    double tvol = 0.0, tx = 0.0, ty = 0.0;
    const int* slot = nie_slot + (size_t)(node >> 5) * maxnie * TILE + (node & 31);
    for (int k = 0; k < maxnie; k++) {
        const int e = slot[k * TILE];
        if (e < 0) break;
        const unsigned lv = el_lev[e];
        const int nu = lv & 255, nlm = (lv >> 8) & 255;
        if (nlm < nz || nz < nu) continue;
        const double ar = el_area[e];
        const int be = tbase(e, nlp);
        const double2 g = AT(txy, be, nz);
        tvol = tvol + ar;
        tx = tx + g.x * ar;
        ty = ty + g.y * ar;
    }
    // tvol == 0: nz lies outside this node's own level range (above its cavity top or below its bottom).  The reference
    // does not compute edge_up_dn_grad there (oce_muscl_adv.F90:313,337: the loops start at ulevels_nod2D(node)) and the
    // entry only ever meets a zero volume flux (no element of the node is wet at nz); 0/0 would turn that 0 into NaN
    if (tvol == 0.0) { gx = 0.0; gy = 0.0; return; }
    gx = tx / tvol; gy = ty / tvol;
}


// diffusive flux through interface z of diff_ver_part_expl_ale (oce_ale_tracer.F90:354-380): surface flux at the top,
// Kv * dT/dz * area below, zero under the last layer
__device__ __forceinline__ double expl_vd2(const KP& p, int tr_num, const double* __restrict__ T, int n, int b, int z, int ul1, int nl1) {
    if (z > nl1) return 0.0;
    if (z == ul1) {
        if (tr_num == 1) return -p.heat_flux[n] / 4.2e6;
        if (tr_num == 2) return p.virtual_salt[n] + p.relax_salt[n] - p.real_salt_flux[n] * p.is_nonlinfs;
        return 0.0;
    }
    const double zinv1 = 1.0 / (AT(p.Z3, b, z - 1) - AT(p.Z3, b, z));
    return (AT(p.Kv, b, z) + 0.0) * (AT(T, b, z - 1) - AT(T, b, z)) * zinv1 * AT(p.area, b, z);
}
__device__ __forceinline__ double expl_vd(const KP& p, const double* __restrict__ T, int n, int b, int z, int ul1, int nl1) { return expl_vd2(p, p.tr_num, T, n, b, z, ul1, nl1); }
