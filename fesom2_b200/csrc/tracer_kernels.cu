// Tracer transport kernels (FP64, sm_100a): AB2, elemental gradients, low-order solution, high-order
// antidiffusive fluxes, Zalesak limiter, update, horizontal diffusion, implicit vertical diffusion.
//
// All edge->node scatters of the reference are node-centric gathers here: each own node walks its incident
// edges in ascending reference edge id, which is the order in which the reference's edge loops add into that
// node, so sums are reproduced term by term without atomics.  Arithmetic follows the Fortran expression order
// (compiled with --fmad=false) — see the file:line notes at each kernel.
#include "common.cuh"

#define LB(n) __launch_bounds__(n)

struct KP {                       // kernel parameter block (by value)
    int nl, nlp, myN, N, myE, Eall, myEd, maxdeg, maxnie;
    double dt, ph, pv, eps_ab, is_nonlinfs;
    int hor, ver, w_split, Redi, use_sw_pene, do_hor_diff, tracer_id, tr_num;
    const EdgeRec* edge; const EdgeHo* edgeho; const int* ne_slot; const int* nie_slot; const int* el_nb;
    const unsigned* el_lev; const double* el_g; const double* el_area; const unsigned* n_lev; const int* n_nbl; const double* n_znb;
    const double *area, *areasvol, *U, *V, *helem, *W, *We, *Wi, *hnode, *hnode_new, *Z3, *zbar3, *Kv, *Ki, *sl1, *sl2, *sl3, *sw3d;
    const double *heat_flux, *water_flux, *virtual_salt, *relax_salt, *real_salt_flux;
    double *T, *Told, *T1;        // this tracer's fields; T1 = tracer 1 (bc_surface)
    double *txy_x, *txy_y, *trz, *LO, *adf_h, *adf_v, *fplus, *fminus, *tstar;
};

__device__ __forceinline__ int tbase(int ent, int nlp) { return (ent >> 5) * (nlp * TILE) + (ent & 31); }
#define AT(p, base, nz) (p)[(base) + (((nz) - 1) << 5)]          // 1-based level

struct ELv { int nu1, nl1, nu2, nl2, nl12, nu12, lo, hi; };
__device__ __forceinline__ ELv unpack_edge_levels(unsigned lev) {
    ELv v; v.nu1 = lev & 255; v.nl1 = (lev >> 8) & 255; v.nu2 = (lev >> 16) & 255; v.nl2 = lev >> 24;
    v.nl12 = min(v.nl1, v.nl2); v.nu12 = max(v.nu1, v.nu2);
    v.lo = v.nu2 > 0 ? min(v.nu1, v.nu2) : v.nu1; v.hi = max(v.nl1, v.nl2);
    return v;
}
// which of the ranges (A)-(E) of oce_adv_tra_hor.F90:129-209 holds nz: bit0 = segment of el(1), bit1 = segment of el(2)
__device__ __forceinline__ int edge_segs(const ELv& v, int nz) {
    const bool inA = nz >= v.nu1 && nz <= v.nu12 - 1;
    const bool inB = v.nu2 > 0 && nz >= v.nu2 && nz <= v.nu12 - 1;
    const bool inC = nz >= v.nu12 && nz <= v.nl12;
    const bool inD = nz >= v.nl12 + 1 && nz <= v.nl1;
    const bool inE = nz >= v.nl12 + 1 && nz <= v.nl2;
    return ((inA | inC | inD) ? 1 : 0) | ((inB | inC | inE) ? 2 : 0);
}
// volume flux through the two median-dual segments of an edge, oce_adv_tra_hor.F90:139,157,176
__device__ __forceinline__ double edge_vflux(const KP& p, const EdgeRec& r, int segs, int nz) {
    double c1 = 0.0, c2 = 0.0;
    if (segs & 1) c1 = (-AT(p.V, r.be1, nz) * r.dX1 + AT(p.U, r.be1, nz) * r.dY1) * AT(p.helem, r.be1, nz);
    if (segs & 2) c2 = (AT(p.V, r.be2, nz) * r.dX2 - AT(p.U, r.be2, nz) * r.dY2) * AT(p.helem, r.be2, nz);
    if (segs == 3) return c1 + c2;
    return (segs & 1) ? c1 : c2;
}
__device__ __forceinline__ double upw1_flux(double t1, double t2, double vflux) {     // oce_adv_tra_hor.F90:143-146
    return -0.5 * (t1 * (vflux + fabs(vflux)) + t2 * (vflux - fabs(vflux)));
}

// ---------------------------------------------------------------------------------------------------
// K1: Adams-Bashforth extrapolation, oce_tracer_mod.F90:64 — over every entry (halo and padding included)
__global__ void LB(256) k_ab2(double* __restrict__ told, const double* __restrict__ t, size_t count, double ca, double cb) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < count; i += stride) told[i] = ca * told[i] + cb * t[i];
}

// K2: tracer_gradient_elements, oce_tracer_mod.F90:35-44 — one thread per own element
__global__ void LB(128) k_grad_elem(KP p, const double* __restrict__ ttf) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= p.myE) return;
    const int b1 = p.el_nb[e], b2 = p.el_nb[p.Eall + e], b3 = p.el_nb[2 * p.Eall + e];
    const double g1 = p.el_g[e], g2 = p.el_g[p.myE + e], g3 = p.el_g[2 * p.myE + e];
    const double g4 = p.el_g[3 * p.myE + e], g5 = p.el_g[4 * p.myE + e], g6 = p.el_g[5 * p.myE + e];
    const unsigned lv = p.el_lev[e];
    const int nu = lv & 255, nlm = (lv >> 8) & 255;
    const int be = tbase(e, p.nlp);
    for (int nz = nu; nz <= nlm; nz++) {
        const double t1 = AT(ttf, b1, nz), t2 = AT(ttf, b2, nz), t3 = AT(ttf, b3, nz);
        AT(p.txy_x, be, nz) = g1 * t1 + g2 * t2 + g3 * t3;
        AT(p.txy_y, be, nz) = g4 * t1 + g5 * t2 + g6 * t3;
    }
}

// tracer_gradient_z, oce_tracer_mod.F90:139-152 — all local nodes
__global__ void LB(128) k_grad_z(KP p, const double* __restrict__ ttf) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= p.N) return;
    const unsigned lv = p.n_lev[n];
    const int nzmin = lv & 255, nzmax = (lv >> 8) & 255;
    const int b = tbase(n, p.nlp);
    for (int nz = nzmin + 1; nz <= nzmax - 1; nz++) {
        const double dz = 0.5 * (AT(p.hnode_new, b, nz - 1) + AT(p.hnode_new, b, nz));
        AT(p.trz, b, nz) = (AT(ttf, b, nz - 1) - AT(ttf, b, nz)) / dz;
    }
    AT(p.trz, b, nzmin) = 0.0;
    AT(p.trz, b, nzmax) = 0.0;
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

// K3: low-order solution + vertical antidiffusive flux.  One thread per own node.
//   fct_LO  : oce_adv_tra_driver.F90:83-122 (adv_tra_hor_upw1 + edge scatter + adv_tra_ver_upw1 + column update)
//   adf_v   : oce_adv_tra_driver.F90:163-179 with init_zero=.false.  (HO(ttfAB, W) - LO(ttf, W))
// dynamic smem: acc[nlp][blockDim]
__global__ void LB(128) k_lo(KP p) {
    extern __shared__ double acc[];
    const int tid = threadIdx.x, BS = blockDim.x;
    const int n = blockIdx.x * BS + tid;
    const bool valid = n < p.myN;
    for (int k = 0; k < p.nlp; k++) acc[k * BS + tid] = 0.0;
    if (!valid) return;
    const int b = tbase(n, p.nlp);
    const unsigned lv = p.n_lev[n];
    const int nzmin = lv & 255, nzmax = (lv >> 8) & 255;
    const double* __restrict__ T = p.T;
    const int* slot = p.ne_slot + (size_t)(n >> 5) * p.maxdeg * TILE + (n & 31);
    for (int j = 0; j < p.maxdeg; j++) {
        const int s = slot[j * TILE];
        if (s < 0) break;
        const EdgeRec r = p.edge[s >> 1];
        const bool second = s & 1;
        const ELv v = unpack_edge_levels(r.lev);
        for (int nz = v.lo; nz <= v.hi; nz++) {
            const int segs = edge_segs(v, nz);
            if (!segs) continue;
            const double vflux = edge_vflux(p, r, segs, nz);
            const double f = upw1_flux(AT(T, r.bn1, nz), AT(T, r.bn2, nz), vflux);
            double a = acc[(nz - 1) * BS + tid];
            a = second ? a - f : a + f;
            acc[(nz - 1) * BS + tid] = a;
        }
    }
    const double* __restrict__ Wlo = p.We;                         // explicit part for the LO solution
    const double* __restrict__ Wad = p.w_split ? p.W : p.We;       // LO part of the antidiffusive flux: full W (:131)
    double fv = ver_upw1(T, Wlo, p.area, b, nzmin, nzmin, nzmax);
    for (int nz = nzmin; nz <= nzmax - 1; nz++) {
        const double fvn = ver_upw1(T, Wlo, p.area, b, nz + 1, nzmin, nzmax);
        const double lo = (AT(T, b, nz) * AT(p.hnode, b, nz) + (acc[(nz - 1) * BS + tid] + (fv - fvn)) * p.dt / AT(p.areasvol, b, nz)) / AT(p.hnode_new, b, nz);
        AT(p.LO, b, nz) = lo;
        fv = fvn;
    }
    const double* __restrict__ TAB = p.Told;
    for (int nz = nzmin; nz <= nzmax; nz++) {
        const double flo = ver_upw1(T, Wad, p.area, b, nz, nzmin, nzmax);
        double ho;
        if (p.ver == 0) ho = ver_qr4c(p, TAB, p.W, b, nz, nzmin, nzmax) - flo;
        else if (p.ver == 1) ho = (nz == nzmax) ? flo : ver_cdiff(p, TAB, p.W, b, nz, nzmin) - flo;
        else ho = ver_upw1(TAB, p.W, p.area, b, nz, nzmin, nzmax) - flo;
        AT(p.adf_v, b, nz) = ho;
    }
}

// Thomas algorithm shared by the two implicit column solves; cp/tp live in shared memory [nlp][BS]
// (oce_ale_tracer.F90:828-844 / oce_adv_tra_ver.F90:201-219)

// K3b: adv_tra_vert_impl (oce_adv_tra_ver.F90:83-227) applied to fct_LO when w_split.  smem: cp,tp [2][nlp][BS]
__global__ void LB(64) k_vert_impl(KP p, double* __restrict__ ttf, const double* __restrict__ W) {
    extern __shared__ double sm[];
    const int tid = threadIdx.x, BS = blockDim.x;
    double* cp = sm; double* tp = sm + (size_t)p.nlp * BS;
    const int n = blockIdx.x * BS + tid;
    if (n >= p.myN) return;
    const int b = tbase(n, p.nlp);
    const unsigned lv = p.n_lev[n];
    const int nzmin = lv & 255, nzmax = (lv >> 8) & 255;
    const double zinv = 1.0 * p.dt;
    double cpm = 0.0, tpm = 0.0;
    for (int nz = nzmin; nz <= nzmax - 1; nz++) {
        double a, bb, c, v_adv;
        const double hn = AT(p.hnode_new, b, nz), asv = AT(p.areasvol, b, nz);
        if (nz == nzmin) {
            a = 0.0;
            v_adv = zinv * AT(p.area, b, nz) / asv;
            bb = hn + AT(W, b, nz) * v_adv;
            v_adv = zinv * AT(p.area, b, nz + 1) / asv;
            bb = bb - fmin(0.0, AT(W, b, nz + 1)) * v_adv;
            c = -fmax(0.0, AT(W, b, nz + 1)) * v_adv;
        } else if (nz <= nzmax - 2) {
            v_adv = zinv * AT(p.area, b, nz) / asv;
            a = fmin(0.0, AT(W, b, nz)) * v_adv;
            bb = hn + fmax(0.0, AT(W, b, nz)) * v_adv;
            v_adv = zinv * AT(p.area, b, nz + 1) / asv;
            bb = bb - fmin(0.0, AT(W, b, nz + 1)) * v_adv;
            c = -fmax(0.0, AT(W, b, nz + 1)) * v_adv;
        } else {
            v_adv = zinv * AT(p.area, b, nz) / asv;
            a = fmin(0.0, AT(W, b, nz)) * v_adv;
            bb = hn + fmax(0.0, AT(W, b, nz)) * v_adv;
            c = 0.0;
        }
        double tr;
        const double dz = hn;
        if (nz == nzmin) tr = -(bb - dz) * AT(ttf, b, nz) - c * AT(ttf, b, nz + 1);
        else if (nz <= nzmax - 2) tr = -a * AT(ttf, b, nz - 1) - (bb - dz) * AT(ttf, b, nz) - c * AT(ttf, b, nz + 1);
        else tr = -a * AT(ttf, b, nz - 1) - (bb - dz) * AT(ttf, b, nz);
        if (nz == nzmin) { cpm = c / bb; tpm = tr / bb; }
        else { const double m = bb - cpm * a; cpm = c / m; tpm = (tr - tpm * a) / m; }
        cp[(nz - 1) * BS + tid] = cpm; tp[(nz - 1) * BS + tid] = tpm;
    }
    double x = tp[(nzmax - 2) * BS + tid];
    AT(ttf, b, nzmax - 1) = AT(ttf, b, nzmax - 1) + x;
    for (int nz = nzmax - 2; nz >= nzmin; nz--) {
        x = tp[(nz - 1) * BS + tid] - cp[(nz - 1) * BS + tid] * x;
        AT(ttf, b, nz) = AT(ttf, b, nz) + x;
    }
}

// ---------------------------------------------------------------------------------------------------
// area-weighted mean of the elemental gradients around a node (fill_up_dn_grad fallback, oce_muscl_adv.F90:313-328)
__device__ __noinline__ void node_mean_grad(const KP& p, int node, int nz, double& gx, double& gy) {
    double tvol = 0.0, tx = 0.0, ty = 0.0;
    const int* slot = p.nie_slot + (size_t)(node >> 5) * p.maxnie * TILE + (node & 31);
    for (int k = 0; k < p.maxnie; k++) {
        const int e = slot[k * TILE];
        if (e < 0) break;
        const unsigned lv = p.el_lev[e];
        const int nu = lv & 255, nlm = (lv >> 8) & 255;
        if (nlm < nz || nz < nu) continue;
        const double ar = p.el_area[e];
        const int be = tbase(e, p.nlp);
        tvol = tvol + ar;
        tx = tx + AT(p.txy_x, be, nz) * ar;
        ty = ty + AT(p.txy_y, be, nz) * ar;
    }
    gx = tx / tvol; gy = ty / tvol;
}

// K4: high-order horizontal flux minus low-order flux = antidiffusive flux adf_h.  One thread per own edge.
//   adv_tra_hor_mfct / _muscl / _upw1 with init_zero=.false. (oce_adv_tra_hor.F90:215-733) on ttfAB,
//   the input flux being adv_tra_hor_upw1(ttf) (oce_adv_tra_driver.F90:83,142-149); the gradients of
//   fill_up_dn_grad (oce_muscl_adv.F90:285-447) are gathered from txy_x/txy_y instead of being stored.
__global__ void LB(128) k_ho_edge(KP p) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= p.myEd) return;
    const EdgeRec r = p.edge[e];
    const EdgeHo h = p.edgeho[e];
    const ELv v = unpack_edge_levels(r.lev);
    const int be = tbase(e, p.nlp);
    const int sh_lo = h.shared & 255, sh_hi = (h.shared >> 8) & 255;
    const bool good = (h.shared >> 16) & 1;
    const int nbl1 = p.n_nbl[h.n1], nbl2 = p.n_nbl[h.n2];
    const double* __restrict__ T = p.T;
    const double* __restrict__ A = p.Told;
    for (int nz = v.lo; nz <= v.hi; nz++) {
        const int segs = edge_segs(v, nz);
        if (!segs) continue;
        const double vflux = edge_vflux(p, r, segs, nz);
        const double flo = upw1_flux(AT(T, r.bn1, nz), AT(T, r.bn2, nz), vflux);
        const double a1 = AT(A, r.bn1, nz), a2 = AT(A, r.bn2, nz);
        double fho;
        if (p.hor == 2) fho = upw1_flux(a1, a2, vflux);
        else {
            double g1, g2, g3, g4;         // edge_up_dn_grad(1:4): x up, x down, y up, y down
            if (good && nz >= sh_lo && nz <= sh_hi - 1) {
                g1 = AT(p.txy_x, h.bup, nz); g2 = AT(p.txy_x, h.bdn, nz);
                g3 = AT(p.txy_y, h.bup, nz); g4 = AT(p.txy_y, h.bdn, nz);
            } else {
                node_mean_grad(p, h.n1, nz, g1, g3);
                node_mean_grad(p, h.n2, nz, g2, g4);
            }
            double Tmean2 = a2 - (2.0 * (a2 - a1) + h.dxa * g2 + h.dyr * g4) / 6.0;
            double Tmean1 = a1 + (2.0 * (a2 - a1) + h.dxa * g1 + h.dyr * g3) / 6.0;
            if (p.hor == 0) {              // MUSCL: c_lo, oce_adv_tra_hor.F90:352-353
                const double c_lo1 = (nbl1 - nz) >= 0 ? 1.0 : 0.0, c_lo2 = (nbl2 - nz) >= 0 ? 1.0 : 0.0;
                Tmean2 = a2 - (2.0 * (a2 - a1) + h.dxa * g2 + h.dyr * g4) / 6.0 * c_lo2;
                Tmean1 = a1 + (2.0 * (a2 - a1) + h.dxa * g1 + h.dyr * g3) / 6.0 * c_lo1;
            }
            const double cHO = (vflux + fabs(vflux)) * Tmean1 + (vflux - fabs(vflux)) * Tmean2;
            fho = -0.5 * (1.0 - p.ph) * cHO - vflux * p.ph * (0.5 * (Tmean1 + Tmean2));
        }
        AT(p.adf_h, be, nz) = fho - flo;
    }
}

// ---------------------------------------------------------------------------------------------------
// K5: Zalesak limiter, parts a1-a3 and b1-b2 of oce_tra_adv_fct (oce_adv_tra_fct.F90:93-276) for one own node
// per thread.  The element min/max pass (a2) is folded into the node pass: the bound of node n at level nz is
// the max/min over n and over every edge neighbour m whose edge has a wet adjacent element at nz — the same set
// of values the reference reaches through UV_rhs(1:2,nz,nod_in_elem2D(:,n)); "-bignumber/+bignumber" enter
// wherever a surrounding element is already below its bottom (nz >= nlevels_nod2D_min).
// dynamic smem: pp, pm, mx, mn  [4][nlp][BS]
__global__ void LB(64) k_fct_bounds(KP p) {
    extern __shared__ double sm[];
    const int tid = threadIdx.x, BS = blockDim.x;
    double* pp = sm; double* pm = sm + (size_t)p.nlp * BS; double* mx = pm + (size_t)p.nlp * BS; double* mn = mx + (size_t)p.nlp * BS;
    const int n = blockIdx.x * BS + tid;
    if (n >= p.myN) return;
    const double flux_eps = 1e-16, bignumber = 1e3;
    const int b = tbase(n, p.nlp);
    const unsigned lv = p.n_lev[n];
    const int nzmin = lv & 255, nzmax = (lv >> 8) & 255, nlmin = (lv >> 16) & 255;
    const double* __restrict__ T = p.T;
    const double* __restrict__ LO = p.LO;
    for (int nz = nzmin; nz <= nzmax - 1; nz++) {
        const double av = AT(p.adf_v, b, nz), avn = AT(p.adf_v, b, nz + 1);
        pp[(nz - 1) * BS + tid] = fmax(0.0, av) + fmax(0.0, -avn);          // :235-236
        pm[(nz - 1) * BS + tid] = fmin(0.0, av) + fmin(0.0, -avn);
        const double lo = AT(LO, b, nz), t = AT(T, b, nz);
        double M = fmax(lo, t), m = fmin(lo, t);                             // a1 :99-100
        if (nz >= nlmin) { M = fmax(M, -bignumber); m = fmin(m, bignumber); }   // a2 :116-121
        mx[(nz - 1) * BS + tid] = M; mn[(nz - 1) * BS + tid] = m;
    }
    const int* slot = p.ne_slot + (size_t)(n >> 5) * p.maxdeg * TILE + (n & 31);
    for (int j = 0; j < p.maxdeg; j++) {
        const int s = slot[j * TILE];
        if (s < 0) break;
        const int e = s >> 1;
        const bool second = s & 1;
        const EdgeRec r = p.edge[e];
        const ELv v = unpack_edge_levels(r.lev);
        const int be = tbase(e, p.nlp);
        const int bo = second ? r.bn1 : r.bn2;
        for (int nz = v.lo; nz <= v.hi; nz++) {
            double f = AT(p.adf_h, be, nz);
            if (second) f = -f;
            pp[(nz - 1) * BS + tid] = pp[(nz - 1) * BS + tid] + fmax(0.0, f);      // b1 :258-261
            pm[(nz - 1) * BS + tid] = pm[(nz - 1) * BS + tid] + fmin(0.0, f);
            if (edge_segs(v, nz)) {
                const double lo = AT(LO, bo, nz), t = AT(T, bo, nz);
                mx[(nz - 1) * BS + tid] = fmax(mx[(nz - 1) * BS + tid], fmax(lo, t));
                mn[(nz - 1) * BS + tid] = fmin(mn[(nz - 1) * BS + tid], fmin(lo, t));
            }
        }
    }
    for (int nz = nzmin; nz <= nzmax - 1; nz++) {
        double qmax = mx[(nz - 1) * BS + tid], qmin = mn[(nz - 1) * BS + tid];
        if (nz > nzmin && nz < nzmax - 1) {                                  // a3, vlimit=1 :153-156
            qmax = fmax(fmax(mx[(nz - 2) * BS + tid], qmax), mx[nz * BS + tid]);
            qmin = fmin(fmin(mn[(nz - 2) * BS + tid], qmin), mn[nz * BS + tid]);
        }
        const double lo = AT(LO, b, nz), asv = AT(p.areasvol, b, nz);
        qmax = qmax - lo; qmin = qmin - lo;
        double flux = pp[(nz - 1) * BS + tid] * p.dt / asv + flux_eps;       // b2 :271-274
        AT(p.fplus, b, nz) = fmin(1.0, qmax / flux);
        flux = pm[(nz - 1) * BS + tid] * p.dt / asv - flux_eps;
        AT(p.fminus, b, nz) = fmin(1.0, qmin / flux);
    }
}

// limited vertical flux at interface nz (b3 vertical, oce_adv_tra_fct.F90:284-311)
__device__ __forceinline__ double limited_fv(const KP& p, int b, int nz, int nzmin, int nzmax) {
    const double flux = AT(p.adf_v, b, nz);
    if (nz == nzmax) return flux;
    double ae = 1.0;
    if (nz == nzmin) {
        if (flux >= 0.0) ae = fmin(ae, AT(p.fplus, b, nz)); else ae = fmin(ae, AT(p.fminus, b, nz));
    } else {
        if (flux >= 0.0) { ae = fmin(ae, AT(p.fminus, b, nz - 1)); ae = fmin(ae, AT(p.fplus, b, nz)); }
        else { ae = fmin(ae, AT(p.fplus, b, nz - 1)); ae = fmin(ae, AT(p.fminus, b, nz)); }
    }
    return ae * flux;
}

// K7a: b3 limiting + oce_tra_adv_flux2dtracer (oce_adv_tra_driver.F90:223-268) + del_ttf assembly
// (oce_ale_tracer.F90:240) + diff_part_hor_redi (:934-1077) + T* update (:294-305).  One thread per own node.
// Writes tstar; tr_arr_old = tr_arr is done by the caller for all entries.  dynamic smem: acc[nlp][BS]
__global__ void LB(128) k_update(KP p) {
    extern __shared__ double acc[];
    const int tid = threadIdx.x, BS = blockDim.x;
    const int n = blockIdx.x * BS + tid;
    for (int k = 0; k < p.nlp; k++) acc[k * BS + tid] = 0.0;
    if (n >= p.myN) return;
    const int b = tbase(n, p.nlp);
    const unsigned lv = p.n_lev[n];
    const int nzmin = lv & 255, nzmax = (lv >> 8) & 255;
    const double* __restrict__ T = p.T;
    const int* slot = p.ne_slot + (size_t)(n >> 5) * p.maxdeg * TILE + (n & 31);
    for (int j = 0; j < p.maxdeg; j++) {                                     // dttf_h
        const int s = slot[j * TILE];
        if (s < 0) break;
        const int e = s >> 1;
        const bool second = s & 1;
        const EdgeRec r = p.edge[e];
        const ELv v = unpack_edge_levels(r.lev);
        const int be = tbase(e, p.nlp);
        for (int nz = v.lo; nz <= v.hi; nz++) {
            const double flux = AT(p.adf_h, be, nz);
            double ae = 1.0;
            if (flux >= 0.0) { ae = fmin(ae, AT(p.fplus, r.bn1, nz)); ae = fmin(ae, AT(p.fminus, r.bn2, nz)); }
            else { ae = fmin(ae, AT(p.fminus, r.bn1, nz)); ae = fmin(ae, AT(p.fplus, r.bn2, nz)); }
            const double fl = ae * flux;
            const double d = fl * p.dt / AT(p.areasvol, b, nz);
            double a = acc[(nz - 1) * BS + tid];
            a = second ? a - d : a + d;
            acc[(nz - 1) * BS + tid] = a;
        }
    }
    {                                                                         // dttf_v, then del_ttf = dttf_h + dttf_v
        double fv = limited_fv(p, b, nzmin, nzmin, nzmax);
        for (int nz = nzmin; nz <= nzmax - 1; nz++) {
            const double fvn = limited_fv(p, b, nz + 1, nzmin, nzmax);
            double dv = 0.0 - AT(T, b, nz) * AT(p.hnode, b, nz) + AT(p.LO, b, nz) * AT(p.hnode_new, b, nz);
            dv = dv + (fv - fvn) * p.dt / AT(p.areasvol, b, nz);
            acc[(nz - 1) * BS + tid] = acc[(nz - 1) * BS + tid] + dv;
            fv = fvn;
        }
    }
    if (p.do_hor_diff) {
        const double isredi = p.Redi ? 1.0 : 0.0;
        for (int j = 0; j < p.maxdeg; j++) {
            const int s = slot[j * TILE];
            if (s < 0) break;
            const bool second = s & 1;
            const EdgeRec r = p.edge[s >> 1];
            const ELv v = unpack_edge_levels(r.lev);
            for (int nz = v.lo; nz <= v.hi; nz++) {
                const int segs = edge_segs(v, nz);
                if (!segs) continue;
                const double Kh = (AT(p.Ki, r.bn1, nz) + AT(p.Ki, r.bn2, nz)) / 2.0;
                double SxTz = 0.0, SyTz = 0.0;
                if (p.Redi) {
                    const double Tz1 = 0.5 * (AT(p.trz, r.bn1, nz) + AT(p.trz, r.bn1, nz + 1));
                    const double Tz2 = 0.5 * (AT(p.trz, r.bn2, nz) + AT(p.trz, r.bn2, nz + 1));
                    SxTz = (Tz1 * AT(p.sl1, r.bn1, nz) + Tz2 * AT(p.sl1, r.bn2, nz)) / 2.0;
                    SyTz = (Tz1 * AT(p.sl2, r.bn1, nz) + Tz2 * AT(p.sl2, r.bn2, nz)) / 2.0;
                }
                double dz, Tx, Ty, c;
                if (segs == 3) {
                    dz = (AT(p.helem, r.be1, nz) + AT(p.helem, r.be2, nz)) / 2.0;
                    Tx = 0.5 * (AT(p.txy_x, r.be1, nz) + AT(p.txy_x, r.be2, nz));
                    Ty = 0.5 * (AT(p.txy_y, r.be1, nz) + AT(p.txy_y, r.be2, nz));
                    const double Fx = Kh * (Tx + SxTz * isredi), Fy = Kh * (Ty + SyTz * isredi);
                    c = ((r.dX2 - r.dX1) * Fy - (r.dY2 - r.dY1) * Fx) * dz;
                } else if (segs == 1) {
                    dz = AT(p.helem, r.be1, nz); Tx = AT(p.txy_x, r.be1, nz); Ty = AT(p.txy_y, r.be1, nz);
                    const double Fx = Kh * (Tx + SxTz * isredi), Fy = Kh * (Ty + SyTz * isredi);
                    c = (-r.dX1 * Fy + r.dY1 * Fx) * dz;
                } else {
                    dz = AT(p.helem, r.be2, nz); Tx = AT(p.txy_x, r.be2, nz); Ty = AT(p.txy_y, r.be2, nz);
                    const double Fx = Kh * (Tx + SxTz * isredi), Fy = Kh * (Ty + SyTz * isredi);
                    c = (r.dX2 * Fy - r.dY2 * Fx) * dz;
                }
                const double rhs = second ? 0.0 - c : 0.0 + c;
                acc[(nz - 1) * BS + tid] = acc[(nz - 1) * BS + tid] + rhs * p.dt / AT(p.areasvol, b, nz);
            }
        }
    }
    for (int nz = nzmin; nz <= nzmax - 1; nz++) {                             // T*  :302-305
        const double t = AT(T, b, nz);
        const double del = acc[(nz - 1) * BS + tid] + t * (AT(p.hnode, b, nz) - AT(p.hnode_new, b, nz));
        AT(p.tstar, b, nz) = t + del / AT(p.hnode_new, b, nz);
    }
}

// ---------------------------------------------------------------------------------------------------
// K7b: diff_ver_part_impl_ale (oce_ale_tracer.F90:509-855): per own column rebuild zbar_n/Z_n from hnode_new,
// assemble a,b,c and the right-hand side from T*, add shortwave penetration and the surface boundary condition
// (bc_surface :1154-1195), Thomas algorithm, T = T* + dT.  One thread per column; dynamic smem: zb, cp, tp [3][nlp][BS]
__global__ void LB(64) k_impl_diff(KP p) {
    extern __shared__ double sm[];
    const int tid = threadIdx.x, BS = blockDim.x;
    double* zb = sm; double* cp = sm + (size_t)p.nlp * BS; double* tp = cp + (size_t)p.nlp * BS;
    const int n = blockIdx.x * BS + tid;
    if (n >= p.myN) return;
    const int b = tbase(n, p.nlp);
    const unsigned lv = p.n_lev[n];
    const int nzmin = lv & 255, nzmax = (lv >> 8) & 255;
    const double* __restrict__ hn = p.hnode_new;
    const double* __restrict__ TS = p.tstar;
    const double dt = p.dt;
    const double isredi = p.Redi ? 1.0 : 0.0;
    const bool do_wimpl = false;      // FCT limiter on => no implicit w here (:424)
#define ZB(nz) zb[((nz) - 1) * BS + tid]
#define ZN(nz) (ZB((nz) + 1) + AT(hn, b, nz) / 2.0)                  /* Z_n(nz) = zbar_n(nz+1) + hnode_new(nz)/2 */
    ZB(nzmax) = p.n_znb[n];
    for (int nz = nzmax - 1; nz >= nzmin; nz--) ZB(nz) = ZB(nz + 1) + AT(hn, b, nz);
    const double vcpw = 4.2e6;
    double cpm = 0.0, tpm = 0.0, zinv1 = 0.0;
    const double zinv = 1.0 * dt;
    for (int nz = nzmin; nz <= nzmax - 1; nz++) {
        double a, bb, c;
        const double asv = AT(p.areasvol, b, nz), h = AT(hn, b, nz);
        if (nz == nzmin) {
            const double zinv2 = 1.0 / (ZN(nz) - ZN(nz + 1));
            double Ty1 = 0.0;
            if (p.Redi) {
                const double s0 = AT(p.sl3, b, nz), s1 = AT(p.sl3, b, nz + 1);
                Ty1 = (ZN(nz) - ZB(nz + 1)) * zinv2 * (s0 * s0) * AT(p.Ki, b, nz) + (ZB(nz + 1) - ZN(nz + 1)) * zinv2 * (s1 * s1) * AT(p.Ki, b, nz + 1);
            }
            Ty1 = Ty1 * isredi;
            a = 0.0;
            c = -(AT(p.Kv, b, nz + 1) + Ty1) * zinv2 * zinv * AT(p.area, b, nz + 1) / asv;
            bb = -c + h;
            zinv1 = zinv2;
        } else if (nz <= nzmax - 2) {
            const double zinv2 = 1.0 / (ZN(nz) - ZN(nz + 1));
            double Ty = 0.0, Ty1 = 0.0;
            if (p.Redi) {
                const double sm1 = AT(p.sl3, b, nz - 1), s0 = AT(p.sl3, b, nz), s1 = AT(p.sl3, b, nz + 1);
                Ty = (ZN(nz - 1) - ZB(nz)) * zinv1 * (sm1 * sm1) * AT(p.Ki, b, nz - 1) + (ZB(nz) - ZN(nz)) * zinv1 * (s0 * s0) * AT(p.Ki, b, nz);
                Ty1 = (ZN(nz) - ZB(nz + 1)) * zinv2 * (s0 * s0) * AT(p.Ki, b, nz) + (ZB(nz + 1) - ZN(nz + 1)) * zinv2 * (s1 * s1) * AT(p.Ki, b, nz + 1);
            }
            Ty = Ty * isredi; Ty1 = Ty1 * isredi;
            a = -(AT(p.Kv, b, nz) + Ty) * zinv1 * zinv * (AT(p.area, b, nz) / asv);
            c = -(AT(p.Kv, b, nz + 1) + Ty1) * zinv2 * zinv * AT(p.area, b, nz + 1) / asv;
            bb = -a - c + h;
            zinv1 = zinv2;
        } else {
            double Ty = 0.0;
            if (p.Redi) {
                const double sm1 = AT(p.sl3, b, nz - 1), s0 = AT(p.sl3, b, nz);
                Ty = (ZN(nz - 1) - ZB(nz)) * zinv1 * (sm1 * sm1) * AT(p.Ki, b, nz - 1) + (ZB(nz) - ZN(nz)) * zinv1 * (s0 * s0) * AT(p.Ki, b, nz);
            }
            Ty = Ty * isredi;
            a = -(AT(p.Kv, b, nz) + Ty) * zinv1 * zinv * (AT(p.area, b, nz) / asv);
            c = 0.0;
            bb = -a + h;
        }
        (void)do_wimpl;
        double tr;
        if (nz == nzmin) tr = -(bb - h) * AT(TS, b, nz) - c * AT(TS, b, nz + 1);
        else if (nz <= nzmax - 2) tr = -a * AT(TS, b, nz - 1) - (bb - h) * AT(TS, b, nz) - c * AT(TS, b, nz + 1);
        else tr = -a * AT(TS, b, nz - 1) - (bb - h) * AT(TS, b, nz);
        if (p.use_sw_pene && p.tr_num == 1)
            tr = tr + (AT(p.sw3d, b, nz) - AT(p.sw3d, b, nz + 1) * AT(p.area, b, nz + 1) / asv) * zinv;
        if (nz == nzmin) {
            double bc = 0.0;
            if (p.tracer_id == 0) {
                const double tsurf = (p.tr_num == 1) ? AT(TS, b, nzmin) : AT(p.T1, b, nzmin);
                bc = -dt * (p.heat_flux[n] / vcpw + tsurf * p.water_flux[n] * p.is_nonlinfs);
            } else if (p.tracer_id == 1)
                bc = dt * (p.virtual_salt[n] + p.relax_salt[n] - p.real_salt_flux[n] * p.is_nonlinfs);
            tr = tr + bc;
        }
        if (nz == nzmin) { cpm = c / bb; tpm = tr / bb; }
        else { const double m = bb - cpm * a; cpm = c / m; tpm = (tr - tpm * a) / m; }
        cp[(nz - 1) * BS + tid] = cpm; tp[(nz - 1) * BS + tid] = tpm;
    }
    double x = tp[(nzmax - 2) * BS + tid];
    AT(p.T, b, nzmax - 1) = AT(TS, b, nzmax - 1) + x;
    for (int nz = nzmax - 2; nz >= nzmin; nz--) {
        x = tp[(nz - 1) * BS + tid] - cp[(nz - 1) * BS + tid] * x;
        AT(p.T, b, nz) = AT(TS, b, nz) + x;
    }
#undef ZB
#undef ZN
}

// explicit vertical diffusion variant is not on the default path (i_vert_diff=.true.); see DESIGN.md

// copy own columns of src into dst for levels nzmin..nzmax-1 (used when i_vert_diff is off)
__global__ void LB(128) k_copy_cols(KP p, double* __restrict__ dst, const double* __restrict__ src) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= p.myN) return;
    const int b = tbase(n, p.nlp);
    const unsigned lv = p.n_lev[n];
    for (int nz = (lv & 255); nz <= (int)((lv >> 8) & 255) - 1; nz++) AT(dst, b, nz) = AT(src, b, nz);
}

// salinity clamp, oce_ale_tracer.F90:176-198 — all local nodes
__global__ void LB(128) k_clamp_salt(KP p, double* __restrict__ S) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= p.N) return;
    const int b = tbase(n, p.nlp);
    const unsigned lv = p.n_lev[n];
    for (int nz = (lv & 255); nz <= (int)((lv >> 8) & 255) - 1; nz++) {
        double s = AT(S, b, nz);
        if (s > 45.0) s = 45.0;
        if (s < 3.0) s = 3.0;
        AT(S, b, nz) = s;
    }
}

// ---------------------------------------------------------------------------------------------------
static KP make_kp(TracerCtx& c, int tr) {
    KP p;
    p.nl = c.nl; p.nlp = c.nlp; p.myN = c.myN; p.N = c.N; p.myE = c.myE; p.Eall = c.Eall; p.myEd = c.myEd; p.maxdeg = c.maxdeg; p.maxnie = c.maxnie;
    p.dt = c.opt.dt; p.ph = c.opt.ph; p.pv = c.opt.pv; p.eps_ab = c.opt.epsilon; p.is_nonlinfs = c.opt.is_nonlinfs;
    p.hor = c.opt.tra_adv_hor; p.ver = c.opt.tra_adv_ver; p.w_split = c.opt.w_split; p.Redi = c.opt.Redi; p.use_sw_pene = c.opt.use_sw_pene;
    p.do_hor_diff = c.opt.do_hor_diff; p.tracer_id = c.opt.tracer_id[tr]; p.tr_num = tr + 1;
    p.edge = c.edge.p; p.edgeho = c.edgeho.p; p.ne_slot = c.ne_slot.p; p.nie_slot = c.nie_slot.p; p.el_nb = c.el_nb.p; p.el_lev = c.el_lev.p;
    p.el_g = c.el_g.p; p.el_area = c.el_area.p; p.n_lev = c.n_lev.p; p.n_nbl = c.n_nbl.p; p.n_znb = c.n_znb.p;
    p.area = c.area.p; p.areasvol = c.areasvol.p; p.U = c.U.p; p.V = c.V.p; p.helem = c.helem.p; p.W = c.W.p; p.We = c.We.p; p.Wi = c.Wi.p;
    p.hnode = c.hnode.p; p.hnode_new = c.hnode_new.p; p.Z3 = c.Z3.p; p.zbar3 = c.zbar3.p; p.Kv = c.Kv.p; p.Ki = c.Ki.p;
    p.sl1 = c.sl1.p; p.sl2 = c.sl2.p; p.sl3 = c.sl3.p; p.sw3d = c.sw3d.p;
    p.heat_flux = c.heat_flux.p; p.water_flux = c.water_flux.p; p.virtual_salt = c.virtual_salt.p; p.relax_salt = c.relax_salt.p; p.real_salt_flux = c.real_salt_flux.p;
    const size_t nf = (size_t)c.Ntiles * c.nlp * TILE;
    p.T = c.T.p + nf * tr; p.Told = c.Told.p + nf * tr; p.T1 = c.T.p;
    p.txy_x = c.txy_x.p; p.txy_y = c.txy_y.p; p.trz = c.trz.p; p.LO = c.LO.p; p.adf_h = c.adf_h.p; p.adf_v = c.adf_v.p;
    p.fplus = c.fplus.p; p.fminus = c.fminus.p; p.tstar = c.tstar.p;
    return p;
}

static bool g_attr_done = false;
static void set_smem_attrs() {
    if (g_attr_done) return;
    const int big = 200 * 1024;
    CUDA_CHECK(cudaFuncSetAttribute(k_lo, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    CUDA_CHECK(cudaFuncSetAttribute(k_vert_impl, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    CUDA_CHECK(cudaFuncSetAttribute(k_fct_bounds, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    CUDA_CHECK(cudaFuncSetAttribute(k_update, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    CUDA_CHECK(cudaFuncSetAttribute(k_impl_diff, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    g_attr_done = true;
}

// solve_tracers_ale (oce_ale_tracer.F90:101-199) for one partition, everything device resident
void tracer_step_all(TracerCtx& c) {
    if (!fb_ok()) return;
    if (!c.committed) { fb_set_error(20, "fb200_solve_tracers before fb200_init_commit"); return; }
    if (c.opt.tra_adv_lim != 0) { fb_set_error(21, "tra_adv_lim='NONE' is not implemented on the device path"); return; }
    if (c.opt.tra_adv_ver == 2) { fb_set_error(22, "tra_adv_ver='PPM' is not implemented on the device path"); return; }
    if (!c.opt.i_vert_diff) { fb_set_error(23, "i_vert_diff=.false. is not implemented on the device path"); return; }
    if (c.opt.dt <= 0.0) { fb_set_error(24, "dt not set (fb200_set_options)"); return; }
    set_smem_attrs();
    c.stream = fb_stream();
    cudaStream_t st = c.stream;
    const size_t nf = (size_t)c.Ntiles * c.nlp * TILE;
    const int gN128 = (c.myN + 127) / 128, gN64 = (c.myN + 63) / 64, gE = (c.myE + 127) / 128, gEd = (c.myEd + 127) / 128, gNall = (c.N + 127) / 128;
    const size_t sm1_128 = (size_t)c.nlp * 128 * sizeof(double);
    const size_t sm_64 = (size_t)c.nlp * 64 * sizeof(double);
    CUDA_CHECK(cudaEventRecord(c.ev0, st));
#define KL(name, ...) do { ProfScope ps_(name, st); __VA_ARGS__; LAUNCH_CHECK(); } while (0)
    for (int tr = 0; tr < c.ntr; tr++) {
        KP p = make_kp(c, tr);
        const double ca = -(0.5 + c.opt.epsilon), cb = 1.5 + c.opt.epsilon;
        KL("k_ab2", k_ab2<<<148 * 8, 256, 0, st>>>(p.Told, p.T, nf, ca, cb));
        KL("k_grad_elem", k_grad_elem<<<gE, 128, 0, st>>>(p, p.Told));
        KL("k_lo", k_lo<<<gN128, 128, sm1_128, st>>>(p));
        if (c.opt.w_split) KL("k_vert_impl", k_vert_impl<<<gN64, 64, 2 * sm_64, st>>>(p, p.LO, p.Wi));
        KL("k_ho_edge", k_ho_edge<<<gEd, 128, 0, st>>>(p));
        KL("k_fct_bounds", k_fct_bounds<<<gN64, 64, 4 * sm_64, st>>>(p));
        if (c.opt.do_hor_diff) KL("k_grad_elem", k_grad_elem<<<gE, 128, 0, st>>>(p, p.T));
        if (c.opt.do_hor_diff && c.opt.Redi) KL("k_grad_z", k_grad_z<<<gNall, 128, 0, st>>>(p, p.T));
        KL("k_update", k_update<<<gN128, 128, sm1_128, st>>>(p));
        {   // tr_arr_old = tr_arr (:274)
            ProfScope ps_("copy_told", st);
            CUDA_CHECK(cudaMemcpyAsync(p.Told, p.T, nf * sizeof(double), cudaMemcpyDeviceToDevice, st));
        }
        KL("k_impl_diff", k_impl_diff<<<gN64, 64, 3 * sm_64, st>>>(p));
    }
    if (c.opt.clamp_salt && c.ntr >= 2) { KP p = make_kp(c, 1); KL("k_clamp_salt", k_clamp_salt<<<gNall, 128, 0, st>>>(p, p.T)); }
#undef KL
    CUDA_CHECK(cudaEventRecord(c.ev1, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    if (fb_ok()) CUDA_CHECK(cudaEventElapsedTime(&c.last_ms, c.ev0, c.ev1));
}

// debug access to work fields (tiled) by name
const double* tracer_debug_field(TracerCtx& c, const char* name, int* kind) {
    std::string s(name);
    *kind = 0;                                    // 0 node field, 1 edge field, 2 element field
    if (s == "fct_LO") return c.LO.p;
    if (s == "adf_v") return c.adf_v.p;
    if (s == "fct_plus") return c.fplus.p;
    if (s == "fct_minus") return c.fminus.p;
    if (s == "tstar") return c.tstar.p;
    if (s == "adf_h") { *kind = 1; return c.adf_h.p; }
    if (s == "tr_xy_x") { *kind = 2; return c.txy_x.p; }
    if (s == "tr_xy_y") { *kind = 2; return c.txy_y.p; }
    return nullptr;
}
