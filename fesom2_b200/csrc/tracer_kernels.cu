// Tracer transport kernels (FP64, sm_100a): AB2, elemental gradients, low-order solution, high-order
// antidiffusive fluxes, Zalesak limiter, update, horizontal diffusion, implicit vertical diffusion.
//
// All edge->node scatters of the reference are node-centric gathers here: each own node walks its incident
// edges in ascending reference edge id, which is the order in which the reference's edge loops add into that
// node, so sums are reproduced term by term without atomics.  Arithmetic follows the Fortran expression order
// (compiled with --fmad=false) — see the file:line notes at each kernel.
#include "common.cuh"
#include "comm.cuh"
#include "tracer_dev.cuh"
#include "tracer_kernels2.cuh"
#include <algorithm>
#include <cstdlib>

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
        AT(p.txy, be, nz) = make_double2(g1 * t1 + g2 * t2 + g3 * t3, g4 * t1 + g5 * t2 + g6 * t3);
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
// Thread mapping of the node kernels: block = one tile of 32 consecutive nodes (threadIdx.x = lane = node) times
// blockDim.y level rows; every thread owns LPT levels nz = threadIdx.y + 1 + q * blockDim.y of its node, so a warp
// reads one coalesced 256-byte line per field and level, all loads of a thread are independent (no serial walk down
// the column) and sums live in registers.  The per-(node, slot) tables ns_* are read once per thread and reused for
// its LPT levels.
#define LPT 3
struct NodeCol { int n, b, nzmin, nzmax, nlmin; bool valid; };
__device__ __forceinline__ NodeCol node_col(const KP& p) {
    NodeCol c;
    c.n = blockIdx.x * TILE + threadIdx.x;
    c.valid = c.n < p.myN;
    c.b = blockIdx.x * (p.nlp * TILE) + threadIdx.x;
    const unsigned lv = c.valid ? p.n_lev[c.n] : 0u;
    c.nzmin = lv & 255; c.nzmax = (lv >> 8) & 255; c.nlmin = (lv >> 16) & 255;
    return c;
}

// K3: low-order solution + vertical antidiffusive flux.
//   fct_LO  : oce_adv_tra_driver.F90:83-122 (adv_tra_hor_upw1 + edge scatter + adv_tra_ver_upw1 + column update)
//   adf_v   : oce_adv_tra_driver.F90:163-179 with init_zero=.false.  (HO(ttfAB, W) - LO(ttf, W)); adf_v(nzmax) is
//             0 - 0 for every scheme and is never written (the buffer is zero-initialised)
template <int VER, bool FCT>                // vertical scheme fixed at compile time: the PPM code stays out of the default kernel
__global__ void __launch_bounds__(256, 5) k_lo(KP p) {
    const NodeCol c = node_col(p);
    if (!c.valid) return;
    const double* __restrict__ T = p.T;
    int nz[LPT]; bool act[LPT]; double acc[LPT], tn[LPT];
    bool any = false;
#pragma unroll
    for (int q = 0; q < LPT; q++) {
        nz[q] = blockIdx.y * (LPT * blockDim.y) + threadIdx.y + 1 + q * blockDim.y;
        act[q] = nz[q] >= c.nzmin && nz[q] <= c.nzmax - 1;
        any |= act[q];
        acc[q] = 0.0;
        tn[q] = act[q] ? AT(T, c.b, nz[q]) : 0.0;
    }
    if (!any) return;
    const double* __restrict__ TAB = p.Told;
    if (!FCT) {                                                    // tra_adv_lim='NONE': adv_flux_ver = HO(ttfAB, Wvel_e), init_zero=.true. (:163-179)
#pragma unroll
        for (int q = 0; q < LPT; q++) {
            if (!act[q]) continue;
            const int z = nz[q];
            double ho;
            if (VER == 0) ho = ver_qr4c(p, TAB, p.We, c.b, z, c.nzmin, c.nzmax);
            else if (VER == 1) ho = ver_cdiff(p, TAB, p.We, c.b, z, c.nzmin);
            else if (VER == 2) ho = ver_ppm(p, TAB, p.We, c.b, z, c.nzmin, c.nzmax);
            else ho = ver_upw1(TAB, p.We, p.area, c.b, z, c.nzmin, c.nzmax);
            AT(p.adf_v, c.b, z) = ho - 0.0;
        }
        return;
    }
    const size_t sbase = (size_t)blockIdx.x * p.maxdeg * TILE + threadIdx.x;
    const int ndeg = p.tile_deg[blockIdx.x];
    for (int j = 0; j < ndeg; j++) {
        const size_t k = sbase + (size_t)j * TILE;
        const SlotRec sr = p.ns_p[k];                                          // empty slots have zero masks
        const bool second = sr.mk & 1ull;
        const int bed = sr.bed;
        const unsigned long long many = sr.mk;
#pragma unroll
        for (int q = 0; q < LPT; q++) {
            if (!act[q] || !((many >> nz[q]) & 1ull)) continue;
            const double f = AT(p.eflx, bed, nz[q]);                           // upwind edge flux, computed once per edge by k_ho_edge
            acc[q] = second ? acc[q] - f : acc[q] + f;
        }
    }
    const double* __restrict__ Wlo = p.We;                         // explicit part for the LO solution
    const double* __restrict__ Wad = p.w_split ? p.W : p.We;       // LO part of the antidiffusive flux: full W (:131)
#pragma unroll
    for (int q = 0; q < LPT; q++) {
        if (!act[q]) continue;
        const int z = nz[q];
        const double fv = ver_upw1(T, Wlo, p.area, c.b, z, c.nzmin, c.nzmax);
        const double fvn = ver_upw1(T, Wlo, p.area, c.b, z + 1, c.nzmin, c.nzmax);
        AT(p.LO, c.b, z) = (tn[q] * AT(p.hnode, c.b, z) + (acc[q] + (fv - fvn)) * p.dt / AT(p.areasvol, c.b, z)) / AT(p.hnode_new, c.b, z);
        const double flo = ver_upw1(T, Wad, p.area, c.b, z, c.nzmin, c.nzmax);
        double ho;
        if (VER == 0) ho = ver_qr4c(p, TAB, p.W, c.b, z, c.nzmin, c.nzmax) - flo;
        else if (VER == 1) ho = ver_cdiff(p, TAB, p.W, c.b, z, c.nzmin) - flo;
        else if (VER == 2) ho = ver_ppm(p, TAB, p.W, c.b, z, c.nzmin, c.nzmax) - flo;
        else ho = ver_upw1(TAB, p.W, p.area, c.b, z, c.nzmin, c.nzmax) - flo;
        AT(p.adf_v, c.b, z) = ho;
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
            bb = bb - dmin(0.0, AT(W, b, nz + 1)) * v_adv;
            c = -dmax(0.0, AT(W, b, nz + 1)) * v_adv;
        } else if (nz <= nzmax - 2) {
            v_adv = zinv * AT(p.area, b, nz) / asv;
            a = dmin(0.0, AT(W, b, nz)) * v_adv;
            bb = hn + dmax(0.0, AT(W, b, nz)) * v_adv;
            v_adv = zinv * AT(p.area, b, nz + 1) / asv;
            bb = bb - dmin(0.0, AT(W, b, nz + 1)) * v_adv;
            c = -dmax(0.0, AT(W, b, nz + 1)) * v_adv;
        } else {
            v_adv = zinv * AT(p.area, b, nz) / asv;
            a = dmin(0.0, AT(W, b, nz)) * v_adv;
            bb = hn + dmax(0.0, AT(W, b, nz)) * v_adv;
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

// K4: high-order horizontal flux minus low-order flux = antidiffusive flux adf_h.  Block = one tile of 32 edges
// times blockDim.y level rows, LPT levels per thread (same mapping as the node kernels).
//   adv_tra_hor_mfct / _muscl / _upw1 with init_zero=.false. (oce_adv_tra_hor.F90:215-733) on ttfAB,
//   the input flux being adv_tra_hor_upw1(ttf) (oce_adv_tra_driver.F90:83,142-149); the gradients of
//   fill_up_dn_grad (oce_muscl_adv.F90:285-447) are gathered from txy instead of being stored.
// One (edge, level): FB = the gradients come from the node-mean fallback (levels outside the range the up/downwind
// triangles share, or an edge without them) instead of txy of the two triangles.
template <bool FCT, bool FB>
__device__ __forceinline__ void ho_edge_level(const KP& p, const EdgeRec& r, const EdgeHo& h, int be, int nz, int segs, int nbl1, int nbl2) {
    const double* __restrict__ T = p.T;
    const double* __restrict__ A = p.Told;
    const double vflux = seg_vflux(p, segs, r.be1, r.be2, r.dX1, r.dY1, r.dX2, r.dY2, nz);
    const double flo = FCT ? upw1_flux(AT(T, r.bn1, nz), AT(T, r.bn2, nz), vflux) : 0.0;
    const double a1 = AT(A, r.bn1, nz), a2 = AT(A, r.bn2, nz);
    double fho;
    if (p.hor == 2) fho = upw1_flux(a1, a2, vflux);
    else {
        double g1, g2, g3, g4;         // edge_up_dn_grad(1:4): x up, x down, y up, y down
        if (!FB) {
            const double2 gu = AT(p.txy, h.bup, nz), gd = AT(p.txy, h.bdn, nz);
            g1 = gu.x; g2 = gd.x; g3 = gu.y; g4 = gd.y;
        } else {
            node_mean_grad(p.nie_slot, p.maxnie, p.el_lev, p.el_area, p.txy, p.nlp, h.n1, nz, g1, g3);
            node_mean_grad(p.nie_slot, p.maxnie, p.el_lev, p.el_area, p.txy, p.nlp, h.n2, nz, g2, g4);
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
    if (FCT) AT(p.eflx, be, nz) = flo;                               // adv_tra_hor_upw1(ttf), gathered by k_lo
}

template <bool FCT>                         // FCT: antidiffusive flux HO - LO and the LO flux; else the HO flux itself (init_zero=.true.)
__global__ void __launch_bounds__(256, 4) k_ho_edge(KP p) {
    const int e = blockIdx.x * TILE + threadIdx.x;
    if (e >= p.myEd) return;
    const EdgeRec r = p.edge[e];
    const unsigned long long many = r.m1 | r.m2;
    if ((many >> min(63u, blockIdx.y * (LPT * blockDim.y) + threadIdx.y + 1)) == 0ull) return;                 // no flux-carrying level at or below this thread's first level
    const EdgeHo h = p.edgeho[e];
    const int be = blockIdx.x * (p.nlp * TILE) + threadIdx.x;
    const int sh_lo = h.shared & 255, sh_hi = (h.shared >> 8) & 255;
    const bool good = (h.shared >> 16) & 1;
    const bool grads = p.hor != 2;                                                               // UPW1 needs no gradients: no fallback levels
    const int nbl1 = p.hor == 0 ? p.n_nbl[h.n1] : 0, nbl2 = p.hor == 0 ? p.n_nbl[h.n2] : 0;     // only MUSCL uses nboundary_lay
#pragma unroll
    for (int q = 0; q < LPT; q++) {
        const int nz = blockIdx.y * (LPT * blockDim.y) + threadIdx.y + 1 + q * blockDim.y;
        if (nz > 63) continue;
        const int segs = mask_segs(r.m1, r.m2, nz);
        if (!segs) continue;
        if (grads && !(good && nz >= sh_lo && nz <= sh_hi - 1)) continue;                        // fallback level: k_ho_edge_fb (list built at commit)
        ho_edge_level<FCT, false>(p, r, h, be, nz, segs, nbl1, nbl2);
    }
}
// the (edge, level) pairs whose gradients are node means (about 2 % of them on a ragged-bottom mesh, but spread over
// 40 % of the edges): one thread per pair from a list built at commit, so no warp of k_ho_edge diverges into the
// element loops of node_mean_grad
template <bool FCT>
__global__ void __launch_bounds__(128) k_ho_edge_fb(KP p, const int2* __restrict__ list, int count) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const int e = list[i].x, nz = list[i].y;
    const EdgeRec r = p.edge[e];
    const EdgeHo h = p.edgeho[e];
    const int nbl1 = p.hor == 0 ? p.n_nbl[h.n1] : 0, nbl2 = p.hor == 0 ? p.n_nbl[h.n2] : 0;
    ho_edge_level<FCT, true>(p, r, h, tbase(e, p.nlp), nz, mask_segs(r.m1, r.m2, nz), nbl1, nbl2);
}

// ---------------------------------------------------------------------------------------------------
// K5: Zalesak limiter, parts a1-a3 and b1-b2 of oce_tra_adv_fct (oce_adv_tra_fct.F90:93-276).  Node-kernel
// mapping (tile x level rows, LPT levels per thread).  The element min/max pass (a2) is folded into the node pass:
// the bound of node n at level nz is the max/min over n and over every edge neighbour m whose edge has a wet
// adjacent element at nz — the same set of values the reference reaches through
// UV_rhs(1:2,nz,nod_in_elem2D(:,n)); "-bignumber/+bignumber" enter wherever a surrounding element is already below
// its bottom (nz >= nlevels_nod2D_min).  The vertical 3-point max/min (a3, vlimit=1) goes through shared memory.
// dynamic smem: mx, mn  [2][nlp][32]
template <int LP>                           // levels per thread: 3 up to 48 layers (16 level rows), 4 up to 62 (the block holds the whole column)
__global__ void __launch_bounds__(512, 2) k_fct_bounds(KP p) {
    extern __shared__ double sm[];
    double* smx = sm; double* smn = sm + (size_t)p.nlp * TILE;
    const NodeCol c = node_col(p);
    const double flux_eps = 1e-16, bignumber = 1e3;
    const double* __restrict__ T = p.T;
    const double* __restrict__ LO = p.LO;
    int nz[LP]; bool act[LP]; double pp[LP], pm[LP], mx[LP], mn[LP], lo[LP];
    bool any = false;
#pragma unroll
    for (int q = 0; q < LP; q++) {
        nz[q] = blockIdx.y * (LP * blockDim.y) + threadIdx.y + 1 + q * blockDim.y;
        act[q] = c.valid && nz[q] >= c.nzmin && nz[q] <= c.nzmax - 1;
        pp[q] = pm[q] = mx[q] = mn[q] = lo[q] = 0.0;
        if (!act[q]) continue;
        any = true;
        const double av = AT(p.adf_v, c.b, nz[q]), avn = AT(p.adf_v, c.b, nz[q] + 1);
        pp[q] = dmax(0.0, av) + dmax(0.0, -avn);                              // :235-236
        pm[q] = dmin(0.0, av) + dmin(0.0, -avn);
        lo[q] = AT(LO, c.b, nz[q]);
        const double t = AT(T, c.b, nz[q]);
        double M = dmax(lo[q], t), m = dmin(lo[q], t);                        // a1 :99-100
        if (nz[q] >= c.nlmin) { M = dmax(M, -bignumber); m = dmin(m, bignumber); }   // a2 :116-121
        mx[q] = M; mn[q] = m;
    }
    if (any) {
        const size_t sbase = (size_t)blockIdx.x * p.maxdeg * TILE + threadIdx.x;
        const int ndeg = p.tile_deg[blockIdx.x];
        for (int j = 0; j < ndeg; j++) {
            const size_t k = sbase + (size_t)j * TILE;
            const SlotRec sr = p.ns_p[k];
            const bool second = sr.mk & 1ull;
            const int bo = sr.bo, bed = sr.bed;
            const unsigned long long many = sr.mk;
#pragma unroll
            for (int q = 0; q < LP; q++) {
                if (!act[q] || !((many >> nz[q]) & 1ull)) continue;
                double f = AT(p.adf_h, bed, nz[q]);
                const double lo_o = AT(LO, bo, nz[q]), t_o = AT(T, bo, nz[q]);
                if (second) f = -f;
                pp[q] = pp[q] + dmax(0.0, f);                                  // b1 :258-261
                pm[q] = pm[q] + dmin(0.0, f);
                mx[q] = dmax(mx[q], dmax(lo_o, t_o));
                mn[q] = dmin(mn[q], dmin(lo_o, t_o));
            }
        }
#pragma unroll
        for (int q = 0; q < LP; q++)
            if (act[q]) { smx[(nz[q] - 1) * TILE + threadIdx.x] = mx[q]; smn[(nz[q] - 1) * TILE + threadIdx.x] = mn[q]; }
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < LP; q++) {
        if (!act[q]) continue;
        const int z = nz[q];
        double qmax = mx[q], qmin = mn[q];
        if (z > c.nzmin && z < c.nzmax - 1) {                                  // a3, vlimit=1 :153-156
            qmax = dmax(dmax(smx[(z - 2) * TILE + threadIdx.x], qmax), smx[z * TILE + threadIdx.x]);
            qmin = dmin(dmin(smn[(z - 2) * TILE + threadIdx.x], qmin), smn[z * TILE + threadIdx.x]);
        }
        const double asv = AT(p.areasvol, c.b, z), rasv = 1.0 / asv;
        qmax = qmax - lo[q]; qmin = qmin - lo[q];
        double flux = div_by(pp[q] * p.dt, asv, rasv) + flux_eps;              // b2 :271-274
        const double rp = dmin(1.0, qmax / flux);
        flux = div_by(pm[q] * p.dt, asv, rasv) - flux_eps;
        AT(p.fpm, c.b, z) = make_double2(rp, dmin(1.0, qmin / flux));
    }
}

// limited vertical flux at interface nz (b3 vertical, oce_adv_tra_fct.F90:284-311)
__device__ __forceinline__ double limited_fv(const KP& p, int b, int nz, int nzmin, int nzmax) {
    const double flux = AT(p.adf_v, b, nz);
    if (nz == nzmax) return flux;
    double ae = 1.0;
    const double2 f0 = AT(p.fpm, b, nz);                   // (fct_plus, fct_minus)
    if (nz == nzmin) {
        if (flux >= 0.0) ae = dmin(ae, f0.x); else ae = dmin(ae, f0.y);
    } else {
        const double2 fm1 = AT(p.fpm, b, nz - 1);
        if (flux >= 0.0) { ae = dmin(ae, fm1.y); ae = dmin(ae, f0.x); }
        else { ae = dmin(ae, fm1.x); ae = dmin(ae, f0.y); }
    }
    return ae * flux;
}

// K6: diff_part_hor_redi (oce_ale_tracer.F90:934-1077), the flux through the two median-dual segments of every own
// edge and level, computed once per edge (edge-kernel mapping) and gathered by k_update.
__global__ void __launch_bounds__(256, 5) k_diff_edge(KP p) {
    const int e = blockIdx.x * TILE + threadIdx.x;
    if (e >= p.myEd) return;
    const EdgeRec r = p.edge[e];
    const unsigned long long many = r.m1 | r.m2;
    if ((many >> min(63u, blockIdx.y * (LPT * blockDim.y) + threadIdx.y + 1)) == 0ull) return;
    const int be = blockIdx.x * (p.nlp * TILE) + threadIdx.x;
    const double isredi = p.Redi ? 1.0 : 0.0;
#pragma unroll
    for (int q = 0; q < LPT; q++) {
        const int z = blockIdx.y * (LPT * blockDim.y) + threadIdx.y + 1 + q * blockDim.y;
        if (z > 63) continue;
        const int segs = mask_segs(r.m1, r.m2, z);
        if (!segs) continue;
        const double Kh = (AT(p.Ki, r.bn1, z) + AT(p.Ki, r.bn2, z)) / 2.0;
        double SxTz = 0.0, SyTz = 0.0;
        if (p.Redi) {
            const double Tz1 = 0.5 * (AT(p.trz, r.bn1, z) + AT(p.trz, r.bn1, z + 1));
            const double Tz2 = 0.5 * (AT(p.trz, r.bn2, z) + AT(p.trz, r.bn2, z + 1));
            SxTz = (Tz1 * AT(p.sl1, r.bn1, z) + Tz2 * AT(p.sl1, r.bn2, z)) / 2.0;
            SyTz = (Tz1 * AT(p.sl2, r.bn1, z) + Tz2 * AT(p.sl2, r.bn2, z)) / 2.0;
        }
        double dz, Tx, Ty, cc;                                              // :1020-1062
        if (segs == 3) {
            const double2 ga = AT(p.txy, r.be1, z), gb = AT(p.txy, r.be2, z);
            dz = (AT(p.helem, r.be1, z) + AT(p.helem, r.be2, z)) / 2.0;
            Tx = 0.5 * (ga.x + gb.x);
            Ty = 0.5 * (ga.y + gb.y);
            const double Fx = Kh * (Tx + SxTz * isredi), Fy = Kh * (Ty + SyTz * isredi);
            cc = ((r.dX2 - r.dX1) * Fy - (r.dY2 - r.dY1) * Fx) * dz;
        } else if (segs == 1) {
            const double2 ga = AT(p.txy, r.be1, z);
            dz = AT(p.helem, r.be1, z); Tx = ga.x; Ty = ga.y;
            const double Fx = Kh * (Tx + SxTz * isredi), Fy = Kh * (Ty + SyTz * isredi);
            cc = (-r.dX1 * Fy + r.dY1 * Fx) * dz;
        } else {
            const double2 gb = AT(p.txy, r.be2, z);
            dz = AT(p.helem, r.be2, z); Tx = gb.x; Ty = gb.y;
            const double Fx = Kh * (Tx + SxTz * isredi), Fy = Kh * (Ty + SyTz * isredi);
            cc = (r.dX2 * Fy - r.dY2 * Fx) * dz;
        }
        AT(p.eflx, be, z) = cc;
    }
}

// K7a: b3 limiting + oce_tra_adv_flux2dtracer (oce_adv_tra_driver.F90:223-268) + del_ttf assembly
// (oce_ale_tracer.F90:240) + diff_part_hor_redi (:934-1077) + T* update (:294-305).  Node-kernel mapping.
// Writes tstar; tr_arr_old = tr_arr is done by the caller for all entries.
// FCT: limited fluxes (b3) and the LO terms of flux2dtracer, else the raw high-order fluxes; EXPL: i_vert_diff=.false.,
// the explicit vertical diffusion joins del_ttf here and no implicit solve follows
template <bool FCT, bool EXPL>
__global__ void __launch_bounds__(256, 5) k_update(KP p) {
    const NodeCol c = node_col(p);
    if (!c.valid) return;
    const double* __restrict__ T = p.T;
    int nz[LPT]; bool act[LPT]; double acc[LPT], fp[LPT], fm[LPT], asv[LPT], rasv[LPT];
    bool any = false;
#pragma unroll
    for (int q = 0; q < LPT; q++) {
        nz[q] = blockIdx.y * (LPT * blockDim.y) + threadIdx.y + 1 + q * blockDim.y;
        act[q] = nz[q] >= c.nzmin && nz[q] <= c.nzmax - 1;
        any |= act[q];
        acc[q] = 0.0; fp[q] = fm[q] = 0.0; asv[q] = rasv[q] = 1.0;
        if (!act[q]) continue;
        if (FCT) { const double2 f = AT(p.fpm, c.b, nz[q]); fp[q] = f.x; fm[q] = f.y; }
        asv[q] = AT(p.areasvol, c.b, nz[q]); rasv[q] = 1.0 / asv[q];
    }
    if (!any) return;
    const size_t sbase = (size_t)blockIdx.x * p.maxdeg * TILE + threadIdx.x;
    const int ndeg = p.tile_deg[blockIdx.x];
    for (int j = 0; j < ndeg; j++) {                                         // dttf_h
        const size_t k = sbase + (size_t)j * TILE;
        const SlotRec sr = p.ns_p[k];
        const bool second = sr.mk & 1ull;
        const int bo = sr.bo, bed = sr.bed;
        const unsigned long long many = sr.mk;
#pragma unroll
        for (int q = 0; q < LPT; q++) {
            if (!act[q] || !((many >> nz[q]) & 1ull)) continue;
            const double flux = AT(p.adf_h, bed, nz[q]);
            double fl = flux;
            if (FCT) {
                const double2 fo = AT(p.fpm, bo, nz[q]);
                // :313-330  flux >= 0: min(fct_plus(n1), fct_minus(n2)) else min(fct_minus(n1), fct_plus(n2)), n1/n2 the
                // edge's nodes; this node is n2 when `second`
                const bool self_plus = (flux >= 0.0) != second;
                double ae = dmin(1.0, self_plus ? fp[q] : fm[q]);
                ae = dmin(ae, self_plus ? fo.y : fo.x);
                fl = ae * flux;
            }
            const double d = div_by(fl * p.dt, asv[q], rasv[q]);
            acc[q] = second ? acc[q] - d : acc[q] + d;
        }
    }
#pragma unroll
    for (int q = 0; q < LPT; q++) {                                          // dttf_v, then del_ttf = dttf_h + dttf_v
        if (!act[q]) continue;
        const int z = nz[q];
        const double fv = FCT ? limited_fv(p, c.b, z, c.nzmin, c.nzmax) : AT(p.adf_v, c.b, z);
        const double fvn = FCT ? limited_fv(p, c.b, z + 1, c.nzmin, c.nzmax) : AT(p.adf_v, c.b, z + 1);
        double dv = 0.0;
        if (FCT) dv = 0.0 - AT(T, c.b, z) * AT(p.hnode, c.b, z) + AT(p.LO, c.b, z) * AT(p.hnode_new, c.b, z);
        dv = dv + div_by((fv - fvn) * p.dt, asv[q], rasv[q]);
        acc[q] = acc[q] + dv;
    }
    if (p.do_hor_diff) {                                                     // diff_part_hor_redi: fluxes from k_diff_edge
        for (int j = 0; j < ndeg; j++) {
            const size_t k = sbase + (size_t)j * TILE;
            const SlotRec sr = p.ns_p[k];
            const bool second = sr.mk & 1ull;
            const int bed = sr.bed;
            const unsigned long long many = sr.mk;
#pragma unroll
            for (int q = 0; q < LPT; q++) {
                if (!act[q] || !((many >> nz[q]) & 1ull)) continue;
                const double cc = AT(p.eflx, bed, nz[q]);
                const double rhs = second ? 0.0 - cc : 0.0 + cc;
                acc[q] = acc[q] + div_by(rhs * p.dt, asv[q], rasv[q]);
            }
        }
    }
    if (EXPL) {                                                              // diff_ver_part_expl_ale :381-392
#pragma unroll
        for (int q = 0; q < LPT; q++) {
            if (!act[q]) continue;
            const int z = nz[q], nl1 = c.nzmax - 1;
            const double vd = expl_vd(p, T, c.n, c.b, z, c.nzmin, nl1);
            const double dzb = AT(p.zbar3, c.b, z) - AT(p.zbar3, c.b, z + 1);
            double term;
            if (z <= nl1 - 1) term = (vd - expl_vd(p, T, c.n, c.b, z + 1, c.nzmin, nl1)) / dzb * p.dt;
            else term = (vd / dzb) * p.dt;
            acc[q] = acc[q] + div_by(term, asv[q], rasv[q]);
        }
    }
    if (p.Redi) {                                                            // diff_ver_part_redi_expl :288, term from k_redi_expl
#pragma unroll
        for (int q = 0; q < LPT; q++) if (act[q]) acc[q] = acc[q] + AT(p.redi_v, c.b, nz[q]);
    }
#pragma unroll
    for (int q = 0; q < LPT; q++) {                                          // T*  :302-305
        if (!act[q]) continue;
        const int z = nz[q];
        const double t = AT(T, c.b, z);
        const double del = acc[q] + t * (AT(p.hnode, c.b, z) - AT(p.hnode_new, c.b, z));
        AT(p.tstar, c.b, z) = t + del / AT(p.hnode_new, c.b, z);
    }
}

// ---------------------------------------------------------------------------------------------------
// diff_ver_part_redi_expl (oce_ale_tracer.F90:860-930): the projection of the horizontal Redi diffusivity onto the
// vertical.  Thread per own column.  Pass 1 (bottom-up) leaves zbar_n(z) in the scratch column zs (:904-914); pass 2
// (top-down) forms tr_xynodes(1:2,z) — the area-weighted sum of tr_xy over nod_in_elem2D in the reference's order, divided
// by 3 and by areasvol (:876-893) —, the product P(z) = slope_tapered(1)*tr_xynodes(1) + slope_tapered(2)*tr_xynodes(2) that
// both lines of vd_flux share (:918-921), vd_flux(z), and writes the del_ttf contribution of level z-1 (:925-927) to out.
// p.txy must hold the gradients of tr_arr (the second tracer_gradient_elements call of init_tracers_AB).
__global__ void LB(128) k_redi_expl(KP p, double* __restrict__ zs, double* __restrict__ out) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= p.myN) return;
    const int b = tbase(n, p.nlp);
    const unsigned lv = p.n_lev[n];
    const int ul1 = lv & 255, nl1 = (int)((lv >> 8) & 255) - 1;
    const double* __restrict__ hn = p.hnode_new;
    const double zbot = p.n_znb[n];
    {
        double zb = zbot;
        for (int z = nl1; z >= ul1; z--) { zb = zb + AT(hn, b, z); AT(zs, b, z) = zb; }
    }
    const int* slot = p.nie_slot + (size_t)(n >> 5) * p.maxnie * TILE + (n & 31);
    double vd_prev = 0.0, P_prev = 0.0, Ki_prev = 0.0, Z_prev = 0.0;            // vd_flux(ul1) = 0
    for (int z = ul1; z <= nl1; z++) {
        double Tx = 0.0, Ty = 0.0;
        for (int k = 0; k < p.maxnie; k++) {
            const int e = slot[k * TILE];
            if (e < 0) break;
            const unsigned le = p.el_lev[e];
            if (z <= (int)((le >> 8) & 255) && z >= (int)(le & 255)) {
                const double ar = p.el_area[e];
                const double2 g = AT(p.txy, tbase(e, p.nlp), z);
                Tx = Tx + g.x * ar;
                Ty = Ty + g.y * ar;
            }
        }
        const double asv = AT(p.areasvol, b, z);
        const double x1 = Tx / 3.0 / asv, x2 = Ty / 3.0 / asv;
        const double P = AT(p.sl1, b, z) * x1 + AT(p.sl2, b, z) * x2;
        const double Ki = AT(p.Ki, b, z);
        const double zb = AT(zs, b, z);                                         // zbar_n(z)
        const double zbn = z + 1 <= nl1 ? AT(zs, b, z + 1) : zbot;              // zbar_n(z+1)
        const double Z = zbn + AT(hn, b, z) / 2.0;                              // Z_n(z)
        double vd = 0.0;
        if (z > ul1) {
            vd = (Z_prev - zb) * P_prev * Ki_prev;
            vd = vd + (zb - Z) * P * Ki;
            vd = vd / (Z_prev - Z) * AT(p.area, b, z);
            AT(out, b, z - 1) = (vd_prev - vd) * p.dt / AT(p.areasvol, b, z - 1);
        }
        vd_prev = vd; P_prev = P; Ki_prev = Ki; Z_prev = Z;
    }
    AT(out, b, nl1) = (vd_prev - 0.0) * p.dt / AT(p.areasvol, b, nl1);          // vd_flux(nl1+1) = 0
}

// diff_part_bh (oce_ale_tracer.F90:1081-1148): biharmonic diffusion as a two-round filter over the edges with two
// triangles.  Thread per own column, per level the incident edges in ascending reference edge id (the order in which
// the reference's edge loop reaches this node).  ROUND 1: dst = temporary_ttf (own nodes; the caller exchanges the halo),
// src = ttf.  ROUND 2: src = temporary_ttf, dst = ttf, updated in place (a thread writes its own column only).
template <int ROUND>
__global__ void LB(128) k_bh(KP p, const double* __restrict__ src, double* __restrict__ dst, double g0, double g1, double g2) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= p.myN) return;
    const int b = tbase(n, p.nlp);
    const unsigned lv = p.n_lev[n];
    const int nlmin_s = (lv >> 16) & 255, ulmax_s = (lv >> 24) & 255;
    const size_t sbase = (size_t)(n >> 5) * p.maxdeg * TILE + (n & 31);
    for (int z = 1; z <= p.nl - 1; z++) {
        double acc = ROUND == 1 ? 0.0 : AT(dst, b, z);
        bool touched = false;
        for (int j = 0; j < p.maxdeg; j++) {
            const size_t k = sbase + (size_t)j * TILE;
            const int es = p.ne_slot[k];
            if (es < 0) break;
            const int4 ni = p.ns_i[k];                                          // other node, element 1, element 2, edge
            if (ni.z < 0) continue;                                             // boundary edge: myList_edge2D(ed) > edge2D_in
            const int no = node_of_base_d(ni.x, p.nlp);
            const unsigned lo = p.n_lev[no];
            const int ul1 = min(ulmax_s, (int)((lo >> 24) & 255)), nl1 = max(nlmin_s, (int)((lo >> 16) & 255)) - 1;
            if (z < ul1 || z > nl1) continue;
            const bool second = es & 1;
            const double len = sqrt(p.el_area[node_of_base_d(ni.y, p.nlp)] + p.el_area[node_of_base_d(ni.z, p.nlp)]);
            const double2 ua = AT(p.UV, ni.y, z), ub = AT(p.UV, ni.z, z);
            const double u1 = ua.x - ub.x, v1 = ua.y - ub.y;
            double vi = u1 * u1 + v1 * v1;
            vi = sqrt(omax(g0, omax(g1 * sqrt(vi), g2 * vi)) * len);
            const double s_self = AT(src, b, z), s_oth = AT(src, ni.x, z);
            double tt = second ? s_oth - s_self : s_self - s_oth;               // value(en(1)) - value(en(2))
            if (ROUND == 1) {
                tt = tt * vi;
                acc = second ? acc + tt : acc - tt;
            } else {
                tt = -tt * vi * p.dt;
                const double q = tt / AT(p.area, b, z);
                acc = second ? acc + q : acc - q;
            }
            touched = true;
        }
        if (ROUND == 1 || touched) AT(dst, b, z) = acc;
    }
}

// ---------------------------------------------------------------------------------------------------
// K7b: diff_ver_part_impl_ale (oce_ale_tracer.F90:509-855).  One warp = one tile of 32 own columns, one thread per
// column; warps are independent (no block barrier), so an SM keeps as many column recurrences in flight as shared
// memory allows (one [nlp][32] array per warp).  Every load address is known up front (level-blocked layout: a warp
// reads 256 contiguous bytes per level), only the arithmetic of the recurrences is serial:
//   pass 1 (bottom-up)  zbar_n(nz) = zbar_n(nz+1) + hnode_new(nz)  -> kept in the output slot T(nz) until pass 2 gets there  (:528-536)
//   pass 2 (top-down)   a, b, c and the right-hand side of level nz (:560-800, shortwave penetration, surface boundary
//                       condition bc_surface :1154-1195) feed the forward elimination of the Thomas algorithm
//                       directly (:828-836); c' -> shared memory, d' -> the output field T
//   pass 3 (bottom-up)  back substitution (:838-844) and T = T* + dT
// zinv1 of level nz is the same expression as zinv2 of level nz-1 and is carried over.
// dynamic smem: cp [blockDim.y][nlp][32]
// KPP = use_kpp_nonlclflx: the non-local transport of the boundary-layer scheme joins the right-hand side (:688-781)
// NT tracers per launch: the coefficients a, b, c (and with them every division and the Thomas factors c') depend on the
// thicknesses, Kv and the areas only, so a second tracer adds its right-hand side and its d' recurrence and nothing else.
struct ImplSet { const double* TS[2]; double* To[2]; int tr_num[2], tracer_id[2]; };
template <bool WIMPL, bool KPP, int NT>     // WIMPL = w_split without FCT: implicit upwind advection by Wvel_i enters a, b, c (:604-616, :668-682, :737-744)
__global__ void __launch_bounds__(64, NT == 1 ? 12 : 10) k_impl_diff(KP p, ImplSet q) {
    const int lane = threadIdx.x;
    const int tile = blockIdx.x * blockDim.y + threadIdx.y;
    const int n = tile * TILE + lane;
    if (n >= p.myN) return;
    const int b = tile * (p.nlp * TILE) + lane;
    double* __restrict__ cps = p.adf_v + b;                  // c' parked in the vertical-flux scratch (consumed by k_update, rewritten by k_lo)
    const unsigned lv = p.n_lev[n];
    const int nzmin = lv & 255, nzmax = (lv >> 8) & 255;
    const double* __restrict__ hn = p.hnode_new;
    double* ZBg = q.To[0];                                // zbar_n(z) waits in the first tracer's output slot of level z until d'(z) replaces it
    const double dt = p.dt;
    const double isredi = p.Redi ? 1.0 : 0.0;
    const double vcpw = 4.2e6;
    const double zinv = 1.0 * dt;
    const double zbot = p.n_znb[n];
    {
        double zb = zbot;
#pragma unroll 4
        for (int z = nzmax - 1; z >= nzmin; z--) { zb = zb + AT(hn, b, z); AT(ZBg, b, z) = zb; }
    }
    __syncwarp();
#define ZBL(z) ((z) >= nzmax ? zbot : AT(ZBg, b, (z)))
    // rolling window around level z: zb0..zb2 = zbar_n(z..z+2); h0, hp = hnode_new(z, z+1); znm, zn0 = Z_n(z-1), Z_n(z);
    // Kv, area, T* at z and z+1.  Everything level z+1 needs is requested at the top of iteration z, before the
    // divisions of level z, so the loads have a whole iteration to arrive.
    int z = nzmin;
    double zb0 = 0.0, zb1 = ZBL(z + 1), zb2 = ZBL(z + 2);
    double h0 = AT(hn, b, z), hp = AT(hn, b, z + 1);
    double zn0 = zb1 + h0 / 2.0, znm = 0.0;
    double kv1 = AT(p.Kv, b, z + 1), ar1 = AT(p.area, b, z + 1), kv0 = 0.0, ar0 = 0.0;
    double tm[NT], t0[NT], tp1[NT], tpm[NT];
#pragma unroll
    for (int t = 0; t < NT; t++) { tm[t] = 0.0; t0[t] = AT(q.TS[t], b, z); tp1[t] = AT(q.TS[t], b, z + 1); tpm[t] = 0.0; }
    double asv = AT(p.areasvol, b, z);
    double zinv1 = 0.0, cpm = 0.0;
    for (; z <= nzmax - 1; z++) {
        double n_zb = zbot, n_asv = 0.0, n_hp = 0.0, n_kv = 0.0, n_ar = 0.0, n_tp[NT];
#pragma unroll
        for (int t = 0; t < NT; t++) n_tp[t] = 0.0;
        if (z + 3 < nzmax) n_zb = AT(ZBg, b, z + 3);
        if (z + 1 <= nzmax - 1) n_asv = AT(p.areasvol, b, z + 1);
        if (z + 2 <= nzmax - 1) {
            n_hp = AT(hn, b, z + 2);
#pragma unroll
            for (int t = 0; t < NT; t++) n_tp[t] = AT(q.TS[t], b, z + 2);
        }
        if (z + 2 <= nzmax) { n_kv = AT(p.Kv, b, z + 2); n_ar = AT(p.area, b, z + 2); }
        const double znp = zb2 + hp / 2.0;                                   // Z_n(z+1)
        const double h = h0;
        const double rasv = 1.0 / asv;                                       // every division by areasvol(z) below: exact-division sequence div_by
        double a, bb, cc, zinv2 = 0.0;
        if (z == nzmin) {
            zinv2 = 1.0 / (zn0 - znp);
            double Ty1 = 0.0;
            if (p.Redi) {
                const double s0 = AT(p.sl3, b, z), s1 = AT(p.sl3, b, z + 1);
                Ty1 = (zn0 - zb1) * zinv2 * (s0 * s0) * AT(p.Ki, b, z) + (zb1 - znp) * zinv2 * (s1 * s1) * AT(p.Ki, b, z + 1);
            }
            Ty1 = Ty1 * isredi;
            a = 0.0;
            cc = div_by(-(kv1 + Ty1) * zinv2 * zinv * ar1, asv, rasv);
            bb = -cc + h;
            if (WIMPL) {
                double v_adv = zinv * div_by(AT(p.area, b, z), asv, rasv);
                bb = bb + AT(p.Wi, b, z) * v_adv;
                v_adv = div_by(zinv * ar1, asv, rasv);
                const double wi1 = AT(p.Wi, b, z + 1);
                bb = bb - omin(0.0, wi1) * v_adv;
                cc = cc - omax(0.0, wi1) * v_adv;
            }
        } else if (z <= nzmax - 2) {
            zinv2 = 1.0 / (zn0 - znp);
            double Ty = 0.0, Ty1 = 0.0;
            if (p.Redi) {
                const double sm1 = AT(p.sl3, b, z - 1), s0 = AT(p.sl3, b, z), s1 = AT(p.sl3, b, z + 1);
                Ty = (znm - zb0) * zinv1 * (sm1 * sm1) * AT(p.Ki, b, z - 1) + (zb0 - zn0) * zinv1 * (s0 * s0) * AT(p.Ki, b, z);
                Ty1 = (zn0 - zb1) * zinv2 * (s0 * s0) * AT(p.Ki, b, z) + (zb1 - znp) * zinv2 * (s1 * s1) * AT(p.Ki, b, z + 1);
            }
            Ty = Ty * isredi; Ty1 = Ty1 * isredi;
            a = -(kv0 + Ty) * zinv1 * zinv * div_by(ar0, asv, rasv);
            cc = div_by(-(kv1 + Ty1) * zinv2 * zinv * ar1, asv, rasv);
            bb = -a - cc + h;
            if (WIMPL) {
                double v_adv = zinv * div_by(ar0, asv, rasv);
                const double wi0 = AT(p.Wi, b, z), wi1 = AT(p.Wi, b, z + 1);
                a = a + omin(0.0, wi0) * v_adv;
                bb = bb + omax(0.0, wi0) * v_adv;
                v_adv = div_by(zinv * ar1, asv, rasv);
                bb = bb - omin(0.0, wi1) * v_adv;
                cc = cc - omax(0.0, wi1) * v_adv;
            }
        } else {
            double Ty = 0.0;
            if (p.Redi) {
                const double sm1 = AT(p.sl3, b, z - 1), s0 = AT(p.sl3, b, z);
                Ty = (znm - zb0) * zinv1 * (sm1 * sm1) * AT(p.Ki, b, z - 1) + (zb0 - zn0) * zinv1 * (s0 * s0) * AT(p.Ki, b, z);
            }
            Ty = Ty * isredi;
            a = -(kv0 + Ty) * zinv1 * zinv * div_by(ar0, asv, rasv);
            cc = 0.0;
            bb = -a + h;
            if (WIMPL) {
                const double v_adv = zinv * div_by(ar0, asv, rasv);
                const double wi0 = AT(p.Wi, b, z);
                a = a + omin(0.0, wi0) * v_adv;
                bb = bb + omax(0.0, wi0) * v_adv;
            }
        }
        double m = bb, rm;                                                   // Thomas pivot :828-835, shared by the tracers
        if (z > nzmin) m = bb - cpm * a;
        rm = 1.0 / m;
#pragma unroll
        for (int t = 0; t < NT; t++) {
            const int tr_num = q.tr_num[t];
            double tr;
            if (z == nzmin) tr = -(bb - h) * t0[t] - cc * tp1[t];
            else if (z <= nzmax - 2) tr = -a * tm[t] - (bb - h) * t0[t] - cc * tp1[t];
            else tr = -a * tm[t] - (bb - h) * t0[t];
            if (KPP && tr_num <= 2) {                                        // :688-781, fesom KPP and cvmix KPP share the formula
                const double* __restrict__ G = tr_num == 1 ? p.kgT : p.kgS;
                const double* __restrict__ B = tr_num == 1 ? p.kb2 : p.kb3;
                const double m0 = z > nzmin ? omin(AT(G, b, z) * AT(B, b, z), 1.0) : 0.0;
                const double m1 = z <= nzmax - 2 ? omin(AT(G, b, z + 1) * AT(B, b, z + 1), 1.0) : 0.0;
                double x;
                if (z == nzmin) x = -m1 * div_by(ar1, asv, rasv);
                else if (z <= nzmax - 2) x = m0 * div_by(ar0, asv, rasv) - m1 * div_by(ar1, asv, rasv);
                else x = m0 * div_by(ar0, asv, rasv);
                if (tr_num == 1) tr = tr + x * p.heat_flux[n] / vcpw * dt;
                else {
                    const double rsss = p.ref_sss_local ? AT(q.TS[t], b, 1) : p.ref_sss;     // tr_arr(1,n,2), :691
                    tr = tr - x * rsss * p.water_flux[n] * dt;
                }
            }
            if (p.use_sw_pene && tr_num == 1)
                tr = tr + (AT(p.sw3d, b, z) - div_by(AT(p.sw3d, b, z + 1) * ar1, asv, rasv)) * zinv;
            if (z == nzmin) {
                double bc = 0.0;
                if (q.tracer_id[t] == 0) {
                    const double tsurf = (tr_num == 1) ? t0[t] : AT(p.T1, b, nzmin);
                    bc = -dt * (p.heat_flux[n] / vcpw + tsurf * p.water_flux[n] * p.is_nonlinfs);
                } else if (q.tracer_id[t] == 1)
                    bc = dt * (p.virtual_salt[n] + p.relax_salt[n] - p.real_salt_flux[n] * p.is_nonlinfs);
                tr = tr + bc;
                tpm[t] = div_by(tr, m, rm);                                  // :829
            } else tpm[t] = div_by(tr - tpm[t] * a, m, rm);                  // :835
        }
        cpm = div_by(cc, m, rm);
        cps[(z - 1) * TILE] = cpm;
#pragma unroll
        for (int t = 0; t < NT; t++) AT(q.To[t], b, z) = tpm[t];
        // shift the window to level z+1
        zb0 = zb1; zb1 = zb2; zb2 = n_zb; znm = zn0; zn0 = znp; h0 = hp; hp = n_hp; zinv1 = zinv2;
        kv0 = kv1; kv1 = n_kv; ar0 = ar1; ar1 = n_ar; asv = n_asv;
#pragma unroll
        for (int t = 0; t < NT; t++) { tm[t] = t0[t]; t0[t] = tp1[t]; tp1[t] = n_tp[t]; }
    }
#undef ZBL
    __syncwarp();
    double x[NT];                                                             // :838-844
#pragma unroll
    for (int t = 0; t < NT; t++) { x[t] = tpm[t]; AT(q.To[t], b, nzmax - 1) = AT(q.TS[t], b, nzmax - 1) + x[t]; }
#pragma unroll 2
    for (int zz = nzmax - 2; zz >= nzmin; zz--) {
        const double cp = cps[(zz - 1) * TILE];
#pragma unroll
        for (int t = 0; t < NT; t++) {
            x[t] = AT(q.To[t], b, zz) - cp * x[t];
            AT(q.To[t], b, zz) = AT(q.TS[t], b, zz) + x[t];
        }
    }
}

// copy own columns of src into dst for levels nzmin..nzmax-1 (used when i_vert_diff is off)
__global__ void LB(128) k_copy_cols(KP p, double* __restrict__ dst, const double* __restrict__ src) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= p.myN) return;
    const int b = tbase(n, p.nlp);
    const unsigned lv = p.n_lev[n];
    for (int nz = (lv & 255); nz <= (int)((lv >> 8) & 255) - 1; nz++) AT(dst, b, nz) = AT(src, b, nz);
}

// relax_to_clim, oce_tracer_mod.F90:87-119 — own nodes, wet levels: T += relax2clim(n) * dt * (clim - T)
__global__ void LB(128) k_relax_clim(KP p, double* __restrict__ T, const double* __restrict__ clim, const double* __restrict__ relax2clim) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= p.myN) return;
    const int b = tbase(n, p.nlp);
    const unsigned lv = p.n_lev[n];
    const double r = relax2clim[n];
    for (int nz = (lv & 255); nz <= (int)((lv >> 8) & 255) - 1; nz++) {
        const double t = AT(T, b, nz);
        AT(T, b, nz) = t + r * p.dt * (AT(clim, b, nz) - t);
    }
}

// relax_zonal_temp, toy_channel_soufflet.F90:78-103 — own + halo nodes, levels 1..nlevels_nod2D-1, always tracer 1:
// T += dt*tau_inv * (Tclim - ((1-a) ztem(nz,nn) + a ztem(nz,nn1))); nn, nn1, a per node come from fb200_set_zonal_relax
__global__ void LB(128) k_relax_zonal(KP p, double* __restrict__ T, const double* __restrict__ clim, const double* __restrict__ ztem,
                                      const int2* __restrict__ bins, const double* __restrict__ wgt, double fac) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= p.N) return;
    const int b = tbase(n, p.nlp), L = p.nl - 1;
    const int nln = (p.n_lev[n] >> 8) & 255;
    const int2 nn = bins[n];
    const double a = wgt[n];
    const double* z0 = ztem + (size_t)(nn.x - 1) * L - 1;
    const double* z1 = ztem + (size_t)(nn.y - 1) * L - 1;
    for (int nz = 1; nz <= nln - 1; nz++) {
        const double tzon = (1.0 - a) * z0[nz] + a * z1[nz];
        AT(T, b, nz) = AT(T, b, nz) + fac * (AT(clim, b, nz) - tzon);
    }
}

// global extrema of one tracer over own wet cells: the quantity check_blowup (write_step_info.F90:220-504) and the
// min/max lines of write_step_info test on the host; with device-resident tracers only two scalars cross PCIe.
// One partial (min, max) per block; the host combines the <= 1184 partials (min/max are order-independent).
__global__ void LB(256) k_extrema(KP p, const double* __restrict__ T, double2* __restrict__ partial) {
    double lo = 1.0e300, hi = -1.0e300;
    for (int n = blockIdx.x * blockDim.x + threadIdx.x; n < p.myN; n += gridDim.x * blockDim.x) {
        const int b = tbase(n, p.nlp);
        const unsigned lv = p.n_lev[n];
        for (int nz = (lv & 255); nz <= (int)((lv >> 8) & 255) - 1; nz++) { const double t = AT(T, b, nz); lo = dmin(lo, t); hi = dmax(hi, t); }
    }
    __shared__ double slo[8], shi[8];
    for (int o = 16; o > 0; o >>= 1) { lo = dmin(lo, __shfl_down_sync(0xffffffffu, lo, o)); hi = dmax(hi, __shfl_down_sync(0xffffffffu, hi, o)); }
    if ((threadIdx.x & 31) == 0) { slo[threadIdx.x >> 5] = lo; shi[threadIdx.x >> 5] = hi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; w++) { lo = dmin(lo, slo[w]); hi = dmax(hi, shi[w]); }
        partial[blockIdx.x] = make_double2(lo, hi);
    }
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
static KP make_kp(TracerCtx& c, int tr, int wslot = 0) {
    KP p;
    p.nl = c.nl; p.nlp = c.nlp; p.myN = c.myN; p.N = c.N; p.myE = c.myE; p.Eall = c.Eall; p.myEd = c.myEd; p.maxdeg = c.maxdeg; p.maxnie = c.maxnie;
    p.dt = c.opt.dt; p.ph = c.opt.ph; p.pv = c.opt.pv; p.eps_ab = c.opt.epsilon; p.is_nonlinfs = c.opt.is_nonlinfs;
    p.hor = c.opt.tra_adv_hor; p.ver = c.opt.tra_adv_ver; p.w_split = c.opt.w_split; p.Redi = c.opt.Redi; p.use_sw_pene = c.opt.use_sw_pene;
    p.do_hor_diff = c.opt.do_hor_diff; p.tracer_id = c.opt.tracer_id[tr]; p.tr_num = tr + 1;
    p.edge = c.edge.p; p.edgeho = c.edgeho.p; p.ne_slot = c.ne_slot.p; p.nie_slot = c.nie_slot.p; p.el_nb = c.el_nb.p; p.el_lev = c.el_lev.p;
    p.tile_deg = c.tile_deg.p; p.ns_i = c.ns_i.p; p.ns_m = c.ns_m.p; p.ns_p = c.ns_p.p; p.ns_d1 = c.ns_d1.p; p.ns_d2 = c.ns_d2.p;
    p.el_g = c.el_g.p; p.el_area = c.el_area.p; p.n_lev = c.n_lev.p; p.n_nbl = c.n_nbl.p; p.n_znb = c.n_znb.p;
    p.area = c.area.p; p.areasvol = c.areasvol.p; p.UV = c.UV.p; p.helem = c.helem.p; p.W = c.W.p; p.We = c.opt.w_split ? c.We.p : c.W.p; p.Wi = c.Wi.p;
    p.hnode = c.hnode.p; p.hnode_new = c.hnode_new.p; p.Z3 = c.Z3.p; p.zbar3 = c.zbar3.p; p.Kv = c.Kv.p; p.Ki = c.Ki.p;
    p.sl1 = c.sl1.p; p.sl2 = c.sl2.p; p.sl3 = c.sl3.p; p.sw3d = c.sw3d.p;
    p.kgT = c.kgT.p; p.kgS = c.kgS.p; p.kb2 = c.kb2.p; p.kb3 = c.kb3.p; p.ref_sss = c.opt.ref_sss; p.ref_sss_local = c.opt.ref_sss_local;
    p.heat_flux = c.heat_flux.p; p.water_flux = c.water_flux.p; p.virtual_salt = c.virtual_salt.p; p.relax_salt = c.relax_salt.p; p.real_salt_flux = c.real_salt_flux.p;
    const size_t nf = (size_t)c.Ntiles * c.nlp * TILE;
    p.T = c.T.p + nf * tr; p.Told = c.Told.p + nf * tr; p.T1 = c.T.p;
    // work fields: one set per tracer of a pass (work slot wslot < c.nwork)
    const size_t ef = (size_t)c.Etiles * c.nlp * TILE, df = (size_t)(c.Edtiles + 1) * c.nlp * TILE, w = (size_t)wslot;   // edge fields: + 1 never-written tile
    p.zero_edge = c.Edtiles * c.nlp * TILE;
    p.txy = c.txy.p + ef * w; p.fpm = c.fpm.p + nf * w; p.trz = c.trz.p + nf * w; p.LO = c.LO.p + nf * w; p.adf_h = c.adf_h.p + df * w;
    p.adf_v = c.adf_v.p + nf * w; p.tstar = c.tstar.p + nf * w; p.eflx = c.eflx.p + df * w; p.redi_v = c.redi_v.p ? c.redi_v.p + nf * w : nullptr;
    return p;
}

void tracer_extrema(TracerCtx& c, int tr, double* tmin, double* tmax) {
    *tmin = 0.0; *tmax = 0.0;
    if (!fb_ok()) return;
    if (!c.committed || tr < 0 || tr >= c.ntr) { fb_set_error(27, "fb200_tracer_extrema: not committed or no such tracer"); return; }
    fb_copy_sync();
    cudaStream_t st = c.stream = fb_stream();
    const int nb = std::min(148 * 8, (c.myN + 255) / 256);
    if (c.extrema_part.n < (size_t)nb) c.extrema_part.alloc(nb);          // once: a monitoring call per step must not allocate and free every time
    DevBuf<double2>& part = c.extrema_part;
    KP p = make_kp(c, tr);
    { ProfScope ps_("k_extrema", st); k_extrema<<<nb, 256, 0, st>>>(p, p.T, part.p); LAUNCH_CHECK(); }
    std::vector<double2> h(nb);
    CUDA_CHECK(cudaMemcpyAsync(h.data(), part.p, nb * sizeof(double2), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    double lo = 1.0e300, hi = -1.0e300;
    for (const double2& v : h) { lo = v.x < lo ? v.x : lo; hi = v.y > hi ? v.y : hi; }
    *tmin = lo; *tmax = hi;
}

// ---------------------------------------------------------------------------------------------------
// fb200_tracer_extrema_noalloc: the extrema scan for callers that cannot call cudaMalloc (code under CUDA graph capture,
// tight serving loops, harnesses that forbid allocation).  No counterpart in the reference.
// The per-block partials of k_extrema and the final (min, max) live in a file-scope device array of fixed
// size instead of a per-call allocation: FB200_EXTREMA_MAX_BLOCKS partials (+ 1 result slot) of 16 bytes = 32 784 bytes that
// every GPU that loads the library holds until the process exits.  In host code the array's NAME is only a host-side
// stand-in: the device address is looked up with cudaGetSymbolAddress once per device (with that device current) and it
// is that pointer the kernels receive.  Every call shares the array, so two calls that overlap on the GPU (different
// streams, different host threads) race on it; calls on one stream are ordered.  Nothing in it has to start at zero: each
// call overwrites partials [0, nblocks) and the result slot before reading them, older contents are never read.
#define FB200_EXTREMA_MAX_BLOCKS 2048
#define FB200_MAX_DEVICES 64
static __device__ __align__(256) double2 g_extrema_scratch[FB200_EXTREMA_MAX_BLOCKS + 1];
static double2* g_extrema_scratch_ptr[FB200_MAX_DEVICES];      // device address of the array on each device, filled on first use

// one block: combines the nb partials of k_extrema in a fixed order (min/max are order-independent anyway) into slot[nb_max]
__global__ void LB(256) k_extrema_final(double2* __restrict__ scratch, int nb, int result_slot) {
    double lo = 1.0e300, hi = -1.0e300;
    for (int i = threadIdx.x; i < nb; i += blockDim.x) { const double2 v = scratch[i]; lo = dmin(lo, v.x); hi = dmax(hi, v.y); }
    __shared__ double slo[8], shi[8];
    for (int o = 16; o > 0; o >>= 1) { lo = dmin(lo, __shfl_down_sync(0xffffffffu, lo, o)); hi = dmax(hi, __shfl_down_sync(0xffffffffu, hi, o)); }
    if ((threadIdx.x & 31) == 0) { slo[threadIdx.x >> 5] = lo; shi[threadIdx.x >> 5] = hi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; w++) { lo = dmin(lo, slo[w]); hi = dmax(hi, shi[w]); }
        scratch[result_slot] = make_double2(lo, hi);
    }
}

void tracer_extrema_noalloc(TracerCtx& c, int tr, int nblocks, double* tmin, double* tmax) {
    *tmin = 0.0; *tmax = 0.0;
    if (!fb_ok()) return;
    if (!c.committed || tr < 0 || tr >= c.ntr) { fb_set_error(27, "fb200_tracer_extrema_noalloc: not committed or no such tracer"); return; }
    if (nblocks > FB200_EXTREMA_MAX_BLOCKS) {
        fb_set_error(28, "fb200_tracer_extrema_noalloc: nblocks = " + std::to_string(nblocks) + " exceeds the static scratch of " +
                             std::to_string(FB200_EXTREMA_MAX_BLOCKS) + " per-block partials");
        return;
    }
    int dev = -1;
    CUDA_CHECK(cudaGetDevice(&dev));
    if (!fb_ok()) return;
    if (dev < 0 || dev >= FB200_MAX_DEVICES) { fb_set_error(28, "fb200_tracer_extrema_noalloc: device ordinal beyond the per-device address table"); return; }
    if (!g_extrema_scratch_ptr[dev]) {                          // once per device, with that device current
        double2* ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_extrema_scratch);
        if (e != cudaSuccess || !ptr) { fb_set_error(100 + (int)e, std::string("cudaGetSymbolAddress(g_extrema_scratch): ") + cudaGetErrorString(e)); return; }
        g_extrema_scratch_ptr[dev] = ptr;
    }
    double2* scratch = g_extrema_scratch_ptr[dev];
    fb_copy_sync();
    cudaStream_t st = c.stream = fb_stream();
    const int want = (c.myN + 255) / 256;
    const int nb = nblocks > 0 ? std::min(nblocks, want) : std::min(148 * 8, want);     // k_extrema strides over the nodes by its grid size
    KP p = make_kp(c, tr);
    { ProfScope ps_("k_extrema", st); k_extrema<<<nb, 256, 0, st>>>(p, p.T, scratch); LAUNCH_CHECK(); }
    { ProfScope ps_("k_extrema_final", st); k_extrema_final<<<1, 256, 0, st>>>(scratch, nb, FB200_EXTREMA_MAX_BLOCKS); LAUNCH_CHECK(); }
    double2 h = make_double2(0.0, 0.0);                          // 16 bytes on the host stack
    CUDA_CHECK(cudaMemcpyAsync(&h, scratch + FB200_EXTREMA_MAX_BLOCKS, sizeof(double2), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    if (!fb_ok()) return;
    *tmin = h.x; *tmax = h.y;
}

// Output means of the resident tracers: update_means (io_meandata.F90:768-794) adds the stream's field to local_values in
// the stream's precision (real64, or real32 with the field rounded to real32 first) and counts; output (:894-900) divides
// by real(addcounter) in that precision and clears.  The whole tiled field is accumulated (dry levels, halo and padding
// included: the reference adds ptr3(1:nl-1, 1:myDim_nod2D) including the entries below the bottom).
template <class A>
__global__ void LB(256) k_mean_acc(A* __restrict__ acc, const double* __restrict__ T, size_t count) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (size_t)gridDim.x * blockDim.x) acc[i] = acc[i] + (A)T[i];
}
template <class A>
__global__ void LB(256) k_mean_out(A* __restrict__ acc, double* __restrict__ out, size_t count, int addcounter) {
    const A d = (A)addcounter;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (size_t)gridDim.x * blockDim.x) {
        out[i] = (double)(acc[i] / d);
        acc[i] = (A)0;
    }
}
void tracer_mean_accumulate(TracerCtx& c, int tr, int accuracy) {
    if (!fb_ok()) return;
    if (!c.committed || tr < 0 || tr >= c.ntr || tr >= 8) { fb_set_error(27, "fb200_mean_accumulate: not committed or no such tracer"); return; }
    if (accuracy != 4 && accuracy != 8) { fb_set_error(27, "fb200_mean_accumulate: accuracy must be 4 (real32) or 8 (real64)"); return; }
    if (c.mean_count[tr] > 0 && c.mean_prec[tr] != accuracy) { fb_set_error(27, "fb200_mean_accumulate: precision changed inside an averaging interval"); return; }
    fb_copy_sync();
    cudaStream_t st = c.stream = fb_stream();
    const size_t nf = (size_t)c.Ntiles * c.nlp * TILE;
    c.mean_prec[tr] = accuracy;
    if (accuracy == 8) {
        if (!c.mean8[tr].p) c.mean8[tr].alloc(nf);             // alloc clears
        ProfScope ps_("k_mean_acc", st); k_mean_acc<double><<<148 * 8, 256, 0, st>>>(c.mean8[tr].p, c.T.p + nf * tr, nf); LAUNCH_CHECK();
    } else {
        if (!c.mean4[tr].p) c.mean4[tr].alloc(nf);
        ProfScope ps_("k_mean_acc", st); k_mean_acc<float><<<148 * 8, 256, 0, st>>>(c.mean4[tr].p, c.T.p + nf * tr, nf); LAUNCH_CHECK();
    }
    c.mean_count[tr]++;
}
// mean_out: (nl-1, myDim_nod2D) real64 or real32, whichever the interval was accumulated in
void tracer_mean_fetch(TracerCtx& c, int tr, void* mean_out) {
    if (!fb_ok()) return;
    if (!c.committed || tr < 0 || tr >= c.ntr || tr >= 8 || c.mean_count[tr] == 0) { fb_set_error(27, "fb200_mean_fetch: nothing accumulated for this tracer"); return; }
    fb_copy_sync();
    cudaStream_t st = c.stream = fb_stream();
    const size_t nf = (size_t)c.Ntiles * c.nlp * TILE;
    const int L = c.nl - 1;
    double* scratch = c.tstar.p;                                // work field of the step, free between steps
    if (c.mean_prec[tr] == 8) {
        { ProfScope ps_("k_mean_out", st); k_mean_out<double><<<148 * 8, 256, 0, st>>>(c.mean8[tr].p, scratch, nf, c.mean_count[tr]); LAUNCH_CHECK(); }
        ctx_pull_node_field(c, scratch, L, c.myN, (double*)mean_out, 1, 0, true);
    } else {
        { ProfScope ps_("k_mean_out", st); k_mean_out<float><<<148 * 8, 256, 0, st>>>(c.mean4[tr].p, scratch, nf, c.mean_count[tr]); LAUNCH_CHECK(); }
        std::vector<double> h((size_t)L * c.myN);
        ctx_pull_node_field(c, scratch, L, c.myN, h.data(), 1, 0, true);
        float* o = (float*)mean_out;
        for (size_t i = 0; i < h.size(); i++) o[i] = (float)h[i];      // exact: the values are real32 numbers
    }
    c.mean_count[tr] = 0;
}

static bool g_attr_done = false;
static void set_smem_attrs() {
    if (g_attr_done) return;
    const int big = 200 * 1024;
    CUDA_CHECK(cudaFuncSetAttribute(k_vert_impl, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    CUDA_CHECK(cudaFuncSetAttribute(k_fct_bounds<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    CUDA_CHECK(cudaFuncSetAttribute(k_fct_bounds<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    g_attr_done = true;
}

// solve_tracers_ale (oce_ale_tracer.F90:101-199) for one partition, everything device resident.
// tracer_step_begin: option checks + start event; tracer_step_tracer: every kernel of ONE tracer on the library stream
// (the tracers are independent of each other: fb200_solve_tracers_ale overlaps the host copies of one with the kernels
// of the other); tracer_step_end: stop event + synchronise.
bool tracer_step_begin(TracerCtx& c) {
    if (!fb_ok()) return false;
    if (!c.committed) { fb_set_error(20, "fb200_solve_tracers before fb200_init_commit"); return false; }
    if (c.opt.dt <= 0.0) { fb_set_error(24, "dt not set (fb200_set_options)"); return false; }
    if (c.nlp > 63) { fb_set_error(25, "nl > 63 is not supported (64-bit level masks; fb200_init_commit applies the same limit)"); return false; }
    if (comm_nranks() > 1 && !c.halo_nod) { fb_set_error(26, "more than one rank but fb200_init_halo was not called"); return false; }
    set_smem_attrs();
    fb_copy_sync();                                             // state pushed with fb200_push_state_early must have landed
    c.stream = fb_stream();
    CUDA_CHECK(cudaEventRecord(c.ev0, c.stream));
    return fb_ok();
}

// which phases run the first-generation kernels of this file instead of the column kernels of tracer_kernels2.cu:
// FB200_V1 = comma-separated list of {grad, edge, lo, bounds, diff, update} or "all" (A/B measurements; results are
// bit-identical either way).  Default: none.
static unsigned v1_phases() {
    static int cached = -1;
    if (cached >= 0) return (unsigned)cached;
    unsigned m = 0;
    const char* e = getenv("FB200_V1");
    if (e) {
        std::string s(e);
        const char* names[6] = {"grad", "edge", "lo", "bounds", "diff", "update"};
        if (s.find("all") != std::string::npos) m = 63;
        for (int k = 0; k < 6; k++) if (s.find(names[k]) != std::string::npos) m |= 1u << k;
    }
    cached = (int)m;
    return m;
}
enum { PH_GRAD = 1, PH_EDGE = 2, PH_LO = 4, PH_BOUNDS = 8, PH_DIFF = 16, PH_UPDATE = 32 };

// number of tracers the pass starting at tracer t0 handles together (1 or 2).  Tracers of a group must not read each
// other inside the pass: bc_surface reads tracer 1 for every tracer with id 0 (oce_ale_tracer.F90:1170), and
// relax_zonal_temp touches tracer 1 after every pass of the tracer loop — in both cases the tracers go one by one.
int tracer_group_size(const TracerCtx& c, int t0) {
    if (c.nwork < 2 || t0 + 1 >= c.ntr || c.opt.relax_zonal) return 1;
    if (getenv("FB200_NT1")) return 1;
    for (int t = t0; t < t0 + 2; t++) if (t > 0 && c.opt.tracer_id[t] == 0) return 1;
    return 2;
}

// every kernel of the tracers t0 .. t0+nt-1 (nt <= MAXNT) on the library stream; work fields of tracer t0+k live in work
// slot k
void tracer_step_group(TracerCtx& c, int t0, int nt) {
    if (!fb_ok()) return;
    cudaStream_t st = c.stream;
    const size_t nf = (size_t)c.Ntiles * c.nlp * TILE;
    const int gN64 = (c.myN + 63) / 64, gE = (c.myE + 127) / 128, gNall = (c.N + 127) / 128;
    const int LPB = c.nl - 1 <= 48 ? 3 : 4;                     // k_fct_bounds: one block = one tile x the whole column, <= 16 level rows
    const dim3 blk(TILE, (c.nl - 1 + LPB - 1) / LPB);
    // first-generation kernels without block-wide exchange: 8 level rows per block (LPT * 8 = 24 levels), gridDim.y level chunks
    const dim3 blk8(TILE, 8);
    const int nchunk = (c.nl - 1 + LPT * 8 - 1) / (LPT * 8);
    const dim3 gN8(c.myNtiles, nchunk);
    const size_t smt = (size_t)c.nlp * TILE * sizeof(double);   // one [nlp][32] shared array
    const size_t sm_64 = (size_t)c.nlp * 64 * sizeof(double);
    const bool multi = comm_nranks() > 1 && c.halo_nod != nullptr;
    const bool fct = c.opt.tra_adv_lim == 0;
    // gradients: the per-tracer kernels of this file (they run at the HBM roofline; the two-tracer variant is no faster) unless FB200_V2_GRAD
    const unsigned v1 = v2_usable(c) ? (v1_phases() | (getenv("FB200_V2_GRAD") ? 0u : (unsigned)PH_GRAD)) : 63u;
#define KL(name, ...) do { ProfScope ps_(name, st); __VA_ARGS__; LAUNCH_CHECK(); } while (0)
#define EACH(...) for (int k_ = 0; k_ < nt; k_++) { KP& p = pk[k_]; (void)p; __VA_ARGS__; }
    if (c.opt.Redi && !c.redi_v.p) c.redi_v.alloc(nf * c.nwork);
    KP pk[MAXNT];
    TrSet s;
    for (int k = 0; k < nt; k++) {
        pk[k] = make_kp(c, t0 + k, k);
        s.T[k] = pk[k].T; s.Told[k] = pk[k].Told; s.LO[k] = pk[k].LO; s.adf_h[k] = pk[k].adf_h; s.adf_v[k] = pk[k].adf_v;
        s.tstar[k] = pk[k].tstar; s.eflx[k] = pk[k].eflx; s.trz[k] = pk[k].trz; s.txy[k] = pk[k].txy; s.fpm[k] = pk[k].fpm;
        s.redi_v[k] = pk[k].redi_v; s.tr_num[k] = t0 + k + 1;
    }
    for (int k = nt; k < MAXNT; k++) {
        s.T[k] = s.T[0]; s.Told[k] = s.Told[0]; s.LO[k] = s.LO[0]; s.adf_h[k] = s.adf_h[0]; s.adf_v[k] = s.adf_v[0]; s.tstar[k] = s.tstar[0];
        s.eflx[k] = s.eflx[0]; s.trz[k] = s.trz[0]; s.txy[k] = s.txy[0]; s.fpm[k] = s.fpm[0]; s.redi_v[k] = s.redi_v[0]; s.tr_num[k] = s.tr_num[0];
    }
    const KP& p0 = pk[0];                                       // the tracer-independent part is the same in every pk[k]
    v2_ab2(c, s, nt, st);                                                              // init_tracers_AB, oce_tracer_mod.F90:64
    if (v1 & PH_GRAD) EACH(KL("k_grad_elem", k_grad_elem<<<gE, 128, 0, st>>>(p, p.Told)))
    else v2_grad_elem(c, p0, s, nt, 0, st);
    if (multi) EACH(halo_exchange_tiled(*c.halo_elem, (double*)p.txy, c.nlp, 2, st))  // exchange_elem(tr_xy), oce_tracer_mod.F90:68,72
    if (v1 & PH_EDGE) {
        EACH(if (fct) KL("k_ho_edge", k_ho_edge<true><<<dim3(c.Edtiles, nchunk), blk8, 0, st>>>(p));
             else KL("k_ho_edge", k_ho_edge<false><<<dim3(c.Edtiles, nchunk), blk8, 0, st>>>(p)))
    } else v2_edge_adv(c, p0, s, nt, fct, st);
    if (c.opt.tra_adv_hor != 2 && c.n_fb > 0) {
        const int gfb = (c.n_fb + 127) / 128;
        EACH(if (fct) KL("k_ho_edge_fb", k_ho_edge_fb<true><<<gfb, 128, 0, st>>>(p, c.fb_list.p, c.n_fb));
             else KL("k_ho_edge_fb", k_ho_edge_fb<false><<<gfb, 128, 0, st>>>(p, c.fb_list.p, c.n_fb)))
    }
    if (v1 & PH_LO) {
#define LO_CASE(V) case V: if (fct) KL("k_lo", k_lo<V, true><<<gN8, blk8, 0, st>>>(p)); else KL("k_lo", k_lo<V, false><<<gN8, blk8, 0, st>>>(p)); break;
        EACH(switch (c.opt.tra_adv_ver) { LO_CASE(0) LO_CASE(1) LO_CASE(2) default: LO_CASE(3) })
#undef LO_CASE
    } else v2_lo(c, p0, s, nt, fct, st);
    if (fct) {
        if (c.opt.w_split) EACH(KL("k_vert_impl", k_vert_impl<<<gN64, 64, 2 * sm_64, st>>>(p, p.LO, p.Wi)))
        if (multi) EACH(halo_exchange_tiled(*c.halo_nod, p.LO, c.nlp, 1, st))          // exchange_nod(fct_LO), oce_adv_tra_driver.F90:134
        if (v1 & PH_BOUNDS) {
            EACH(if (LPB == 3) KL("k_fct_bounds", k_fct_bounds<3><<<c.myNtiles, blk, 2 * smt, st>>>(p));
                 else KL("k_fct_bounds", k_fct_bounds<4><<<c.myNtiles, blk, 2 * smt, st>>>(p)))
        } else v2_fct_bounds(c, p0, s, nt, st);
        if (multi) EACH(halo_exchange_tiled(*c.halo_nod, (double*)p.fpm, c.nlp, 2, st)) // exchange_nod(fct_plus, fct_minus), oce_adv_tra_fct.F90:279
    }
    if (c.opt.do_hor_diff || c.opt.Redi) {                      // tracer_gradient_elements(tr_arr), oce_tracer_mod.F90:79
        if (v1 & PH_GRAD) EACH(KL("k_grad_elem", k_grad_elem<<<gE, 128, 0, st>>>(p, p.T)))
        else v2_grad_elem(c, p0, s, nt, 1, st);
    }
    if (c.opt.do_hor_diff && c.opt.Redi) {
        if (v1 & PH_GRAD) EACH(KL("k_grad_z", k_grad_z<<<gNall, 128, 0, st>>>(p, p.T)))
        else v2_grad_z(c, p0, s, nt, st);
    }
    if (c.opt.do_hor_diff) {
        if (v1 & PH_DIFF) EACH(KL("k_diff_edge", k_diff_edge<<<dim3(c.Edtiles, nchunk), blk8, 0, st>>>(p)))
        else v2_edge_diff(c, p0, s, nt, st);
    }
    if (c.opt.Redi)                                                                     // diff_ver_part_redi_expl, oce_ale_tracer.F90:288
        EACH(KL("k_redi_expl", k_redi_expl<<<(c.myN + 127) / 128, 128, 0, st>>>(p, p.tstar, (double*)p.redi_v)))
    if (v1 & PH_UPDATE) {
        EACH(if (c.opt.i_vert_diff) {
                 if (fct) KL("k_update", k_update<true, false><<<gN8, blk8, 0, st>>>(p));
                 else KL("k_update", k_update<false, false><<<gN8, blk8, 0, st>>>(p));
             } else {
                 if (fct) KL("k_update", k_update<true, true><<<gN8, blk8, 0, st>>>(p));
                 else KL("k_update", k_update<false, true><<<gN8, blk8, 0, st>>>(p));
             })
    } else v2_update(c, p0, s, nt, fct, !c.opt.i_vert_diff, st);
    {   // tr_arr_old = tr_arr (:274); the fields of the group are contiguous
        ProfScope ps_(nt == 2 ? "copy_told_x2" : "copy_told", st);
        CUDA_CHECK(cudaMemcpyAsync(pk[0].Told, pk[0].T, nf * nt * sizeof(double), cudaMemcpyDeviceToDevice, st));
    }
    if (c.opt.i_vert_diff) {                                    // diff_ver_part_impl_ale: the tracers of the pass share the coefficients
        const bool wimpl = c.opt.w_split && !fct, kpp = c.opt.use_kpp_nonlclflx && c.kgT.p;
        static const bool one_by_one = getenv("FB200_NT_IMPL") && getenv("FB200_NT_IMPL")[0] == '1';
        const int step = (nt == 2 && !one_by_one) ? 2 : 1;
        for (int k0 = 0; k0 < nt; k0 += step) {
            ImplSet q;
            for (int t = 0; t < 2; t++) {
                const int kk = k0 + (t < step ? t : 0);
                q.TS[t] = pk[kk].tstar; q.To[t] = pk[kk].T; q.tr_num[t] = t0 + kk + 1; q.tracer_id[t] = c.opt.tracer_id[t0 + kk];
            }
            const KP& p = pk[k0];                               // c' scratch: the first tracer's vertical-flux field
            const dim3 gI((c.myNtiles + 1) / 2), bI(TILE, 2);
#define IMPL(W, K) do { if (step == 2) KL("k_impl_diff_x2", k_impl_diff<W, K, 2><<<gI, bI, 0, st>>>(p, q)); else KL("k_impl_diff", k_impl_diff<W, K, 1><<<gI, bI, 0, st>>>(p, q)); } while (0)
            if (wimpl && kpp) IMPL(true, true); else if (wimpl) IMPL(true, false); else if (kpp) IMPL(false, true); else IMPL(false, false);
#undef IMPL
        }
    }
    for (int k = 0; k < nt; k++) {
        KP& p = pk[k];
        const int tr = t0 + k;
        if (!c.opt.i_vert_diff) KL("k_copy_cols", k_copy_cols<<<(c.myN + 127) / 128, 128, 0, st>>>(p, p.T, p.tstar));      // T = T*, no implicit solve
        if (c.opt.smooth_bh_tra) {                                                          // diff_part_bh, oce_ale_tracer.F90:321-323
            if (!c.bh_tmp.p) c.bh_tmp.alloc(nf);
            const int gB = (c.myN + 127) / 128;
            KL("k_bh", k_bh<1><<<gB, 128, 0, st>>>(p, p.T, c.bh_tmp.p, c.opt.gamma0_tra, c.opt.gamma1_tra, c.opt.gamma2_tra));
            if (multi) halo_exchange_tiled(*c.halo_nod, c.bh_tmp.p, c.nlp, 1, st);          // exchange_nod(temporary_ttf), :1121
            KL("k_bh", k_bh<2><<<gB, 128, 0, st>>>(p, c.bh_tmp.p, p.T, c.opt.gamma0_tra, c.opt.gamma1_tra, c.opt.gamma2_tra));
        }
        if (c.opt.relax_zonal)                                                              // relax_zonal_temp: tracer 1 on every pass, oce_ale_tracer.F90:150-151
            KL("k_relax_zonal", k_relax_zonal<<<gNall, 128, 0, st>>>(p, c.T.p, c.Tclim.p, c.zon_ztem.p, c.zon_nn.p, c.zon_a.p, c.opt.dt * c.zon_fac));
        else if (c.opt.clim_relax > 1.0e-8 && tr < 2 && c.relax2clim.p)                     // relax_to_clim(tr_num), :152-154
            KL("k_relax_clim", k_relax_clim<<<(c.myN + 127) / 128, 128, 0, st>>>(p, p.T, tr == 0 ? c.Tclim.p : c.Sclim.p, c.relax2clim.p));
        if (multi) halo_exchange_tiled(*c.halo_nod, p.T, c.nlp, 1, st);                    // exchange_nod(tr_arr), oce_ale_tracer.F90:155
        if (c.opt.clamp_salt && tr == 1) KL("k_clamp_salt", k_clamp_salt<<<gNall, 128, 0, st>>>(p, p.T));   // tr_arr(:,:,2) only, :176-198
    }
#undef EACH
#undef KL
}
void tracer_step_tracer(TracerCtx& c, int tr) { tracer_step_group(c, tr, 1); }

void tracer_step_end(TracerCtx& c) {
    CUDA_CHECK(cudaEventRecord(c.ev1, c.stream));
    CUDA_CHECK(cudaStreamSynchronize(c.stream));
    if (fb_ok()) CUDA_CHECK(cudaEventElapsedTime(&c.last_ms, c.ev0, c.ev1));
}

void tracer_step_all(TracerCtx& c) {
    if (!tracer_step_begin(c)) return;
    for (int tr = 0; tr < c.ntr;) { const int nt = tracer_group_size(c, tr); tracer_step_group(c, tr, nt); tr += nt; }
    tracer_step_end(c);
}

// debug access to work fields (tiled) by name
const double* tracer_debug_field(TracerCtx& c, const char* name, int* kind, int* ncomp, int* comp) {
    std::string s(name);
    *kind = 0; *ncomp = 1; *comp = 0;             // kind: 0 node field, 1 edge field, 2 element field
    // work fields of the LAST tracer of the step: its work slot is its position in the last pass
    size_t w = 0;
    for (int t0 = 0; t0 < c.ntr;) { const int nt = tracer_group_size(c, t0); w = (size_t)(nt - 1); t0 += nt; }
    const size_t nf = (size_t)c.Ntiles * c.nlp * TILE * w, ef = (size_t)c.Etiles * c.nlp * TILE * w, df = (size_t)(c.Edtiles + 1) * c.nlp * TILE * w;
    if (s == "fct_LO") return c.LO.p + nf;
    if (s == "adf_v") return c.adf_v.p + nf;
    if (s == "fct_plus") { *ncomp = 2; return (const double*)(c.fpm.p + nf); }
    if (s == "fct_minus") { *ncomp = 2; *comp = 1; return (const double*)(c.fpm.p + nf); }
    if (s == "tstar") return c.tstar.p + nf;
    if (s == "adf_h") { *kind = 1; return c.adf_h.p + df; }
    if (s == "tr_xy_x") { *kind = 2; *ncomp = 2; return (const double*)(c.txy.p + ef); }
    if (s == "tr_xy_y") { *kind = 2; *ncomp = 2; *comp = 1; return (const double*)(c.txy.p + ef); }
    if (s == "helem") { *kind = 2; return c.helem.p; }
    if (s == "hnode") return c.hnode.p;
    if (s == "hnode_new") return c.hnode_new.p;
    if (s == "Z_3d_n") return c.Z3.p;
    if (s == "Wvel") { *kind = 3; return c.W.p; }                  // kind 3: node field with nl levels
    if (s == "Wvel_e") { *kind = 3; return c.We.p; }
    if (s == "Wvel_i") { *kind = 3; return c.Wi.p; }
    if (s == "zbar_3d_n") { *kind = 3; return c.zbar3.p; }
    return nullptr;
}

// ---------------------------------------------------------------------------------------------------
// self-test of div_by against IEEE division on pseudo-random operand pairs spanning the magnitudes of the tracer
// path — even samples: fluxes 1e-30..1e12 over areas 1e4..1e12; odd samples: tracer differences / thickness sums
// 1e-18..1e12 over layer thicknesses, level-depth differences and matrix pivots 1e-6..1e12; every 7th sample divides by the
// constants 3 and 6 of the reconstruction formulas — plus exact-quotient and tiny/zero numerators
__global__ void k_selftest_div(long long n, unsigned long long* mismatches) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long stride = (long long)gridDim.x * blockDim.x;
    unsigned long long bad = 0;
    for (; i < n; i += stride) {
        unsigned long long h = (unsigned long long)i * 0x9E3779B97F4A7C15ull + 0x632BE59BD9B4E019ull;
        h ^= h >> 29; h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 32;
        unsigned long long g = h * 0x94D049BB133111EBull; g ^= g >> 31;
        const double mx = 1.0 + (double)(h & 0xFFFFFFFFFFFFFull) * 2.220446049250313e-16;      // [1,2)
        const double mb = 1.0 + (double)(g & 0xFFFFFFFFFFFFFull) * 2.220446049250313e-16;
        int ex = (int)((h >> 52) % 140) - 100, eb = (int)((g >> 52) % 28) + 12;
        if (i & 1) { ex = (int)((h >> 52) % 100) - 60; eb = (int)((g >> 52) % 61) - 20; }
        double x = ldexp(mx, ex), b = ldexp(mb, eb);
        if (i % 7 == 3) b = (i & 8) ? 3.0 : 6.0;
        if ((i & 15) == 0) x = b * (double)(int)(h >> 58);          // exact quotients
        if ((i & 1023) == 1) x = 0.0;
        if (g >> 63) x = -x;
        const double rb = 1.0 / b;
        if (div_by(x, b, rb) != x / b) bad++;
    }
    if (bad) atomicAdd(mismatches, bad);
}
extern "C" void fb200_selftest_div(long long* n, long long* mismatches) {
    unsigned long long* d = nullptr;
    *mismatches = -1;
    if (cudaMalloc(&d, sizeof(unsigned long long)) != cudaSuccess) { fb_set_error(12, "no CUDA device: libfesom_b200 has no CPU fallback"); return; }
    cudaMemset(d, 0, sizeof(unsigned long long));
    k_selftest_div<<<148 * 8, 256>>>(*n, d); LAUNCH_CHECK();
    unsigned long long h = 0;
    CUDA_CHECK(cudaMemcpy(&h, d, sizeof(h), cudaMemcpyDeviceToHost));
    cudaFree(d);
    *mismatches = (long long)h;
}

// ===================================================================================================
// Device-side producers of the tracer step's inputs: compute_hbar_ale (oce_ale.F90:1585-1676),
// vert_vel_ale (:1692-2204, linfs / plain zlevel / zstar) and update_thickness_ale (:800-993), restated after the
// host-side mirror fesom2_b200/host/ale.cpp, which the GPU tests compare them with bit for bit.  With them resident the
// host sends UV (after update_vel) and the surface fluxes only; Wvel*, hnode_new, helem, hnode, zbar_3d_n and Z_3d_n never
// cross PCIe.  Edge loops become gathers in ascending edge id per node (ne_slot order), like the tracer kernels.
// With several partitions the reference's four exchanges (ssh_rhs_old :1654, hbar :1663, Wvel and hnode_new :2135-2136) run
// over NCCL between the kernels; Wvel_e / Wvel_i of the halo nodes, which the reference recomputes from the exchanged
// Wvel and hnode_new (:2140-2203, same arithmetic on the same numbers), are exchanged as well.  Under linfs the reference
// leaves partial edge sums in the halo entries of ssh_rhs_old (never read); here they are zero.
__device__ __forceinline__ int node_of_base(int base, int nlp) { return (base / (nlp * TILE)) * TILE + base % (nlp * TILE); }

// per edge: c1 + c2 of compute_hbar_ale (:1612-1640), the transport through the edge's two segments summed over the column
__global__ void LB(128) k_ale_edge_sum(KP p, double* __restrict__ esum) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= p.myEd) return;
    const EdgeRec r = p.edge[e];
    double c1 = 0.0, c2 = 0.0;
    for (int nz = 1; nz <= p.nl - 1; nz++)
        if ((r.m1 >> nz) & 1ull) { const double2 uv = AT(p.UV, r.be1, nz); c1 = c1 + (uv.y * r.dX1 - uv.x * r.dY1) * AT(p.helem, r.be1, nz); }
    if (r.be2 >= 0)
        for (int nz = 1; nz <= p.nl - 1; nz++)
            if ((r.m2 >> nz) & 1ull) { const double2 uv = AT(p.UV, r.be2, nz); c2 = c2 - (uv.y * r.dX2 - uv.x * r.dY2) * AT(p.helem, r.be2, nz); }
    esum[e] = c1 + c2;
}
// ssh_rhs_old, hbar_old = hbar, hbar += ssh_rhs_old * dt / areasvol(ulevel)   (:1641-1663)
__global__ void LB(128) k_ale_hbar(KP p, const double* __restrict__ esum, double* __restrict__ sro, double* __restrict__ hbar,
                                   double* __restrict__ hbar_old, int ale) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= p.N) return;
    const double hb = hbar[n];
    hbar_old[n] = hb;
    if (n >= p.myN) { sro[n] = 0.0; return; }
    double s = 0.0;
    const int* slot = p.ne_slot + (size_t)(n >> 5) * p.maxdeg * TILE + (n & 31);
    for (int j = 0; j < p.maxdeg; j++) {
        const int v = slot[j * TILE];
        if (v < 0) break;
        const double cc = esum[v >> 1];
        s = (v & 1) ? s - cc : s + cc;
    }
    const int nzmin = p.n_lev[n] & 255;
    const double asv = AT(p.areasvol, tbase(n, p.nlp), nzmin);
    if (ale != 0) s = s - p.water_flux[n] * asv;
    sro[n] = s;
    hbar[n] = hb + s * p.dt / asv;
}
// dhe (:1665-1675)
__global__ void LB(128) k_ale_dhe(KP p, const double* __restrict__ hbar, const double* __restrict__ hbar_old, double* __restrict__ dhe) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= p.myE) return;
    if ((p.el_lev[e] & 255) > 1) { dhe[e] = 0.0; return; }
    const int n1 = node_of_base(p.el_nb[e], p.nlp), n2 = node_of_base(p.el_nb[p.Eall + e], p.nlp), n3 = node_of_base(p.el_nb[2 * p.Eall + e], p.nlp);
    dhe[e] = ((hbar[n1] - hbar_old[n1]) + (hbar[n2] - hbar_old[n2]) + (hbar[n3] - hbar_old[n3])) / 3.0;
}
// horizontal divergence per (node, level): the edge loop of vert_vel_ale (:1745-1790) as a gather; node-kernel mapping
__global__ void __launch_bounds__(256, 4) k_ale_wdiv(KP p, double* __restrict__ W) {
    const NodeCol c = node_col(p);
    if (!c.valid) return;
    int nz[LPT]; bool act[LPT]; double acc[LPT];
    bool any = false;
#pragma unroll
    for (int q = 0; q < LPT; q++) {
        nz[q] = blockIdx.y * (LPT * blockDim.y) + threadIdx.y + 1 + q * blockDim.y;
        act[q] = nz[q] >= c.nzmin && nz[q] <= c.nzmax - 1;
        any |= act[q];
        acc[q] = 0.0;
    }
    if (!any) return;
    const size_t sbase = (size_t)blockIdx.x * p.maxdeg * TILE + threadIdx.x;
    const int ndeg = p.tile_deg[blockIdx.x];
    for (int j = 0; j < ndeg; j++) {
        const size_t k = sbase + (size_t)j * TILE;
        const ulonglong2 m = p.ns_m[k];
        if ((m.x | m.y) == 0ull) continue;
        const int4 ni = p.ns_i[k];
        const bool second = p.ns_p[k].mk & 1ull;
        const double2 d1 = p.ns_d1[k], d2 = p.ns_d2[k];
#pragma unroll
        for (int q = 0; q < LPT; q++) {
            if (!act[q]) continue;
            if ((m.x >> nz[q]) & 1ull) {
                const double2 uv = AT(p.UV, ni.y, nz[q]);
                const double c1 = (uv.y * d1.x - uv.x * d1.y) * AT(p.helem, ni.y, nz[q]);
                acc[q] = second ? acc[q] - c1 : acc[q] + c1;
            }
            if ((m.y >> nz[q]) & 1ull) {
                const double2 uv = AT(p.UV, ni.z, nz[q]);
                const double c2 = -(uv.y * d2.x - uv.x * d2.y) * AT(p.helem, ni.z, nz[q]);
                acc[q] = second ? acc[q] - c2 : acc[q] + c2;
            }
        }
    }
#pragma unroll
    for (int q = 0; q < LPT; q++) if (act[q]) AT(W, c.b, nz[q]) = acc[q];
}
// column part of vert_vel_ale: bottom-up sum, division by the area, surface terms of zlevel / zstar with hnode_new, Courant
// number and the explicit/implicit split (:1795-2204).  One thread per own column.
__global__ void LB(128) k_ale_wcol(KP p, double* __restrict__ W, double* __restrict__ We, double* __restrict__ Wi, double* __restrict__ hnode_new,
                                   double* __restrict__ CFL, const double* __restrict__ hbar, const double* __restrict__ hbar_old,
                                   int ale, int w_split, double w_max_cfl) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= p.myN) return;
    const int b = tbase(n, p.nlp);
    const unsigned lv = p.n_lev[n];
    const int nzmin = lv & 255, nln = (lv >> 8) & 255, nlmin = (lv >> 16) & 255;
    for (int nz = nln - 1; nz >= nzmin; nz--) AT(W, b, nz) = AT(W, b, nz) + AT(W, b, nz + 1);
    for (int nz = nzmin; nz <= nln - 1; nz++) AT(W, b, nz) = AT(W, b, nz) / AT(p.area, b, nz);
    if (ale == 1) {
        if (nzmin == 1) {
            const double dh = hbar[n] - hbar_old[n];
            AT(W, b, nzmin) = AT(W, b, nzmin) - dh / p.dt;
            AT(hnode_new, b, nzmin) = AT(p.hnode, b, nzmin) + dh;
        }
        AT(W, b, nzmin) = AT(W, b, nzmin) - p.water_flux[n];
    } else if (ale == 2) {
        const int nzmax = nlmin - 1;
        if (nzmin == 1) {
            const double dd1 = AT(p.zbar3, b, nzmax);
            double dd = AT(p.zbar3, b, nzmin) - dd1;
            dd = (hbar[n] - hbar_old[n]) / dd;
            const double dddt = dd / p.dt;
            for (int nz = nzmin; nz <= nzmax - 1; nz++) {
                AT(W, b, nz) = AT(W, b, nz) - (AT(p.zbar3, b, nz) - dd1) * dddt;
                AT(hnode_new, b, nz) = AT(p.hnode, b, nz) + (AT(p.zbar3, b, nz) - AT(p.zbar3, b, nz + 1)) * dd;
            }
        }
        AT(W, b, nzmin) = AT(W, b, nzmin) - p.water_flux[n];
    }
    if (!w_split) return;                              // We == W and Wi == 0 are not read without w_split
    AT(CFL, b, 1) = 0.0;
    for (int nz = nzmin; nz <= nln - 1; nz++) {
        const double c1 = fabs(AT(W, b, nz) * p.dt / AT(hnode_new, b, nz));
        const double c2 = fabs(AT(W, b, nz + 1) * p.dt / AT(hnode_new, b, nz));
        AT(CFL, b, nz) = AT(CFL, b, nz) + c1;
        AT(CFL, b, nz + 1) = c2;
    }
    for (int nz = nzmin; nz <= nln; nz++) {
        double c1 = 1.0, c2 = 0.0;
        const double cfl = AT(CFL, b, nz);
        if (cfl > w_max_cfl) {
            const double dd = omax(cfl - w_max_cfl, 0.0) / omax(w_max_cfl, 1.e-12);
            c1 = 1.0 / (1.0 + dd); c2 = dd / (1.0 + dd);
        }
        AT(We, b, nz) = c1 * AT(W, b, nz);
        AT(Wi, b, nz) = c2 * AT(W, b, nz);
    }
}
// update_thickness_ale, node part (:905-970): hnode = hnode_new, zbar_3d_n and Z_3d_n rebuilt from below
__global__ void LB(128) k_ale_thick(KP p, double* __restrict__ hnode, double* __restrict__ zbar3, double* __restrict__ Z3, int ale) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= p.N) return;
    const int b = tbase(n, p.nlp);
    const unsigned lv = p.n_lev[n];
    const int uln = lv & 255, nlmin = (lv >> 16) & 255, ulmax = (lv >> 24) & 255;
    if (ale == 1) {
        const int nz = ulmax;
        if (nz > 1) return;
        const double hn = AT(p.hnode_new, b, nz);
        AT(hnode, b, nz) = hn;
        AT(zbar3, b, nz) = AT(zbar3, b, nz + 1) + hn;
        AT(Z3, b, nz) = AT(zbar3, b, nz + 1) + hn / 2.0;
    } else if (ale == 2) {
        if (uln > 1) return;
        for (int nz = nlmin - 2; nz >= uln; nz--) {
            const double hn = AT(p.hnode_new, b, nz);
            AT(hnode, b, nz) = hn;
            AT(zbar3, b, nz) = AT(zbar3, b, nz + 1) + hn;
            AT(Z3, b, nz) = AT(zbar3, b, nz + 1) + hn / 2.0;
        }
    }
}
// update_thickness_ale, element part (:972-990)
__global__ void LB(128) k_ale_helem(KP p, double* __restrict__ helem, int ale) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= p.myE) return;
    const unsigned lv = p.el_lev[e];
    const int nu = lv & 255, nlm = (lv >> 8) & 255;
    if (nu > 1) return;
    const int b1 = p.el_nb[e], b2 = p.el_nb[p.Eall + e], b3 = p.el_nb[2 * p.Eall + e], be = tbase(e, p.nlp);
    if (ale == 1) AT(helem, be, nu) = (AT(p.hnode_new, b1, nu) + AT(p.hnode_new, b2, nu) + AT(p.hnode_new, b3, nu)) / 3.0;
    else if (ale == 2)
        for (int nz = nu; nz <= nlm - 1; nz++) AT(helem, be, nz) = (AT(p.hnode, b1, nz) + AT(p.hnode, b2, nz) + AT(p.hnode, b3, nz)) / 3.0;
}

static bool ale_ready(TracerCtx& c, const char* who) {
    if (!fb_ok()) return false;
    if (!c.committed || c.which_ale < 0) { fb_set_error(28, std::string(who) + ": fb200_init_commit / fb200_ale_init missing"); return false; }
    if (comm_nranks() > 1 && !c.halo_nod) { fb_set_error(26, std::string(who) + ": more than one rank but fb200_init_halo was not called"); return false; }
    return true;
}

extern "C" {

// which_ale: 0 linfs, 1 zlevel, 2 zstar (oce_ale.F90:82-795 sets the thicknesses up on the host; push them once with
// fb200_push_step_state).  hbar(N) = the current surface elevation.
void fb200_ale_init(int* which_ale, double* w_max_cfl, double* hbar) {
    TracerCtx& c = tctx();
    if (!c.committed) { fb_set_error(28, "fb200_ale_init before fb200_init_commit"); return; }
    if (*which_ale < 0 || *which_ale > 2) { fb_set_error(28, "fb200_ale_init: which_ale must be 0 (linfs), 1 (zlevel) or 2 (zstar)"); return; }
    fb_copy_sync();
    c.stream = fb_stream();
    c.which_ale = *which_ale; c.w_max_cfl = *w_max_cfl;
    if (!c.hbar.p) {
        c.hbar.alloc(c.N); c.hbar_old.alloc(c.N); c.ssh_rhs_old.alloc(c.N); c.dhe.alloc(c.myE); c.esum.alloc(c.myEd);
        if (c.opt.w_split) c.CFL.alloc((size_t)c.Ntiles * c.nlp * TILE);
    }
    CUDA_CHECK(cudaMemcpyAsync(c.hbar.p, hbar, (size_t)c.N * sizeof(double), cudaMemcpyHostToDevice, c.stream));
    CUDA_CHECK(cudaMemcpyAsync(c.hbar_old.p, hbar, (size_t)c.N * sizeof(double), cudaMemcpyHostToDevice, c.stream));
    CUDA_CHECK(cudaStreamSynchronize(c.stream));
}

// compute_hbar_ale + vert_vel_ale on the device from the velocity the host has just updated (update_vel) and the
// freshwater flux: H2D UV(2,nl-1,elems) and water_flux(N); resident afterwards: Wvel, Wvel_e, Wvel_i, hnode_new; back to the
// host (it needs them for the next SSH right-hand side and stiffness update): hbar(N), ssh_rhs_old(N), dhe(myDim_elem2D).
void fb200_ale_step_begin(double* UV, double* water_flux, double* hbar_out, double* ssh_rhs_old_out, double* dhe_out) {
    TracerCtx& c = tctx();
    if (!ale_ready(c, "fb200_ale_step_begin")) return;
    fb_copy_sync();
    cudaStream_t st = c.stream = fb_stream();
    const int L = c.nl - 1;
    if (UV) {                                          // NULL: the resident velocity (fb200_update_vel has just produced it)
        ctx_push_node_field(c, UV, L, c.E, (double*)c.UV.p, 2, 0, 2, 0);
        ctx_push_node_field(c, UV, L, c.E, (double*)c.UV.p, 2, 1, 2, 1);
    }
    CUDA_CHECK(cudaMemcpyAsync(c.water_flux.p, water_flux, (size_t)c.N * sizeof(double), cudaMemcpyHostToDevice, st));
    KP p = make_kp(c, 0);
    const int ale = c.which_ale;
#define KL(name, ...) do { ProfScope ps_(name, st); __VA_ARGS__; LAUNCH_CHECK(); } while (0)
    KL("k_ale_edge_sum", k_ale_edge_sum<<<(c.myEd + 127) / 128, 128, 0, st>>>(p, c.esum.p));
    KL("k_ale_hbar", k_ale_hbar<<<(c.N + 127) / 128, 128, 0, st>>>(p, c.esum.p, c.ssh_rhs_old.p, c.hbar.p, c.hbar_old.p, ale));
    const bool multi = comm_nranks() > 1;
    if (multi) {
        if (ale != 0) halo_exchange_flat(*c.halo_nod, c.ssh_rhs_old.p, st);           // exchange_nod(ssh_rhs_old), :1654
        halo_exchange_flat(*c.halo_nod, c.hbar.p, st);                                // exchange_nod(hbar), :1663
    }
    KL("k_ale_dhe", k_ale_dhe<<<(c.myE + 127) / 128, 128, 0, st>>>(p, c.hbar.p, c.hbar_old.p, c.dhe.p));
    CUDA_CHECK(cudaMemsetAsync(c.W.p, 0, c.W.n * sizeof(double), st));
    const int nchunk = (c.nl - 1 + LPT * 8 - 1) / (LPT * 8);
    KL("k_ale_wdiv", k_ale_wdiv<<<dim3(c.myNtiles, nchunk), dim3(TILE, 8), 0, st>>>(p, c.W.p));
    KL("k_ale_wcol", k_ale_wcol<<<(c.myN + 127) / 128, 128, 0, st>>>(p, c.W.p, c.We.p, c.Wi.p, c.hnode_new.p, c.CFL.p, c.hbar.p, c.hbar_old.p,
                                                                   ale, c.opt.w_split, c.w_max_cfl));
    if (multi) {
        halo_exchange_tiled(*c.halo_nod, c.W.p, c.nlp, 1, st);                        // exchange_nod(Wvel), :2135
        if (ale != 0) halo_exchange_tiled(*c.halo_nod, c.hnode_new.p, c.nlp, 1, st);  // exchange_nod(hnode_new), :2136
        if (c.opt.w_split) { halo_exchange_tiled(*c.halo_nod, c.We.p, c.nlp, 1, st); halo_exchange_tiled(*c.halo_nod, c.Wi.p, c.nlp, 1, st); }
    }
    CUDA_CHECK(cudaMemcpyAsync(hbar_out, c.hbar.p, (size_t)c.N * sizeof(double), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaMemcpyAsync(ssh_rhs_old_out, c.ssh_rhs_old.p, (size_t)c.N * sizeof(double), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaMemcpyAsync(dhe_out, c.dhe.p, (size_t)c.myE * sizeof(double), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
}

// update_thickness_ale on the device, after the tracer step: hnode, zbar_3d_n, Z_3d_n, helem for the next step
void fb200_ale_step_end(void) {
    TracerCtx& c = tctx();
    if (!ale_ready(c, "fb200_ale_step_end")) return;
    cudaStream_t st = c.stream = fb_stream();
    KP p = make_kp(c, 0);
    const int ale = c.which_ale;
    if (ale == 0) return;                              // linfs: the thicknesses never change
    KL("k_ale_thick", k_ale_thick<<<(c.N + 127) / 128, 128, 0, st>>>(p, c.hnode.p, c.zbar3.p, c.Z3.p, ale));
    KL("k_ale_helem", k_ale_helem<<<(c.myE + 127) / 128, 128, 0, st>>>(p, c.helem.p, ale));
#undef KL
    CUDA_CHECK(cudaStreamSynchronize(st));
}

}  // extern "C"

// ===================================================================================================
// SSH assembly on the device: compute_ssh_rhs_ale (oce_ale.F90:1478-1572), update_stiff_mat_ale
// (:1371-1470) and update_vel (oce_dyn.F90:101-127), restated after the host-side mirror fesom2_b200/host/setup.cpp.  They
// work on the state the device-side ALE producers keep resident (UV, helem, water_flux, ssh_rhs_old, dhe), so
// fb200_ale_init must have been called.  What still crosses PCIe is the momentum step's UV_rhs (the momentum step itself
// is outside the path); the stiffness values, the right-hand side and — after update_vel — UV are returned so that the
// host's psolve call and momentum step see them.
// per edge: c1 + c2 of compute_ssh_rhs_ale (:1518-1547)
__global__ void LB(128) k_sshrhs_edge_sum(KP p, const double2* __restrict__ UVr, double alpha, double* __restrict__ esum) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= p.myEd) return;
    const EdgeRec r = p.edge[e];
    double c1 = 0.0, c2 = 0.0;
    for (int nz = 1; nz <= p.nl - 1; nz++)
        if ((r.m1 >> nz) & 1ull) {
            const double2 uv = AT(p.UV, r.be1, nz), ur = AT(UVr, r.be1, nz);
            c1 = c1 + alpha * ((uv.y + ur.y) * r.dX1 - (uv.x + ur.x) * r.dY1) * AT(p.helem, r.be1, nz);
        }
    if (r.be2 >= 0)
        for (int nz = 1; nz <= p.nl - 1; nz++)
            if ((r.m2 >> nz) & 1ull) {
                const double2 uv = AT(p.UV, r.be2, nz), ur = AT(UVr, r.be2, nz);
                c2 = c2 - alpha * ((uv.y + ur.y) * r.dX2 - (uv.x + ur.x) * r.dY2) * AT(p.helem, r.be2, nz);
            }
    esum[e] = c1 + c2;
}
// per own node: signed sum over the incident edges in ascending edge id, water flux and the (1 - alpha) part (:1549-1568)
__global__ void LB(128) k_sshrhs_node(KP p, const double* __restrict__ esum, const double* __restrict__ sro, double* __restrict__ rhs,
                                      double alpha, int ale) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= p.N) return;
    if (n >= p.myN) { rhs[n] = 0.0; return; }
    double s = 0.0;
    const int* slot = p.ne_slot + (size_t)(n >> 5) * p.maxdeg * TILE + (n & 31);
    for (int j = 0; j < p.maxdeg; j++) {
        const int v = slot[j * TILE];
        if (v < 0) break;
        const double cc = esum[v >> 1];
        s = (v & 1) ? s - cc : s + cc;
    }
    const int nzmin = p.n_lev[n] & 255;
    if (ale != 0) {
        const double asv = AT(p.areasvol, tbase(n, p.nlp), nzmin);
        if (nzmin > 1) s = s - alpha * p.water_flux[n] * asv;
        else s = s - alpha * p.water_flux[n] * asv + (1.0 - alpha) * sro[n];
    } else if (nzmin <= 1) s = s + (1.0 - alpha) * sro[n];
    rhs[n] = s;
}
// update_stiff_mat_ale (:1398-1468) as a row gather: one thread per own row walks its edges in ascending edge id and, per
// edge, the two elements and their three corners — the order in which the reference's edge loop reaches the entries of
// that row — so every entry receives its contributions in the reference's order without atomics.
__global__ void LB(128) k_stiff_update(KP p, const int* __restrict__ rowptr, const int* __restrict__ colind, const double* __restrict__ dhe,
                                       double factor, double* __restrict__ values) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= p.myN) return;
    const int q0 = rowptr[n] - rowptr[0], q1 = rowptr[n + 1] - rowptr[0];
    const int* slot = p.ne_slot + (size_t)(n >> 5) * p.maxdeg * TILE + (n & 31);
    for (int j = 0; j < p.maxdeg; j++) {
        const int v = slot[j * TILE];
        if (v < 0) break;
        const EdgeRec r = p.edge[v >> 1];
        const bool second = v & 1;
        for (int i = 0; i < 2; i++) {
            const int be = i == 0 ? r.be1 : r.be2;
            if (be < 0) continue;
            const int elem = node_of_base(be, p.nlp);
            const double dX = i == 0 ? r.dX1 : r.dX2, dY = i == 0 ? r.dY1 : r.dY2;
            const double dh = dhe[elem];
            for (int k = 0; k < 3; k++) {
                double fy = -dh * (p.el_g[k * p.myE + elem] * dY - p.el_g[(k + 3) * p.myE + elem] * dX);
                if (i == 1) fy = -fy;
                if (second) fy = -fy;
                const int col = node_of_base(p.el_nb[k * p.Eall + elem], p.nlp) + 1;
                for (int q = q0; q < q1; q++)
                    if (colind[q] == col) { values[q] = values[q] + fy * factor; break; }
            }
        }
    }
}
// update_vel (oce_dyn.F90:101-124): UV += UV_rhs - g theta dt grad(d_eta) on own elements
__global__ void LB(128) k_update_vel(KP p, const double2* __restrict__ UVr, const double* __restrict__ d_eta, double g, double theta,
                                     double2* __restrict__ UV) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= p.myE) return;
    const int n1 = node_of_base(p.el_nb[e], p.nlp), n2 = node_of_base(p.el_nb[p.Eall + e], p.nlp), n3 = node_of_base(p.el_nb[2 * p.Eall + e], p.nlp);
    const double e1 = -(g * theta * p.dt * d_eta[n1]), e2 = -(g * theta * p.dt * d_eta[n2]), e3 = -(g * theta * p.dt * d_eta[n3]);
    const double Fx = p.el_g[e] * e1 + p.el_g[p.myE + e] * e2 + p.el_g[2 * p.myE + e] * e3;
    const double Fy = p.el_g[3 * p.myE + e] * e1 + p.el_g[4 * p.myE + e] * e2 + p.el_g[5 * p.myE + e] * e3;
    const unsigned lv = p.el_lev[e];
    const int nu = lv & 255, nlm = (lv >> 8) & 255, be = tbase(e, p.nlp);
    for (int nz = nu; nz <= nlm; nz++) {
        const double2 uv = AT(UV, be, nz), ur = AT(UVr, be, nz);
        AT(UV, be, nz) = make_double2(uv.x + ur.x + Fx, uv.y + ur.y + Fy);
    }
}
// UV in the reference layout (2, nl-1, elems) for the trip back to the host
__global__ void LB(256) k_uv_to_ref(const double2* __restrict__ UV, double2* __restrict__ out, int E, int L, int nlp) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)E * L) return;
    const int e = (int)(i / L), nz = (int)(i - (size_t)e * L) + 1;
    out[i] = AT(UV, tbase(e, nlp), nz);
}

static void push_uv_rhs(TracerCtx& c, double* UV_rhs) {
    const size_t ne = (size_t)c.Etiles * c.nlp * TILE;
    if (c.UVr.n < ne) c.UVr.alloc(ne);
    const int L = c.nl - 1;
    ctx_push_node_field(c, UV_rhs, L, c.E, (double*)c.UVr.p, 2, 0, 2, 0);
    ctx_push_node_field(c, UV_rhs, L, c.E, (double*)c.UVr.p, 2, 1, 2, 1);
}

extern "C" {

// once, after fb200_ale_init: the CSR structure of ssh_stiff as init_stiff_mat_ale leaves it (rowptr(myDim_nod2D+1) 1-based,
// colind_loc(nnz) LOCAL 1-based node numbers, values(nnz)), the host's current dhe(myDim_elem2D) and ssh_rhs_old(N) (what its
// last compute_hbar_ale left; from then on fb200_ale_step_begin produces them), g and the namelist's alpha, theta
// (oce_ale.F90:1088-1354)
void fb200_ssh_assembly_init(int* rowptr, int* colind_loc, double* values, double* dhe, double* ssh_rhs_old, double* g, double* alpha,
                             double* theta) {
    TracerCtx& c = tctx();
    if (!ale_ready(c, "fb200_ssh_assembly_init")) return;
    fb_copy_sync();
    cudaStream_t st = c.stream = fb_stream();
    const int nnz = rowptr[c.myN] - rowptr[0];
    c.sa_rowptr.alloc(c.myN + 1); c.sa_colind.alloc(nnz); c.sa_values.alloc(nnz); c.sa_rhs.alloc(c.N);
    CUDA_CHECK(cudaMemcpyAsync(c.sa_rowptr.p, rowptr, (size_t)(c.myN + 1) * sizeof(int), cudaMemcpyHostToDevice, st));
    CUDA_CHECK(cudaMemcpyAsync(c.sa_colind.p, colind_loc, (size_t)nnz * sizeof(int), cudaMemcpyHostToDevice, st));
    CUDA_CHECK(cudaMemcpyAsync(c.sa_values.p, values, (size_t)nnz * sizeof(double), cudaMemcpyHostToDevice, st));
    CUDA_CHECK(cudaMemcpyAsync(c.dhe.p, dhe, (size_t)c.myE * sizeof(double), cudaMemcpyHostToDevice, st));
    CUDA_CHECK(cudaMemcpyAsync(c.ssh_rhs_old.p, ssh_rhs_old, (size_t)c.N * sizeof(double), cudaMemcpyHostToDevice, st));
    c.sa_g = *g; c.sa_alpha = *alpha; c.sa_theta = *theta;
    CUDA_CHECK(cudaStreamSynchronize(st));
}

// update_stiff_mat_ale from the resident dhe of the last fb200_ale_step_begin; values_out(nnz) receives the updated matrix
// (what the host hands to psolve); not called under linfs, as in the reference (oce_ale.F90:2686)
void fb200_update_stiff_mat(double* values_out) {
    TracerCtx& c = tctx();
    if (!ale_ready(c, "fb200_update_stiff_mat")) return;
    if (!c.sa_values.p) { fb_set_error(28, "fb200_update_stiff_mat before fb200_ssh_assembly_init"); return; }
    fb_copy_sync();
    cudaStream_t st = c.stream = fb_stream();
    KP p = make_kp(c, 0);
    const double factor = c.sa_g * c.opt.dt * c.sa_alpha * c.sa_theta;
    { ProfScope ps_("k_stiff_update", st); k_stiff_update<<<(c.myN + 127) / 128, 128, 0, st>>>(p, c.sa_rowptr.p, c.sa_colind.p, c.dhe.p, factor, c.sa_values.p); LAUNCH_CHECK(); }
    CUDA_CHECK(cudaMemcpyAsync(values_out, c.sa_values.p, c.sa_values.n * sizeof(double), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
}

// compute_ssh_rhs_ale: H2D UV_rhs(2, nl-1, elems) of the momentum step; ssh_rhs_out(N) (own entries, halo exchanged)
void fb200_compute_ssh_rhs(double* UV_rhs, double* ssh_rhs_out) {
    TracerCtx& c = tctx();
    if (!ale_ready(c, "fb200_compute_ssh_rhs")) return;
    if (!c.sa_rhs.p) { fb_set_error(28, "fb200_compute_ssh_rhs before fb200_ssh_assembly_init"); return; }
    fb_copy_sync();
    cudaStream_t st = c.stream = fb_stream();
    push_uv_rhs(c, UV_rhs);
    KP p = make_kp(c, 0);
    { ProfScope ps_("k_sshrhs_edge_sum", st); k_sshrhs_edge_sum<<<(c.myEd + 127) / 128, 128, 0, st>>>(p, c.UVr.p, c.sa_alpha, c.esum.p); LAUNCH_CHECK(); }
    { ProfScope ps_("k_sshrhs_node", st); k_sshrhs_node<<<(c.N + 127) / 128, 128, 0, st>>>(p, c.esum.p, c.ssh_rhs_old.p, c.sa_rhs.p, c.sa_alpha, c.which_ale); LAUNCH_CHECK(); }
    if (comm_nranks() > 1) halo_exchange_flat(*c.halo_nod, c.sa_rhs.p, st);          // exchange_nod(ssh_rhs), :1570
    CUDA_CHECK(cudaMemcpyAsync(ssh_rhs_out, c.sa_rhs.p, (size_t)c.N * sizeof(double), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
}

// update_vel with the UV_rhs of the last fb200_compute_ssh_rhs and the SSH solution d_eta(N): the resident UV becomes the
// velocity of the new step (pass UV = NULL to the following fb200_ale_step_begin) and UV_out(2, nl-1, elems) returns it to
// the host's momentum code.  eta_n += d_eta stays on the host (one line of oce_dyn.F90:125).
void fb200_update_vel(double* d_eta, double* UV_out) {
    TracerCtx& c = tctx();
    if (!ale_ready(c, "fb200_update_vel")) return;
    if (!c.UVr.p) { fb_set_error(28, "fb200_update_vel before fb200_compute_ssh_rhs"); return; }
    fb_copy_sync();
    cudaStream_t st = c.stream = fb_stream();
    if (c.sa_eta.n < (size_t)c.N) c.sa_eta.alloc(c.N);
    CUDA_CHECK(cudaMemcpyAsync(c.sa_eta.p, d_eta, (size_t)c.N * sizeof(double), cudaMemcpyHostToDevice, st));
    KP p = make_kp(c, 0);
    { ProfScope ps_("k_update_vel", st); k_update_vel<<<(c.myE + 127) / 128, 128, 0, st>>>(p, c.UVr.p, c.sa_eta.p, c.sa_g, c.sa_theta, c.UV.p); LAUNCH_CHECK(); }
    if (comm_nranks() > 1) halo_exchange_tiled(*c.halo_elem, (double*)c.UV.p, c.nlp, 2, st);      // exchange_elem(UV), oce_dyn.F90:126
    const int L = c.nl - 1;
    const size_t cnt = (size_t)c.Eall * L;              // UV(2, nl-1, myDim+eDim+eXDim): exchange_elem fills every halo element
    if (c.stage.n < 2 * cnt) c.stage.alloc(2 * cnt);
    { ProfScope ps_("k_uv_to_ref", st); k_uv_to_ref<<<(unsigned)((cnt + 255) / 256), 256, 0, st>>>(c.UV.p, (double2*)c.stage.p, c.Eall, L, c.nlp); LAUNCH_CHECK(); }
    CUDA_CHECK(cudaMemcpyAsync(UV_out, c.stage.p, 2 * cnt * sizeof(double), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
}

}  // extern "C"
