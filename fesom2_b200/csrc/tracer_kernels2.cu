// Tracer transport kernels, second generation: the same arithmetic as tracer_kernels.cu (every expression keeps the
// reference's operation order, --fmad=false), restructured around two observations of the round-1 profiles
// (profiles/r01_ncu_tracer_1m.txt): the node and edge kernels were bound by instruction issue and gather latency, not
// by DRAM — 600-1100 warp instructions per (warp, level), most of them index arithmetic, mask tests and per-(slot, level)
// bookkeeping repeated for each of a thread's three scattered levels — and every tracer re-read the tracer-independent
// data (slot records, masks, volume fluxes from UV * helem, areasvol, thicknesses, W) that the tracers share.
//
//  * COLUMN THREADS.  One thread owns one (node | edge | element) column and walks down its levels; one warp = one tile of
//    32 consecutive entities, so every field access of a warp at a level is one coalesced 256-byte line of the
//    level-blocked layout.  Tile bases, slot records, geometry and level ranges are set up once per column instead of
//    once per three levels; vertical stencils (3-point limiter bounds, limited vertical fluxes) live in registers that
//    slide down the column — no shared-memory exchange, no block barrier.
//  * SLOT RECORDS IN SHARED MEMORY.  The per-(node, slot) records of a tile (other node, edge, level mask | sign) are
//    staged once per warp in shared memory and read back conflict-free at every level.
//  * ALL TRACERS OF A GROUP IN ONE PASS (NT = 1 or 2).  A thread carries NT accumulators; the record, the mask test, the
//    volume flux, areasvol, the thicknesses and the vertical velocities are loaded once and serve every tracer.
//
// Sums still run over the incident edges in ascending reference edge id (slot order), one term at a time, exactly like
// the reference's edge loops add into a node, so results stay bit-identical to the oracle (tests/test_gpu_tracer.py,
// tests/test_emulated.py).
#include "common.cuh"
#include "comm.cuh"
#include "tracer_dev.cuh"
#include "tracer_kernels2.cuh"
#include <cstdlib>
#include <string>

// compiler barrier between the load block and the arithmetic of a level: keeps the loads together at the top (without it
// the compiler may sink each gather down to its first use to save registers, which serialises the round trips again)
#define LOADS_DONE() asm volatile("" ::: "memory")
// correctly rounded reciprocals of the constants the reconstruction formulas divide by (see div_by in tracer_dev.cuh)
#define R3 (1.0 / 3.0)
#define R6 (1.0 / 6.0)
// minimum resident blocks per SM of the node kernels (register budget = 65536 / (128 * MB)); -D overrides for A/B builds
#ifndef LO_MB
#define LO_MB 3
#endif
#ifndef FCTB_MB
#define FCTB_MB 5
#endif
#ifndef UPD_MB
#define UPD_MB 4
#endif
#ifndef EDGE_MB
#define EDGE_MB 6
#endif
#define WPB 4                                 // warps (= tiles) per block of the column kernels
__device__ __forceinline__ int warp_min_i(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ int warp_max_i(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// one own-node column of a column kernel: warp = tile
struct Col { int tile, lane, n, b, nzmin, nzmax, nlmin; bool valid; };
__device__ __forceinline__ bool col_setup(const KP& p, Col& c) {
    c.lane = threadIdx.x & 31;
    c.tile = blockIdx.x * WPB + (threadIdx.x >> 5);
    if (c.tile * TILE >= p.myN) return false;                      // warp-uniform
    c.n = c.tile * TILE + c.lane;
    c.valid = c.n < p.myN;
    c.b = c.tile * (p.nlp * TILE) + c.lane;
    const unsigned lv = c.valid ? p.n_lev[c.n] : 0u;
    c.nzmin = lv & 255; c.nzmax = (lv >> 8) & 255; c.nlmin = (lv >> 16) & 255;
    return true;
}
#define SB_ 8                                 // = SB below
__host__ __device__ inline int slot_cap(int maxdeg) { return ((maxdeg + SB_ - 1) / SB_) * SB_; }
// stage the tile's slot records in this warp's part of shared memory; returns the warp's record array [cap][32]
__device__ __forceinline__ const SlotRec* stage_slots(const KP& p, const Col& c, SlotRec* sm, int& ndeg) {
    ndeg = p.tile_deg[c.tile];
    const int cap = slot_cap(p.maxdeg);                            // the warp's array holds whole batches: slots beyond the tile's degree get zero masks
    SlotRec* rec = sm + (size_t)(threadIdx.x >> 5) * cap * TILE;
    const SlotRec* src = p.ns_p + (size_t)c.tile * p.maxdeg * TILE + c.lane;
    SlotRec none; none.bo = c.b; none.bed = p.zero_edge; none.mk = 0ull;
    for (int j = 0; j < cap; j++) rec[j * TILE + c.lane] = j < ndeg ? src[j * TILE] : none;
    __syncwarp();
    return rec + c.lane;
}

// ---------------------------------------------------------------------------------------------------
// Adams-Bashforth extrapolation of every tracer of the group (contiguous fields), oce_tracer_mod.F90:64
__global__ void LB(256) k_ab2_g(double* __restrict__ told, const double* __restrict__ t, size_t count, double ca, double cb) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < count; i += stride) told[i] = ca * told[i] + cb * t[i];
}

// tracer_gradient_elements (oce_tracer_mod.F90:35-44) of NT fields in one pass: the corner bases and gradient_sca are read
// once.  which = 0: gradients of ttfAB (Told), 1: of ttf (T)
template <int NT>
__global__ void LB(128) k_grad_elem2(KP p, TrSet s, int which) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= p.myE) return;
    const int b1 = p.el_nb[e], b2 = p.el_nb[p.Eall + e], b3 = p.el_nb[2 * p.Eall + e];
    const double g1 = p.el_g[e], g2 = p.el_g[p.myE + e], g3 = p.el_g[2 * p.myE + e];
    const double g4 = p.el_g[3 * p.myE + e], g5 = p.el_g[4 * p.myE + e], g6 = p.el_g[5 * p.myE + e];
    const unsigned lv = p.el_lev[e];
    const int nu = lv & 255, nlm = (lv >> 8) & 255;
    const int be = tbase(e, p.nlp);
    const double* __restrict__ F[NT];
#pragma unroll
    for (int t = 0; t < NT; t++) F[t] = which ? s.T[t] : s.Told[t];
#pragma unroll 4
    for (int nz = nu; nz <= nlm; nz++) {
#pragma unroll
        for (int t = 0; t < NT; t++) {
            const double t1 = AT(F[t], b1, nz), t2 = AT(F[t], b2, nz), t3 = AT(F[t], b3, nz);
            AT(s.txy[t], be, nz) = make_double2(g1 * t1 + g2 * t2 + g3 * t3, g4 * t1 + g5 * t2 + g6 * t3);
        }
    }
}

// tracer_gradient_z (oce_tracer_mod.F90:139-152), all local nodes, NT fields
template <int NT>
__global__ void LB(128) k_grad_z2(KP p, TrSet s) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= p.N) return;
    const unsigned lv = p.n_lev[n];
    const int nzmin = lv & 255, nzmax = (lv >> 8) & 255;
    const int b = tbase(n, p.nlp);
    for (int nz = nzmin + 1; nz <= nzmax - 1; nz++) {
        const double dz = 0.5 * (AT(p.hnode_new, b, nz - 1) + AT(p.hnode_new, b, nz));
#pragma unroll
        for (int t = 0; t < NT; t++) AT(s.trz[t], b, nz) = (AT(s.T[t], b, nz - 1) - AT(s.T[t], b, nz)) / dz;
    }
#pragma unroll
    for (int t = 0; t < NT; t++) { AT(s.trz[t], b, nzmin) = 0.0; AT(s.trz[t], b, nzmax) = 0.0; }
}

// ---------------------------------------------------------------------------------------------------
// LOAD BATCHING.  The first column kernels walked the slots (and the two segments of an edge) through `if (active) { load;
// use; }` regions: correct, and slow — every region became its own basic block, so a warp issued one gather, waited out its
// latency, issued the next (SASS: one BSSY/BSYNC pair per slot), eight dependent round trips per level.  Below, every
// address is made valid first (inactive slots and absent elements read a harmless resident entry), ALL loads of a level
// are issued unconditionally in one straight-line block, and only then the values are combined with selects, in slot
// order.  One round trip per level instead of one per slot.

// High-order horizontal flux minus low-order flux (adf_h) and the low-order flux itself (eflx), per own edge and level:
// adv_tra_hor_mfct / _muscl / _upw1 with init_zero=.false. on ttfAB against adv_tra_hor_upw1(ttf)
// (oce_adv_tra_hor.F90:57-733, oce_adv_tra_driver.F90:83,142-149), gradients of fill_up_dn_grad gathered from txy.
// Thread per edge column; the volume flux through the edge's two segments is formed once per level for all tracers.
// Levels whose gradients are node means go to the list kernel k_ho_edge_fb (tracer_kernels.cu), as before.
template <int NT, bool FCT>
__global__ void __launch_bounds__(128, EDGE_MB) k_edge_adv2(KP p, TrSet s) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = e < p.myEd;
    EdgeRec r; EdgeHo h;
    unsigned long long many = 0ull;
    if (valid) { r = p.edge[e]; h = p.edgeho[e]; many = r.m1 | r.m2; }
    const int zlo = warp_min_i(many ? __ffsll((long long)many) - 1 : 255), zhi = warp_max_i(many ? 63 - __clzll((long long)many) : 0);
    if (!valid || !many) return;
    const int be = tbase(e, p.nlp);
    const int sh_lo = h.shared & 255, sh_hi = (h.shared >> 8) & 255;
    const bool good = (h.shared >> 16) & 1;
    const bool grads = p.hor != 2;
    const int nbl1 = p.hor == 0 ? p.n_nbl[h.n1] : 0, nbl2 = p.hor == 0 ? p.n_nbl[h.n2] : 0;
    const int be2s = r.be2 >= 0 ? r.be2 : r.be1;                     // boundary edge: a resident stand-in, never used
    const int bups = h.bup >= 0 ? h.bup : r.be1, bdns = h.bdn >= 0 ? h.bdn : r.be1;
    for (int nz = zlo; nz <= zhi; nz++) {
        const int segs = mask_segs(r.m1, r.m2, nz);
        if (!segs) continue;
        if (grads && !(good && nz >= sh_lo && nz <= sh_hi - 1)) continue;                        // fallback level: k_ho_edge_fb
        // ---- all loads of this level
        const double2 uv1 = AT(p.UV, r.be1, nz), uv2 = AT(p.UV, be2s, nz);
        const double h1 = AT(p.helem, r.be1, nz), h2 = AT(p.helem, be2s, nz);
        double t1[NT], t2[NT], a1[NT], a2[NT];
        double2 gu[NT], gd[NT];
#pragma unroll
        for (int t = 0; t < NT; t++) {
            t1[t] = t2[t] = 0.0;
            if (FCT) { t1[t] = AT(s.T[t], r.bn1, nz); t2[t] = AT(s.T[t], r.bn2, nz); }
            a1[t] = AT(s.Told[t], r.bn1, nz); a2[t] = AT(s.Told[t], r.bn2, nz);
            gu[t] = gd[t] = make_double2(0.0, 0.0);
            if (grads) { gu[t] = AT(s.txy[t], bups, nz); gd[t] = AT(s.txy[t], bdns, nz); }
        }
        LOADS_DONE();
        // ---- arithmetic (oce_adv_tra_hor.F90:139,157,176: volume flux through the two median-dual segments)
        const double c1 = (-uv1.y * r.dX1 + uv1.x * r.dY1) * h1, c2 = (uv2.y * r.dX2 - uv2.x * r.dY2) * h2;
        const double vflux = segs == 3 ? c1 + c2 : ((segs & 1) ? c1 : c2);
#pragma unroll
        for (int t = 0; t < NT; t++) {
            const double flo = FCT ? upw1_flux(t1[t], t2[t], vflux) : 0.0;
            double fho;
            if (p.hor == 2) fho = upw1_flux(a1[t], a2[t], vflux);
            else {
                const double g1 = gu[t].x, g2 = gd[t].x, g3 = gu[t].y, g4 = gd[t].y;             // edge_up_dn_grad(1:4)
                // x / 6.0 as the exact-division sequence with the constant's reciprocal (div_by, tracer_dev.cuh)
                const double e2 = div_by(2.0 * (a2[t] - a1[t]) + h.dxa * g2 + h.dyr * g4, 6.0, R6), e1 = div_by(2.0 * (a2[t] - a1[t]) + h.dxa * g1 + h.dyr * g3, 6.0, R6);
                double Tmean2 = a2[t] - e2;
                double Tmean1 = a1[t] + e1;
                if (p.hor == 0) {                                                                // MUSCL: c_lo, oce_adv_tra_hor.F90:352-353
                    const double c_lo1 = (nbl1 - nz) >= 0 ? 1.0 : 0.0, c_lo2 = (nbl2 - nz) >= 0 ? 1.0 : 0.0;
                    Tmean2 = a2[t] - e2 * c_lo2;
                    Tmean1 = a1[t] + e1 * c_lo1;
                }
                const double cHO = (vflux + fabs(vflux)) * Tmean1 + (vflux - fabs(vflux)) * Tmean2;
                fho = -0.5 * (1.0 - p.ph) * cHO - vflux * p.ph * (0.5 * (Tmean1 + Tmean2));
            }
            AT(s.adf_h[t], be, nz) = fho - flo;
            if (FCT) AT(s.eflx[t], be, nz) = flo;
        }
    }
}

// diff_part_hor_redi (oce_ale_tracer.F90:934-1077): diffusive flux through the two segments of every own edge and level
// -> eflx (gathered by k_update2).  Thread per edge column; Kh and the layer thickness are formed once for all tracers.
template <int NT>
__global__ void __launch_bounds__(128, EDGE_MB) k_edge_diff2(KP p, TrSet s) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = e < p.myEd;
    EdgeRec r;
    unsigned long long many = 0ull;
    if (valid) { r = p.edge[e]; many = r.m1 | r.m2; }
    const int zlo = warp_min_i(many ? __ffsll((long long)many) - 1 : 255), zhi = warp_max_i(many ? 63 - __clzll((long long)many) : 0);
    if (!valid || !many) return;
    const int be = tbase(e, p.nlp);
    const double isredi = p.Redi ? 1.0 : 0.0;
    const int be2s = r.be2 >= 0 ? r.be2 : r.be1;
    for (int z = zlo; z <= zhi; z++) {
        const int segs = mask_segs(r.m1, r.m2, z);
        if (!segs) continue;
        // ---- all loads of this level
        const double ki1 = AT(p.Ki, r.bn1, z), ki2 = AT(p.Ki, r.bn2, z);
        const double he1 = AT(p.helem, r.be1, z), he2 = AT(p.helem, be2s, z);
        double s11 = 0.0, s12 = 0.0, s21 = 0.0, s22 = 0.0;
        if (p.Redi) { s11 = AT(p.sl1, r.bn1, z); s12 = AT(p.sl1, r.bn2, z); s21 = AT(p.sl2, r.bn1, z); s22 = AT(p.sl2, r.bn2, z); }
        double2 ga[NT], gb[NT];
        double z1a[NT], z1b[NT], z2a[NT], z2b[NT];
#pragma unroll
        for (int t = 0; t < NT; t++) {
            ga[t] = AT(s.txy[t], r.be1, z); gb[t] = AT(s.txy[t], be2s, z);
            z1a[t] = z1b[t] = z2a[t] = z2b[t] = 0.0;
            if (p.Redi) {
                z1a[t] = AT(s.trz[t], r.bn1, z); z1b[t] = AT(s.trz[t], r.bn1, z + 1);
                z2a[t] = AT(s.trz[t], r.bn2, z); z2b[t] = AT(s.trz[t], r.bn2, z + 1);
            }
        }
        LOADS_DONE();
        // ---- arithmetic
        const double Kh = (ki1 + ki2) / 2.0;
        double dz;
        if (segs == 3) dz = (he1 + he2) / 2.0;
        else if (segs == 1) dz = he1;
        else dz = he2;
#pragma unroll
        for (int t = 0; t < NT; t++) {
            double SxTz = 0.0, SyTz = 0.0;
            if (p.Redi) {
                const double Tz1 = 0.5 * (z1a[t] + z1b[t]);
                const double Tz2 = 0.5 * (z2a[t] + z2b[t]);
                SxTz = (Tz1 * s11 + Tz2 * s12) / 2.0;
                SyTz = (Tz1 * s21 + Tz2 * s22) / 2.0;
            }
            double Tx, Ty, cc;                                                  // :1020-1062
            if (segs == 3) {
                Tx = 0.5 * (ga[t].x + gb[t].x);
                Ty = 0.5 * (ga[t].y + gb[t].y);
                const double Fx = Kh * (Tx + SxTz * isredi), Fy = Kh * (Ty + SyTz * isredi);
                cc = ((r.dX2 - r.dX1) * Fy - (r.dY2 - r.dY1) * Fx) * dz;
            } else if (segs == 1) {
                Tx = ga[t].x; Ty = ga[t].y;
                const double Fx = Kh * (Tx + SxTz * isredi), Fy = Kh * (Ty + SyTz * isredi);
                cc = (-r.dX1 * Fy + r.dY1 * Fx) * dz;
            } else {
                Tx = gb[t].x; Ty = gb[t].y;
                const double Fx = Kh * (Tx + SxTz * isredi), Fy = Kh * (Ty + SyTz * isredi);
                cc = (r.dX2 * Fy - r.dY2 * Fx) * dz;
            }
            AT(s.eflx[t], be, z) = cc;
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// The slot batch of the node kernels: records of slots j0 .. j0+SB-1 from shared memory, activity at level z, and addresses
// that are valid either way (an inactive slot reads the thread's own column / edge entry 0 of the level).
#define SB 8                                  // slots per batch (tiles whose nodes have more incident edges take further batches)
// ZERO ENTRIES.  An edge field (adf_h, eflx) is written at the (edge, level) pairs of the edge's level mask only; everywhere
// else it keeps the zero of its allocation, and one extra tile behind the last edge tile is never written at all.  A slot that
// is inactive at level z therefore needs no special case in the flux sums: it reads its own entry (zero) — or, for a slot
// beyond the node's degree, an entry of the extra tile — and adds +-0, which leaves every partial sum unchanged bit for bit
// (a sum that starts at +0 never becomes -0 under round-to-nearest).  Only the max/min gathers of the limiter need the
// activity bit: there an inactive slot reads the thread's own column instead of the other node's.
template <int NS>
struct SlotBatchT {
    bool on[NS], second[NS];
    int bo[NS], bed[NS], bo_act[NS];
    __device__ __forceinline__ void setup(const SlotRec* rec, int j0, int z, int own_b) {
#pragma unroll
        for (int j = 0; j < NS; j++) {
            const int4 v = *reinterpret_cast<const int4*>(&rec[(j0 + j) * TILE]);     // one 16-byte shared-memory load (bo, bed, mask lo, mask hi)
            const unsigned long long mk = ((unsigned long long)(unsigned)v.w << 32) | (unsigned)v.z;
            on[j] = (mk >> z) & 1ull;
            second[j] = v.z & 1;
            bo[j] = v.x;                                                       // other node (own column for a slot beyond the node's degree)
            bed[j] = v.y;                                                      // edge (the never-written extra tile for a slot beyond the degree)
            bo_act[j] = on[j] ? v.x : own_b;
        }
    }
};
typedef SlotBatchT<SB> SlotBatch;

// value form of adv_tra_ver_upw1 at interface nz (oce_adv_tra_ver.F90:255-281): t0 = T(nz), tm1 = T(nz-1)
__device__ __forceinline__ double ver_upw1_w(double t0, double tm1, double w, double ar, int nz, int nzmin, int nzmax) {
    if (nz == nzmax) return 0.0;
    if (nz == nzmin) return -w * t0 * ar;
    return -0.5 * (t0 * (w + fabs(w)) + tm1 * (w - fabs(w))) * ar;
}
// value form of adv_tra_ver_qr4c at interface nz (oce_adv_tra_ver.F90:314-356).  The three slopes divide by the level-depth
// differences dzc = Z(nz-1) - Z(nz), dzu = Z(nz) - Z(nz+1), dzd = Z(nz-2) - Z(nz-1), which do not depend on the tracer and
// recur from level to level: the caller keeps (difference, reciprocal) pairs in a window — one true division per level — and
// the quotients are formed by the exact-division sequence div_by (tracer_dev.cuh), like the two divisions by 3.
__device__ __forceinline__ double ver_qr4c_w(double t0, double tm1, double tm2, double tp1, double z0, double zm1,
                                             double dzc, double rdzc, double dzu, double rdzu, double dzd, double rdzd,
                                             double zb, double w, double ar, double pv, int nz, int nzmin, int nzmax) {
    if (nz == nzmax) return 0.0;
    if (nz == nzmin) return -t0 * w * ar;
    if (nz == nzmin + 1 || nz == nzmax - 1) return -0.5 * (tm1 + t0) * w * ar;
    const double qc = div_by(tm1 - t0, dzc, rdzc);
    const double qu = div_by(t0 - tp1, dzu, rdzu);
    const double qd = div_by(tm2 - tm1, dzd, rdzd);
    const double Tmean1 = t0 + div_by((2 * qc + qu) * (zb - z0), 3.0, R3);
    const double Tmean2 = tm1 + div_by((2 * qc + qd) * (zb - zm1), 3.0, R3);
    const double Tmean = (w + fabs(w)) * Tmean1 + (w - fabs(w)) * Tmean2;
    return (-0.5 * (1.0 - pv) * Tmean - pv * (0.5 * (Tmean1 + Tmean2)) * w) * ar;
}

// Low-order solution + vertical antidiffusive flux, per own node column (see k_lo in tracer_kernels.cu for the formulas):
//   fct_LO : oce_adv_tra_driver.F90:83-122;  adf_v : :163-179 with init_zero=.false.
// The column values the vertical stencils need slide down the column in registers — windows over levels z-2 .. z+1 of T,
// ttfAB and Z_3d_n, over z, z+1 of W and area; every iteration loads one new level (z+2) of each.
template <int NT, int VER, bool FCT, int S0>
__device__ __forceinline__ void lo_body(const KP& p, const TrSet& s, const Col& c, const SlotRec* rec, int ndeg, int zlo, int zhi) {
    const double* __restrict__ Wlo = p.We;                         // explicit part for the LO solution
    const double* __restrict__ Wad = p.w_split ? p.W : p.We;       // LO part of the antidiffusive flux: full W (:131)
    const double* __restrict__ Who = FCT ? p.W : p.We;             // high-order flux: W under FCT, Wvel_e without (:163-179)
    const int b = c.b, nzmin = c.nzmin, nzmax = c.nzmax, nlp = p.nlp;
    // windows at the first level z = nzmin: index k holds level z-2+k (clamped into the field)
    double tw[NT][4], aw[NT][4], zw[4], wlo[2], wad[2], who[2], ar[2];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int zz = min(max(nzmin - 2 + k, 1), nlp);
        zw[k] = AT(p.Z3, b, zz);
#pragma unroll
        for (int t = 0; t < NT; t++) { tw[t][k] = AT(s.T[t], b, zz); aw[t][k] = AT(s.Told[t], b, zz); }
    }
#pragma unroll
    for (int k = 0; k < 2; k++) {
        const int zz = min(nzmin + k, nlp);
        wlo[k] = AT(Wlo, b, zz); wad[k] = AT(Wad, b, zz); who[k] = AT(Who, b, zz); ar[k] = AT(p.area, b, zz);
    }
    // level-depth differences of the QR4C slopes and their reciprocals: index 0 = Z(z-2)-Z(z-1), 1 = Z(z-1)-Z(z), 2 = Z(z)-Z(z+1)
    double dzw[3], rdzw[3];
#pragma unroll
    for (int k = 0; k < 3; k++) { dzw[k] = zw[k] - zw[k + 1]; rdzw[k] = 1.0 / dzw[k]; }
    double fv_z[NT];                                               // upwind flux (Wlo) through interface z, left by the iteration above
#pragma unroll
    for (int t = 0; t < NT; t++) fv_z[t] = 0.0;
    bool have_fv = false;
    for (int z = zlo; z <= zhi; z++) {
        if (z < nzmin || z > nzmax - 1) continue;
        // ---- loads of this level: level z+2 for the windows, the level's own scalars, the slot batch
        const int z2 = min(z + 2, nlp);
        const double n_z3 = AT(p.Z3, b, z2);
        const double n_wlo = AT(Wlo, b, z2), n_wad = AT(Wad, b, z2), n_who = AT(Who, b, z2), n_ar = AT(p.area, b, z2);
        const double zb = AT(p.zbar3, b, z);
        double n_t[NT], n_a[NT];
#pragma unroll
        for (int t = 0; t < NT; t++) { n_t[t] = AT(s.T[t], b, z2); n_a[t] = AT(s.Told[t], b, z2); }
        double hn = 0.0, hnn = 1.0, asv = 1.0;
        double acc[NT];
#pragma unroll
        for (int t = 0; t < NT; t++) acc[t] = 0.0;
        double rhnn = 1.0, rasv = 1.0;
        if (FCT) {
            hn = AT(p.hnode, b, z); hnn = AT(p.hnode_new, b, z); asv = AT(p.areasvol, b, z);
            rhnn = 1.0 / hnn; rasv = 1.0 / asv;                    // shared by the tracers of the pass
            for (int j0 = 0; j0 < (S0 < SB ? S0 : ndeg); j0 += SB) {
                SlotBatchT<S0> sb;
                sb.setup(rec, j0, z, b);
                double f[NT][S0];
#pragma unroll
                for (int j = 0; j < S0; j++)
#pragma unroll
                    for (int t = 0; t < NT; t++) f[t][j] = AT(s.eflx[t], sb.bed[j], z);        // upwind edge flux, once per edge by k_edge_adv2
                LOADS_DONE();
#pragma unroll
                for (int j = 0; j < S0; j++)
#pragma unroll
                    for (int t = 0; t < NT; t++) acc[t] = sb.second[j] ? acc[t] - f[t][j] : acc[t] + f[t][j];   // inactive slot: f = 0
            }
        }
        // ---- arithmetic; window index: [0] = z-2, [1] = z-1, [2] = z, [3] = z+1
#pragma unroll
        for (int t = 0; t < NT; t++) {
            const double t0 = tw[t][2], tm1 = tw[t][1], tp1 = tw[t][3];
            if (FCT) {
                const double fv = have_fv ? fv_z[t] : ver_upw1_w(t0, tm1, wlo[0], ar[0], z, nzmin, nzmax);
                const double fvn = ver_upw1_w(tp1, t0, wlo[1], ar[1], z + 1, nzmin, nzmax);
                AT(s.LO[t], b, z) = div_by(t0 * hn + div_by((acc[t] + (fv - fvn)) * p.dt, asv, rasv), hnn, rhnn);
                fv_z[t] = fvn;
            }
            const double flo = FCT ? ver_upw1_w(t0, tm1, wad[0], ar[0], z, nzmin, nzmax) : 0.0;
            double ho;
            if (VER == 0) ho = ver_qr4c_w(aw[t][2], aw[t][1], aw[t][0], aw[t][3], zw[2], zw[1], dzw[1], rdzw[1], dzw[2], rdzw[2], dzw[0], rdzw[0],
                                          zb, who[0], ar[0], p.pv, z, nzmin, nzmax);
            else if (VER == 1) ho = ver_cdiff(p, s.Told[t], Who, b, z, nzmin);
            else if (VER == 2) ho = ver_ppm(p, s.Told[t], Who, b, z, nzmin, nzmax);
            else ho = ver_upw1_w(aw[t][2], aw[t][1], who[0], ar[0], z, nzmin, nzmax);
            AT(s.adf_v[t], b, z) = ho - flo;
            tw[t][0] = tw[t][1]; tw[t][1] = tw[t][2]; tw[t][2] = tw[t][3]; tw[t][3] = n_t[t];
            aw[t][0] = aw[t][1]; aw[t][1] = aw[t][2]; aw[t][2] = aw[t][3]; aw[t][3] = n_a[t];
        }
        zw[0] = zw[1]; zw[1] = zw[2]; zw[2] = zw[3]; zw[3] = n_z3;
        dzw[0] = dzw[1]; dzw[1] = dzw[2]; dzw[2] = zw[2] - zw[3];
        rdzw[0] = rdzw[1]; rdzw[1] = rdzw[2]; rdzw[2] = VER == 0 ? 1.0 / dzw[2] : 0.0;
        wlo[0] = wlo[1]; wlo[1] = n_wlo; wad[0] = wad[1]; wad[1] = n_wad; who[0] = who[1]; who[1] = n_who; ar[0] = ar[1]; ar[1] = n_ar;
        have_fv = true;
    }
}
template <int NT, int VER, bool FCT>
__global__ void __launch_bounds__(32 * WPB, LO_MB) k_lo2(KP p, TrSet s) {
    extern __shared__ SlotRec smrec[];
    Col c;
    if (!col_setup(p, c)) return;
    int ndeg = 0;
    const SlotRec* rec = FCT ? stage_slots(p, c, smrec, ndeg) : nullptr;
    const int zlo = warp_min_i(c.valid ? c.nzmin : 255), zhi = warp_max_i(c.valid ? c.nzmax - 1 : 0);
    if (!c.valid) return;
    if (!FCT || ndeg <= 6) lo_body<NT, VER, FCT, 6>(p, s, c, rec, ndeg, zlo, zhi);
    else if (ndeg == 7) lo_body<NT, VER, FCT, 7>(p, s, c, rec, ndeg, zlo, zhi);
    else lo_body<NT, VER, FCT, SB>(p, s, c, rec, ndeg, zlo, zhi);
}

// ---------------------------------------------------------------------------------------------------
// Zalesak limiter, parts a1-a3 and b1-b2 of oce_tra_adv_fct (oce_adv_tra_fct.F90:93-276), per own node column.  Level z+1 is
// gathered before level z is finished, so the 3-point vertical max/min (a3, vlimit=1) is a register window.
template <int NT, int S0>
__device__ __forceinline__ void fct_bounds_body(const KP& p, const TrSet& s, const Col& c, const SlotRec* rec, int ndeg, int zlo, int zhi) {
    const double flux_eps = 1e-16, bignumber = 1e3;
    const int b = c.b;
    // window: index 0 = level z-2, 1 = level z-1 (to be finished), 2 = level z (just gathered)
    double Mx[NT][3], Mn[NT][3], pp1[NT], pm1[NT], lo1[NT], avz[NT];
#pragma unroll
    for (int t = 0; t < NT; t++) { Mx[t][0] = Mx[t][1] = Mx[t][2] = 0.0; Mn[t][0] = Mn[t][1] = Mn[t][2] = 0.0; pp1[t] = pm1[t] = lo1[t] = avz[t] = 0.0; }
    bool have_av = false;
    for (int z = zlo; z <= zhi + 1; z++) {
        const bool act = z >= c.nzmin && z <= c.nzmax - 1;
        const int zf = z - 1;                                                      // level finished in this iteration
        const bool fin = zf >= c.nzmin && zf <= c.nzmax - 1;
        if (!act && !fin) continue;
        double pp[NT], pm[NT], lo[NT];
        const double asvf = AT(p.areasvol, b, fin ? zf : z);
        if (act) {
            // ---- loads of level z
            double av[NT], avn[NT], tv[NT];
#pragma unroll
            for (int t = 0; t < NT; t++) {
                av[t] = have_av ? avz[t] : AT(s.adf_v[t], b, z);
                avn[t] = AT(s.adf_v[t], b, z + 1);
                lo[t] = AT(s.LO[t], b, z);
                tv[t] = AT(s.T[t], b, z);
            }
#pragma unroll
            for (int t = 0; t < NT; t++) {
                pp[t] = dmax(0.0, av[t]) + dmax(0.0, -avn[t]);                    // :235-236
                pm[t] = dmin(0.0, av[t]) + dmin(0.0, -avn[t]);
                double M = dmax(lo[t], tv[t]), m = dmin(lo[t], tv[t]);            // a1 :99-100
                if (z >= c.nlmin) { M = dmax(M, -bignumber); m = dmin(m, bignumber); }   // a2 :116-121
                Mx[t][2] = M; Mn[t][2] = m;
                avz[t] = avn[t];
            }
            for (int j0 = 0; j0 < (S0 < SB ? S0 : ndeg); j0 += SB) {
                SlotBatchT<S0> sb;
                sb.setup(rec, j0, z, b);
                double f[NT][S0], lo_o[NT][S0], t_o[NT][S0];
#pragma unroll
                for (int j = 0; j < S0; j++)
#pragma unroll
                    for (int t = 0; t < NT; t++) {
                        f[t][j] = AT(s.adf_h[t], sb.bed[j], z);
                        lo_o[t][j] = AT(s.LO[t], sb.bo_act[j], z);
                        t_o[t][j] = AT(s.T[t], sb.bo_act[j], z);
                    }
                LOADS_DONE();
#pragma unroll
                for (int j = 0; j < S0; j++)
#pragma unroll
                    for (int t = 0; t < NT; t++) {
                        const double fs = sb.second[j] ? -f[t][j] : f[t][j];               // inactive slot: +-0, and the own column below
                        pp[t] = pp[t] + dmax(0.0, fs);                                       // b1 :258-261
                        pm[t] = pm[t] + dmin(0.0, fs);
                        Mx[t][2] = dmax(Mx[t][2], dmax(lo_o[t][j], t_o[t][j]));
                        Mn[t][2] = dmin(Mn[t][2], dmin(lo_o[t][j], t_o[t][j]));
                    }
            }
            have_av = true;
        }
        if (fin) {
            const double asv = asvf, rasv = 1.0 / asv;
#pragma unroll
            for (int t = 0; t < NT; t++) {
                double qmax = Mx[t][1], qmin = Mn[t][1];
                if (zf > c.nzmin && zf < c.nzmax - 1) {                            // a3, vlimit=1 :153-156
                    qmax = dmax(dmax(Mx[t][0], qmax), Mx[t][2]);
                    qmin = dmin(dmin(Mn[t][0], qmin), Mn[t][2]);
                }
                qmax = qmax - lo1[t]; qmin = qmin - lo1[t];
                double flux = div_by(pp1[t] * p.dt, asv, rasv) + flux_eps;         // b2 :271-274
                const double rp = dmin(1.0, qmax / flux);
                flux = div_by(pm1[t] * p.dt, asv, rasv) - flux_eps;
                AT(s.fpm[t], b, zf) = make_double2(rp, dmin(1.0, qmin / flux));
            }
        }
#pragma unroll
        for (int t = 0; t < NT; t++) {
            Mx[t][0] = Mx[t][1]; Mx[t][1] = Mx[t][2]; Mn[t][0] = Mn[t][1]; Mn[t][1] = Mn[t][2];
            if (act) { pp1[t] = pp[t]; pm1[t] = pm[t]; lo1[t] = lo[t]; }
        }
    }
}
template <int NT>
__global__ void __launch_bounds__(32 * WPB, FCTB_MB) k_fct_bounds2(KP p, TrSet s) {
    extern __shared__ SlotRec smrec[];
    Col c;
    if (!col_setup(p, c)) return;
    int ndeg = 0;
    const SlotRec* rec = stage_slots(p, c, smrec, ndeg);
    const int zlo = warp_min_i(c.valid ? c.nzmin : 255), zhi = warp_max_i(c.valid ? c.nzmax - 1 : 0);
    if (!c.valid) return;
    // the tile's largest node degree picks the slot count of the (unrolled) batch: 6 on a regular triangulation, where 8 slots would
    // spend a quarter of the kernel's instructions on empty slots
    if (ndeg <= 6) fct_bounds_body<NT, 6>(p, s, c, rec, ndeg, zlo, zhi);
    else if (ndeg == 7) fct_bounds_body<NT, 7>(p, s, c, rec, ndeg, zlo, zhi);
    else fct_bounds_body<NT, SB>(p, s, c, rec, ndeg, zlo, zhi);
}

// ---------------------------------------------------------------------------------------------------
// b3 limiting + oce_tra_adv_flux2dtracer (oce_adv_tra_driver.F90:223-268) + del_ttf assembly (oce_ale_tracer.F90:240) +
// the fluxes of diff_part_hor_redi (:934-1077) + explicit vertical diffusion / Redi term + T* update (:294-305), per own
// node column.  The limited vertical flux of interface z+1 is kept for level z+1 (register window over fct_plus/minus).
// The diffusive fluxes are gathered together with the advective ones and wait in registers for their turn in the sum (the
// reference adds them after the vertical part).
__device__ __forceinline__ double limited_fv2(double flux, double2 fm1, double2 f0, bool top, bool bottom) {
    if (bottom) return flux;                                  // interface nzmax: never limited (and zero)
    double ae = 1.0;
    if (top) { if (flux >= 0.0) ae = dmin(ae, f0.x); else ae = dmin(ae, f0.y); }
    else if (flux >= 0.0) { ae = dmin(ae, fm1.y); ae = dmin(ae, f0.x); }
    else { ae = dmin(ae, fm1.x); ae = dmin(ae, f0.y); }
    return ae * flux;
}
template <int NT, bool FCT, bool EXPL, int S0>
__device__ __forceinline__ void update_body(const KP& p, const TrSet& s, const Col& c, const SlotRec* rec, int ndeg, int zlo, int zhi) {
    const int b = c.b;
    const bool hd = p.do_hor_diff;
    // window over the own limiter factors and the vertical flux: fpm(z-1), fpm(z); adf_v(z) and the limited fv(z) of the iteration above
    double2 fprev[NT], fcur[NT];
    double fv_z[NT], av_z[NT];
#pragma unroll
    for (int t = 0; t < NT; t++) { fprev[t] = make_double2(0.0, 0.0); fcur[t] = make_double2(0.0, 0.0); fv_z[t] = 0.0; av_z[t] = 0.0; }
    bool primed = false;
    for (int z = zlo; z <= zhi; z++) {
        if (z < c.nzmin || z > c.nzmax - 1) continue;
        // ---- loads of this level
        const double asv = AT(p.areasvol, b, z);
        const double hn = AT(p.hnode, b, z), hnn = AT(p.hnode_new, b, z);
        double tv[NT], lov[NT], avn[NT], rv[NT];
        double2 fnext[NT];
        const int zn = z + 1 <= c.nzmax - 1 ? z + 1 : z;               // fpm(z+1) is not used at the bottom: read level z again
#pragma unroll
        for (int t = 0; t < NT; t++) {
            tv[t] = AT(s.T[t], b, z);
            lov[t] = FCT ? AT(s.LO[t], b, z) : 0.0;
            avn[t] = AT(s.adf_v[t], b, z + 1);
            if (!primed) { av_z[t] = AT(s.adf_v[t], b, z); if (FCT) fcur[t] = AT(s.fpm[t], b, z); }
            fnext[t] = FCT ? AT(s.fpm[t], b, zn) : make_double2(0.0, 0.0);
            rv[t] = p.Redi ? AT(s.redi_v[t], b, z) : 0.0;
        }
        const double rasv = 1.0 / asv;
        double acc[NT];
#pragma unroll
        for (int t = 0; t < NT; t++) acc[t] = 0.0;
        // first batch of slots: advective and diffusive fluxes gathered together
        SlotBatchT<S0> sb0;
        double dfl[NT][S0];
        {
            sb0.setup(rec, 0, z, b);
            double ah[NT][S0];
            double2 fo[NT][S0];
#pragma unroll
            for (int j = 0; j < S0; j++)
#pragma unroll
                for (int t = 0; t < NT; t++) {
                    ah[t][j] = AT(s.adf_h[t], sb0.bed[j], z);
                    dfl[t][j] = hd ? AT(s.eflx[t], sb0.bed[j], z) : 0.0;
                    fo[t][j] = FCT ? AT(s.fpm[t], sb0.bo[j], z) : make_double2(0.0, 0.0);
                }
            LOADS_DONE();
#pragma unroll
            for (int j = 0; j < S0; j++)                                         // dttf_h
#pragma unroll
                for (int t = 0; t < NT; t++) {
                    const double flux = ah[t][j];
                    double fl = flux;
                    if (FCT) {
                        // :313-330  flux >= 0: min(fct_plus(n1), fct_minus(n2)) else min(fct_minus(n1), fct_plus(n2)); this node is n2 when `second`
                        const bool self_plus = (flux >= 0.0) != sb0.second[j];
                        double ae = dmin(1.0, self_plus ? fcur[t].x : fcur[t].y);
                        ae = dmin(ae, self_plus ? fo[t][j].y : fo[t][j].x);
                        fl = ae * flux;
                    }
                    const double d = div_by(fl * p.dt, asv, rasv);                    // inactive slot: flux = 0 -> d = 0
                    acc[t] = sb0.second[j] ? acc[t] - d : acc[t] + d;
                }
        }
        for (int j0 = SB; S0 == SB && j0 < ndeg; j0 += SB) {                                 // nodes with more than SB incident edges
            SlotBatch sb;
            sb.setup(rec, j0, z, b);
#pragma unroll
            for (int j = 0; j < S0; j++)
#pragma unroll
                for (int t = 0; t < NT; t++) {
                    const double flux = AT(s.adf_h[t], sb.bed[j], z);
                    double fl = flux;
                    if (FCT) {
                        const double2 fo = AT(s.fpm[t], sb.bo[j], z);
                        const bool self_plus = (flux >= 0.0) != sb.second[j];
                        double ae = dmin(1.0, self_plus ? fcur[t].x : fcur[t].y);
                        ae = dmin(ae, self_plus ? fo.y : fo.x);
                        fl = ae * flux;
                    }
                    const double d = div_by(fl * p.dt, asv, rasv);
                    acc[t] = sb.second[j] ? acc[t] - d : acc[t] + d;
                }
        }
#pragma unroll
        for (int t = 0; t < NT; t++) {                                           // dttf_v, then del_ttf = dttf_h + dttf_v
            double fv, fvn;
            if (FCT) {
                fv = primed ? fv_z[t] : limited_fv2(av_z[t], fprev[t], fcur[t], z == c.nzmin, false);
                fvn = limited_fv2(avn[t], fcur[t], fnext[t], false, z + 1 == c.nzmax);
            } else { fv = av_z[t]; fvn = avn[t]; }
            double dv = 0.0;
            if (FCT) dv = 0.0 - tv[t] * hn + lov[t] * hnn;
            dv = dv + div_by((fv - fvn) * p.dt, asv, rasv);
            acc[t] = acc[t] + dv;
            fv_z[t] = fvn; av_z[t] = avn[t];
        }
        if (hd) {                                                                // diff_part_hor_redi: fluxes from k_edge_diff2
#pragma unroll
            for (int j = 0; j < S0; j++)
#pragma unroll
                for (int t = 0; t < NT; t++) {
                    const double cc = dfl[t][j];
                    const double rhs = sb0.second[j] ? 0.0 - cc : 0.0 + cc;
                    acc[t] = acc[t] + div_by(rhs * p.dt, asv, rasv);
                }
            for (int j0 = SB; S0 == SB && j0 < ndeg; j0 += SB) {
                SlotBatch sb;
                sb.setup(rec, j0, z, b);
#pragma unroll
                for (int j = 0; j < S0; j++)
#pragma unroll
                    for (int t = 0; t < NT; t++) {
                        const double cc = AT(s.eflx[t], sb.bed[j], z);
                        const double rhs = sb.second[j] ? 0.0 - cc : 0.0 + cc;
                        acc[t] = acc[t] + div_by(rhs * p.dt, asv, rasv);
                    }
            }
        }
#pragma unroll
        for (int t = 0; t < NT; t++) {
            if (EXPL) {                                                          // diff_ver_part_expl_ale :381-392
                const double* __restrict__ T = s.T[t];
                const int nl1 = c.nzmax - 1, trn = s.tr_num[t];
                const double vd = expl_vd2(p, trn, T, c.n, b, z, c.nzmin, nl1);
                const double dzb = AT(p.zbar3, b, z) - AT(p.zbar3, b, z + 1);
                double term;
                if (z <= nl1 - 1) term = (vd - expl_vd2(p, trn, T, c.n, b, z + 1, c.nzmin, nl1)) / dzb * p.dt;
                else term = (vd / dzb) * p.dt;
                acc[t] = acc[t] + div_by(term, asv, rasv);
            }
            if (p.Redi) acc[t] = acc[t] + rv[t];                                 // diff_ver_part_redi_expl :288
            const double del = acc[t] + tv[t] * (hn - hnn);                      // T*  :302-305
            AT(s.tstar[t], b, z) = tv[t] + del / hnn;
            if (FCT) { fprev[t] = fcur[t]; fcur[t] = fnext[t]; }
        }
        primed = true;
    }
}
template <int NT, bool FCT, bool EXPL>
__global__ void __launch_bounds__(32 * WPB, UPD_MB) k_update2(KP p, TrSet s) {
    extern __shared__ SlotRec smrec[];
    Col c;
    if (!col_setup(p, c)) return;
    int ndeg = 0;
    const SlotRec* rec = stage_slots(p, c, smrec, ndeg);
    const int zlo = warp_min_i(c.valid ? c.nzmin : 255), zhi = warp_max_i(c.valid ? c.nzmax - 1 : 0);
    if (!c.valid) return;
    if (ndeg <= 6) update_body<NT, FCT, EXPL, 6>(p, s, c, rec, ndeg, zlo, zhi);
    else if (ndeg == 7) update_body<NT, FCT, EXPL, 7>(p, s, c, rec, ndeg, zlo, zhi);
    else update_body<NT, FCT, EXPL, SB>(p, s, c, rec, ndeg, zlo, zhi);
}

// ---------------------------------------------------------------------------------------------------
// host side: launches
#define KL(name, ...) do { ProfScope ps_(name, st); __VA_ARGS__; LAUNCH_CHECK(); } while (0)
static bool g_attr2 = false;
template <class K> static void smem_attr(K k) { CUDA_CHECK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024)); }
static void set_attrs2() {
    if (g_attr2) return;
    g_attr2 = true;
#define A_LO(V) smem_attr(k_lo2<1, V, true>); smem_attr(k_lo2<2, V, true>); smem_attr(k_lo2<1, V, false>); smem_attr(k_lo2<2, V, false>);
    A_LO(0) A_LO(1) A_LO(2) A_LO(3)
#undef A_LO
    smem_attr(k_fct_bounds2<1>); smem_attr(k_fct_bounds2<2>);
    smem_attr(k_update2<1, true, false>); smem_attr(k_update2<2, true, false>); smem_attr(k_update2<1, false, false>); smem_attr(k_update2<2, false, false>);
    smem_attr(k_update2<1, true, true>); smem_attr(k_update2<2, true, true>); smem_attr(k_update2<1, false, true>); smem_attr(k_update2<2, false, true>);
}
static size_t slot_smem(const TracerCtx& c) { return (size_t)WPB * slot_cap(c.maxdeg) * TILE * sizeof(SlotRec); }
static int col_grid(const TracerCtx& c) { return (c.myNtiles + WPB - 1) / WPB; }

// tracers per launch of each phase: the edge kernels share the volume flux between two tracers, the node kernels carry
// 8 slots x 3 gathers per tracer in registers and run one tracer per launch by default.  FB200_NT_<PHASE>=1|2 overrides
// (A/B measurements; results do not depend on it).
static int phase_nt(const char* name, int dflt) {
    const std::string key = std::string("FB200_NT_") + name;
    const char* e = getenv(key.c_str());
    if (e && (e[0] == '1' || e[0] == '2')) return e[0] - '0';
    return dflt;
}
static TrSet one_of(const TrSet& s, int k) {
    TrSet o = s;
    o.T[0] = s.T[k]; o.Told[0] = s.Told[k]; o.LO[0] = s.LO[k]; o.adf_h[0] = s.adf_h[k]; o.adf_v[0] = s.adf_v[k]; o.tstar[0] = s.tstar[k];
    o.eflx[0] = s.eflx[k]; o.trz[0] = s.trz[k]; o.txy[0] = s.txy[k]; o.fpm[0] = s.fpm[k]; o.redi_v[0] = s.redi_v[k]; o.tr_num[0] = s.tr_num[k];
    return o;
}
bool v2_usable(const TracerCtx& c) { return slot_smem(c) <= 160 * 1024; }

void v2_ab2(TracerCtx& c, const TrSet& s, int nt, cudaStream_t st) {
    const size_t nf = (size_t)c.Ntiles * c.nlp * TILE;
    const double ca = -(0.5 + c.opt.epsilon), cb = 1.5 + c.opt.epsilon;
    // the fields of a group are contiguous (tracer-major) in c.T / c.Told
    KL(nt == 2 ? "k_ab2_x2" : "k_ab2", k_ab2_g<<<148 * 8, 256, 0, st>>>(s.Told[0], s.T[0], nf * nt, ca, cb));
}
void v2_grad_elem(TracerCtx& c, const KP& p, const TrSet& s, int nt, int which, cudaStream_t st) {
    static const int pnt = phase_nt("GRAD", 2);
    if (nt == 2 && pnt == 1) { v2_grad_elem(c, p, one_of(s, 0), 1, which, st); v2_grad_elem(c, p, one_of(s, 1), 1, which, st); return; }
    const int g = (c.myE + 127) / 128;
    if (nt == 2) KL("k_grad_elem_x2", k_grad_elem2<2><<<g, 128, 0, st>>>(p, s, which));
    else KL("k_grad_elem", k_grad_elem2<1><<<g, 128, 0, st>>>(p, s, which));
}
void v2_grad_z(TracerCtx& c, const KP& p, const TrSet& s, int nt, cudaStream_t st) {
    const int g = (c.N + 127) / 128;
    if (nt == 2) KL("k_grad_z_x2", k_grad_z2<2><<<g, 128, 0, st>>>(p, s));
    else KL("k_grad_z", k_grad_z2<1><<<g, 128, 0, st>>>(p, s));
}
void v2_edge_adv(TracerCtx& c, const KP& p, const TrSet& s, int nt, bool fct, cudaStream_t st) {
    static const int pnt = phase_nt("EDGE", 2);
    if (nt == 2 && pnt == 1) { v2_edge_adv(c, p, one_of(s, 0), 1, fct, st); v2_edge_adv(c, p, one_of(s, 1), 1, fct, st); return; }
    const int g = (c.myEd + 127) / 128;
    if (nt == 2) { if (fct) KL("k_ho_edge_x2", k_edge_adv2<2, true><<<g, 128, 0, st>>>(p, s)); else KL("k_ho_edge_x2", k_edge_adv2<2, false><<<g, 128, 0, st>>>(p, s)); }
    else { if (fct) KL("k_ho_edge", k_edge_adv2<1, true><<<g, 128, 0, st>>>(p, s)); else KL("k_ho_edge", k_edge_adv2<1, false><<<g, 128, 0, st>>>(p, s)); }
}
void v2_edge_diff(TracerCtx& c, const KP& p, const TrSet& s, int nt, cudaStream_t st) {
    static const int pnt = phase_nt("DIFF", 2);
    if (nt == 2 && pnt == 1) { v2_edge_diff(c, p, one_of(s, 0), 1, st); v2_edge_diff(c, p, one_of(s, 1), 1, st); return; }
    const int g = (c.myEd + 127) / 128;
    if (nt == 2) KL("k_diff_edge_x2", k_edge_diff2<2><<<g, 128, 0, st>>>(p, s));
    else KL("k_diff_edge", k_edge_diff2<1><<<g, 128, 0, st>>>(p, s));
}
void v2_lo(TracerCtx& c, const KP& p, const TrSet& s, int nt, bool fct, cudaStream_t st) {
    static const int pnt = phase_nt("LO", 2);
    if (nt == 2 && pnt == 1) { v2_lo(c, p, one_of(s, 0), 1, fct, st); v2_lo(c, p, one_of(s, 1), 1, fct, st); return; }
    set_attrs2();
    const int g = col_grid(c);
    const size_t sm = fct ? slot_smem(c) : 0;
#define LO2(NT, V) do { if (fct) KL(NT == 2 ? "k_lo_x2" : "k_lo", k_lo2<NT, V, true><<<g, 32 * WPB, sm, st>>>(p, s)); else KL(NT == 2 ? "k_lo_x2" : "k_lo", k_lo2<NT, V, false><<<g, 32 * WPB, sm, st>>>(p, s)); } while (0)
#define LO2V(NT) switch (c.opt.tra_adv_ver) { case 0: LO2(NT, 0); break; case 1: LO2(NT, 1); break; case 2: LO2(NT, 2); break; default: LO2(NT, 3); }
    if (nt == 2) { LO2V(2) } else { LO2V(1) }
#undef LO2V
#undef LO2
}
void v2_fct_bounds(TracerCtx& c, const KP& p, const TrSet& s, int nt, cudaStream_t st) {
    static const int pnt = phase_nt("BOUNDS", 1);
    if (nt == 2 && pnt == 1) { v2_fct_bounds(c, p, one_of(s, 0), 1, st); v2_fct_bounds(c, p, one_of(s, 1), 1, st); return; }
    set_attrs2();
    if (nt == 2) KL("k_fct_bounds_x2", k_fct_bounds2<2><<<col_grid(c), 32 * WPB, slot_smem(c), st>>>(p, s));
    else KL("k_fct_bounds", k_fct_bounds2<1><<<col_grid(c), 32 * WPB, slot_smem(c), st>>>(p, s));
}
void v2_update(TracerCtx& c, const KP& p, const TrSet& s, int nt, bool fct, bool expl, cudaStream_t st) {
    static const int pnt = phase_nt("UPDATE", 1);
    if (nt == 2 && pnt == 1) { v2_update(c, p, one_of(s, 0), 1, fct, expl, st); v2_update(c, p, one_of(s, 1), 1, fct, expl, st); return; }
    set_attrs2();
    const int g = col_grid(c);
    const size_t sm = slot_smem(c);
#define UP2(NT) do { \
        if (fct && !expl) KL(NT == 2 ? "k_update_x2" : "k_update", k_update2<NT, true, false><<<g, 32 * WPB, sm, st>>>(p, s)); \
        else if (fct) KL(NT == 2 ? "k_update_x2" : "k_update", k_update2<NT, true, true><<<g, 32 * WPB, sm, st>>>(p, s)); \
        else if (!expl) KL(NT == 2 ? "k_update_x2" : "k_update", k_update2<NT, false, false><<<g, 32 * WPB, sm, st>>>(p, s)); \
        else KL(NT == 2 ? "k_update_x2" : "k_update", k_update2<NT, false, true><<<g, 32 * WPB, sm, st>>>(p, s)); } while (0)
    if (nt == 2) UP2(2); else UP2(1);
#undef UP2
}
#undef KL
