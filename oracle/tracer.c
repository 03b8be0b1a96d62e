/* ORACLE (test infrastructure) — tracer transport restated from the reference, routine by routine.
 * See oracle.h for the rules.  do_Xmoment is fixed to 1 (the only value the reference's call site passes,
 * oce_ale_tracer.F90:236), so "x**do_Xmoment" is x. */
#include "oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define NL (o->nl)
#define LM (o->nl - 1)
#define T2(a, nz, n) (a)[(size_t)((n) - 1) * (size_t)LM + (size_t)((nz) - 1)]               /* (nl-1, N)   */
#define W2(a, nz, n) (a)[(size_t)((n) - 1) * (size_t)NL + (size_t)((nz) - 1)]               /* (nl,   N)   */
#define V3(a, c, nz, e) (a)[((size_t)((e) - 1) * (size_t)LM + (size_t)((nz) - 1)) * 2 + ((c) - 1)]   /* (2,nl-1,E) */
#define G4(a, c, nz, e) (a)[((size_t)((e) - 1) * (size_t)LM + (size_t)((nz) - 1)) * 4 + ((c) - 1)]   /* (4,nl-1,Ed)*/
#define S3(a, c, nz, n) (a)[((size_t)((n) - 1) * (size_t)LM + (size_t)((nz) - 1)) * 3 + ((c) - 1)]   /* (3,nl-1,N) */
#define I2(a, ld, i, j) (a)[(size_t)((j) - 1) * (size_t)(ld) + (size_t)((i) - 1)]
static const double r_earth = 6367500.0;   /* oce_modules.F90:16 */
static const double vcpw = 4.2e6;          /* oce_modules.F90:18 */
static inline double dmax(double a, double b) { return a > b ? a : b; }
static inline double dmin(double a, double b) { return a < b ? a : b; }
static inline int imax(int a, int b) { return a > b ? a : b; }
static inline int imin(int a, int b) { return a < b ? a : b; }

/* oce_tracer_mod.F90:19-45 */
void o_tracer_gradient_elements(oracle_t* o, const double* ttf) {
    for (int elem = 1; elem <= o->myDim_elem2D; elem++) {
        const int* en = &I2(o->elem2D_nodes, 3, 1, elem);
        int nzmin = o->ulevels[elem - 1], nzmax = o->nlevels[elem - 1];
        const double* g = &I2(o->gradient_sca, 6, 1, elem);
        for (int nz = nzmin; nz <= nzmax - 1; nz++) {
            V3(o->tr_xy, 1, nz, elem) = g[0] * T2(ttf, nz, en[0]) + g[1] * T2(ttf, nz, en[1]) + g[2] * T2(ttf, nz, en[2]);
            V3(o->tr_xy, 2, nz, elem) = g[3] * T2(ttf, nz, en[0]) + g[4] * T2(ttf, nz, en[1]) + g[5] * T2(ttf, nz, en[2]);
        }
    }
}

/* oce_tracer_mod.F90:124-153 */
void o_tracer_gradient_z(oracle_t* o, const double* ttf) {
    int N = o->myDim_nod2D + o->eDim_nod2D;
    for (int n = 1; n <= N; n++) {
        int nzmax = o->nlevels_nod2D[n - 1], nzmin = o->ulevels_nod2D[n - 1];
        for (int nz = nzmin + 1; nz <= nzmax - 1; nz++) {
            double dz = 0.5 * (T2(o->hnode_new, nz - 1, n) + T2(o->hnode_new, nz, n));
            W2(o->tr_z, nz, n) = (T2(ttf, nz - 1, n) - T2(ttf, nz, n)) / dz;
        }
        W2(o->tr_z, nzmin, n) = 0.0;
        W2(o->tr_z, nzmax, n) = 0.0;
    }
}

static void node_mean_grad(oracle_t* o, int node, int nz, double* gx, double* gy) {   /* oce_muscl_adv.F90:313-328 */
    double tvol = 0.0, tx = 0.0, ty = 0.0;
    for (int k = 1; k <= o->nod_in_elem2D_num[node - 1]; k++) {
        int elem = I2(o->nod_in_elem2D, o->nod_in_elem2D_max, k, node);
        if (o->nlevels[elem - 1] - 1 < nz || nz < o->ulevels[elem - 1]) continue;
        tvol = tvol + o->elem_area[elem - 1];
        tx = tx + V3(o->tr_xy, 1, nz, elem) * o->elem_area[elem - 1];
        ty = ty + V3(o->tr_xy, 2, nz, elem) * o->elem_area[elem - 1];
    }
    *gx = tx / tvol; *gy = ty / tvol;
}

/* oce_muscl_adv.F90:285-447 */
void o_fill_up_dn_grad(oracle_t* o) {
    double* g = o->edge_up_dn_grad;
    for (int edge = 1; edge <= o->myDim_edge2D; edge++) {
        int ed[2] = {I2(o->edges, 2, 1, edge), I2(o->edges, 2, 2, edge)};
        int up = I2(o->edge_up_dn_tri, 2, 1, edge), dn = I2(o->edge_up_dn_tri, 2, 2, edge);
        if (up != 0 && dn != 0) {
            int nzmin = imax(o->ulevels_nod2D_max[ed[0] - 1], o->ulevels_nod2D_max[ed[1] - 1]);
            int nzmax = imin(o->nlevels_nod2D_min[ed[0] - 1], o->nlevels_nod2D_min[ed[1] - 1]);
            for (int nz = o->ulevels_nod2D[ed[0] - 1]; nz <= nzmin - 1; nz++) node_mean_grad(o, ed[0], nz, &G4(g, 1, nz, edge), &G4(g, 3, nz, edge));
            for (int nz = o->ulevels_nod2D[ed[1] - 1]; nz <= nzmin - 1; nz++) node_mean_grad(o, ed[1], nz, &G4(g, 2, nz, edge), &G4(g, 4, nz, edge));
            for (int nz = nzmin; nz <= nzmax - 1; nz++) {
                G4(g, 1, nz, edge) = V3(o->tr_xy, 1, nz, up); G4(g, 2, nz, edge) = V3(o->tr_xy, 1, nz, dn);
                G4(g, 3, nz, edge) = V3(o->tr_xy, 2, nz, up); G4(g, 4, nz, edge) = V3(o->tr_xy, 2, nz, dn);
            }
            for (int nz = nzmax; nz <= o->nlevels_nod2D[ed[0] - 1] - 1; nz++) node_mean_grad(o, ed[0], nz, &G4(g, 1, nz, edge), &G4(g, 3, nz, edge));
            for (int nz = nzmax; nz <= o->nlevels_nod2D[ed[1] - 1] - 1; nz++) node_mean_grad(o, ed[1], nz, &G4(g, 2, nz, edge), &G4(g, 4, nz, edge));
        } else {
            for (int nz = o->ulevels_nod2D[ed[0] - 1]; nz <= o->nlevels_nod2D[ed[0] - 1] - 1; nz++) node_mean_grad(o, ed[0], nz, &G4(g, 1, nz, edge), &G4(g, 3, nz, edge));
            for (int nz = o->ulevels_nod2D[ed[1] - 1]; nz <= o->nlevels_nod2D[ed[1] - 1] - 1; nz++) node_mean_grad(o, ed[1], nz, &G4(g, 2, nz, edge), &G4(g, 4, nz, edge));
        }
    }
}

/* common edge prologue of the three horizontal schemes (oce_adv_tra_hor.F90:89-130) */
typedef struct { int en1, en2, el1, el2, nl1, nu1, nl2, nu2, nl12, nu12; double dX1, dY1, dX2, dY2, a; } edge_t;
static void edge_setup(oracle_t* o, int edge, edge_t* e) {
    e->en1 = I2(o->edges, 2, 1, edge); e->en2 = I2(o->edges, 2, 2, edge);
    e->el1 = I2(o->edge_tri, 2, 1, edge); e->el2 = I2(o->edge_tri, 2, 2, edge);
    e->nl1 = o->nlevels[e->el1 - 1] - 1; e->nu1 = o->ulevels[e->el1 - 1];
    e->dX1 = I2(o->edge_cross_dxdy, 4, 1, edge); e->dY1 = I2(o->edge_cross_dxdy, 4, 2, edge);
    e->a = r_earth * o->elem_cos[e->el1 - 1];
    e->nl2 = 0; e->nu2 = 0; e->dX2 = 0.0; e->dY2 = 0.0;
    if (e->el2 > 0) {
        e->dX2 = I2(o->edge_cross_dxdy, 4, 3, edge); e->dY2 = I2(o->edge_cross_dxdy, 4, 4, edge);
        e->nl2 = o->nlevels[e->el2 - 1] - 1; e->nu2 = o->ulevels[e->el2 - 1];
        e->a = 0.5 * (e->a + r_earth * o->elem_cos[e->el2 - 1]);
    }
    e->nl12 = imin(e->nl1, e->nl2); e->nu12 = imax(e->nu1, e->nu2);
}
#define VF1(vel, nz) ((-V3(vel, 2, nz, e.el1) * e.dX1 + V3(vel, 1, nz, e.el1) * e.dY1) * T2(o->helem, nz, e.el1))
#define VF2(vel, nz) ((V3(vel, 2, nz, e.el2) * e.dX2 - V3(vel, 1, nz, e.el2) * e.dY2) * T2(o->helem, nz, e.el2))

/* oce_adv_tra_hor.F90:57-211 */
void o_adv_tra_hor_upw1(oracle_t* o, const double* ttf, const double* vel, double* flux, int init_zero) {
    if (init_zero) memset(flux, 0, sizeof(double) * (size_t)LM * (size_t)o->myDim_edge2D);
    for (int edge = 1; edge <= o->myDim_edge2D; edge++) {
        edge_t e; edge_setup(o, edge, &e);
        double vflux;
#define UPW(nz) T2(flux, nz, edge) = -0.5 * (T2(ttf, nz, e.en1) * (vflux + fabs(vflux)) + T2(ttf, nz, e.en2) * (vflux - fabs(vflux))) - T2(flux, nz, edge)
        for (int nz = e.nu1; nz <= e.nu12 - 1; nz++) { vflux = VF1(vel, nz); UPW(nz); }                         /* (A) */
        if (e.nu2 > 0) for (int nz = e.nu2; nz <= e.nu12 - 1; nz++) { vflux = VF2(vel, nz); UPW(nz); }           /* (B) */
        for (int nz = e.nu12; nz <= e.nl12; nz++) { vflux = VF1(vel, nz) + VF2(vel, nz); UPW(nz); }              /* (C) */
        for (int nz = e.nl12 + 1; nz <= e.nl1; nz++) { vflux = VF1(vel, nz); UPW(nz); }                          /* (D) */
        for (int nz = e.nl12 + 1; nz <= e.nl2; nz++) { vflux = VF2(vel, nz); UPW(nz); }                          /* (E) */
#undef UPW
    }
}

/* oce_adv_tra_hor.F90:215-481 (MUSCL, mfct=0) and :485-733 (MFCT, mfct=1: no c_lo factor) */
void o_adv_tra_hor_muscl(oracle_t* o, const double* ttf, const double* vel, double num_ord, double* flux, int init_zero, int mfct) {
    if (init_zero) memset(flux, 0, sizeof(double) * (size_t)LM * (size_t)o->myDim_edge2D);
    const double* g = o->edge_up_dn_grad;
    for (int edge = 1; edge <= o->myDim_edge2D; edge++) {
        edge_t e; edge_setup(o, edge, &e);
        double dx = I2(o->edge_dxdy, 2, 1, edge), dy = I2(o->edge_dxdy, 2, 2, edge);
        double vflux, Tmean1, Tmean2, cHO;
#define RECON(nz)                                                                                                        \
        if (mfct) {                                                                                                      \
            Tmean2 = T2(ttf, nz, e.en2) - (2.0 * (T2(ttf, nz, e.en2) - T2(ttf, nz, e.en1)) + dx * e.a * G4(g, 2, nz, edge) + dy * r_earth * G4(g, 4, nz, edge)) / 6.0; \
            Tmean1 = T2(ttf, nz, e.en1) + (2.0 * (T2(ttf, nz, e.en2) - T2(ttf, nz, e.en1)) + dx * e.a * G4(g, 1, nz, edge) + dy * r_earth * G4(g, 3, nz, edge)) / 6.0; \
        } else {                                                                                                         \
            double c_lo1 = (double)imax((o->nboundary_lay[e.en1 - 1] - (nz)) >= 0 ? 1 : -1, 0);                          \
            double c_lo2 = (double)imax((o->nboundary_lay[e.en2 - 1] - (nz)) >= 0 ? 1 : -1, 0);                          \
            Tmean2 = T2(ttf, nz, e.en2) - (2.0 * (T2(ttf, nz, e.en2) - T2(ttf, nz, e.en1)) + dx * e.a * G4(g, 2, nz, edge) + dy * r_earth * G4(g, 4, nz, edge)) / 6.0 * c_lo2; \
            Tmean1 = T2(ttf, nz, e.en1) + (2.0 * (T2(ttf, nz, e.en2) - T2(ttf, nz, e.en1)) + dx * e.a * G4(g, 1, nz, edge) + dy * r_earth * G4(g, 3, nz, edge)) / 6.0 * c_lo1; \
        }
#define HOFLUX(nz)                                                                                       \
        cHO = (vflux + fabs(vflux)) * Tmean1 + (vflux - fabs(vflux)) * Tmean2;                           \
        T2(flux, nz, edge) = -0.5 * (1.0 - num_ord) * cHO - vflux * num_ord * (0.5 * (Tmean1 + Tmean2)) - T2(flux, nz, edge)
        for (int nz = e.nu1; nz <= e.nu12 - 1; nz++) { RECON(nz) vflux = VF1(vel, nz); HOFLUX(nz); }
        if (e.nu2 > 0) for (int nz = e.nu2; nz <= e.nu12 - 1; nz++) { RECON(nz) vflux = VF2(vel, nz); HOFLUX(nz); }
        for (int nz = e.nu12; nz <= e.nl12; nz++) { RECON(nz) vflux = VF1(vel, nz) + VF2(vel, nz); HOFLUX(nz); }
        for (int nz = e.nl12 + 1; nz <= e.nl1; nz++) { RECON(nz) vflux = VF1(vel, nz); HOFLUX(nz); }
        for (int nz = e.nl12 + 1; nz <= e.nl2; nz++) { RECON(nz) vflux = VF2(vel, nz); HOFLUX(nz); }
#undef RECON
#undef HOFLUX
    }
}

/* oce_adv_tra_ver.F90:231-282 */
void o_adv_tra_ver_upw1(oracle_t* o, const double* ttf, const double* W, double* flux, int init_zero) {
    if (init_zero) memset(flux, 0, sizeof(double) * (size_t)NL * (size_t)o->myDim_nod2D);
    for (int n = 1; n <= o->myDim_nod2D; n++) {
        int nzmax = o->nlevels_nod2D[n - 1], nzmin = o->ulevels_nod2D[n - 1];
        int nz = nzmin;
        W2(flux, nz, n) = -W2(W, nz, n) * T2(ttf, nz, n) * W2(o->area, nz, n) - W2(flux, nz, n);
        nz = nzmax;
        W2(flux, nz, n) = 0.0 - W2(flux, nz, n);
        for (nz = nzmin + 1; nz <= nzmax - 1; nz++)
            W2(flux, nz, n) = -0.5 * (T2(ttf, nz, n) * (W2(W, nz, n) + fabs(W2(W, nz, n))) +
                                      T2(ttf, nz - 1, n) * (W2(W, nz, n) - fabs(W2(W, nz, n)))) * W2(o->area, nz, n) - W2(flux, nz, n);
    }
}

/* oce_adv_tra_ver.F90:286-357 */
void o_adv_tra_ver_qr4c(oracle_t* o, const double* ttf, const double* W, double num_ord, double* flux, int init_zero) {
    if (init_zero) memset(flux, 0, sizeof(double) * (size_t)NL * (size_t)o->myDim_nod2D);
    const double* Z = o->Z_3d_n; const double* zb = o->zbar_3d_n;
    for (int n = 1; n <= o->myDim_nod2D; n++) {
        int nzmax = o->nlevels_nod2D[n - 1], nzmin = o->ulevels_nod2D[n - 1];
        int nz = nzmin;
        W2(flux, nz, n) = -T2(ttf, nz, n) * W2(W, nz, n) * W2(o->area, nz, n) - W2(flux, nz, n);
        nz = nzmin + 1;
        W2(flux, nz, n) = -0.5 * (T2(ttf, nz - 1, n) + T2(ttf, nz, n)) * W2(W, nz, n) * W2(o->area, nz, n) - W2(flux, nz, n);
        nz = nzmax - 1;
        W2(flux, nz, n) = -0.5 * (T2(ttf, nz - 1, n) + T2(ttf, nz, n)) * W2(W, nz, n) * W2(o->area, nz, n) - W2(flux, nz, n);
        nz = nzmax;
        W2(flux, nz, n) = 0.0 - W2(flux, nz, n);
        for (nz = nzmin + 2; nz <= nzmax - 2; nz++) {
            double qc = (T2(ttf, nz - 1, n) - T2(ttf, nz, n)) / (T2(Z, nz - 1, n) - T2(Z, nz, n));
            double qu = (T2(ttf, nz, n) - T2(ttf, nz + 1, n)) / (T2(Z, nz, n) - T2(Z, nz + 1, n));
            double qd = (T2(ttf, nz - 2, n) - T2(ttf, nz - 1, n)) / (T2(Z, nz - 2, n) - T2(Z, nz - 1, n));
            double Tmean1 = T2(ttf, nz, n) + (2 * qc + qu) * (W2(zb, nz, n) - T2(Z, nz, n)) / 3.0;
            double Tmean2 = T2(ttf, nz - 1, n) + (2 * qc + qd) * (W2(zb, nz, n) - T2(Z, nz - 1, n)) / 3.0;
            double Tmean = (W2(W, nz, n) + fabs(W2(W, nz, n))) * Tmean1 + (W2(W, nz, n) - fabs(W2(W, nz, n))) * Tmean2;
            W2(flux, nz, n) = (-0.5 * (1.0 - num_ord) * Tmean - num_ord * (0.5 * (Tmean1 + Tmean2)) * W2(W, nz, n)) * W2(o->area, nz, n) - W2(flux, nz, n);
        }
    }
}

static inline double fsign(double a, double b) { return (b >= 0.0 && !(b == 0.0 && signbit(b))) ? fabs(a) : -fabs(a); }   /* Fortran SIGN */

/* oce_adv_tra_ver.F90:361-538 */
void o_adv_tra_vert_ppm(oracle_t* o, const double* ttf, const double* W, double* flux, int init_zero) {
    if (init_zero) memset(flux, 0, sizeof(double) * (size_t)NL * (size_t)o->myDim_nod2D);
    double* tv = (double*)calloc((size_t)NL + 2, sizeof(double));
    double* tvert = (double*)calloc((size_t)NL + 2, sizeof(double));
    const double* hn = o->hnode_new;
    const double dt = o->dt;
    for (int n = 1; n <= o->myDim_nod2D; n++) {
        int nzmax = o->nlevels_nod2D[n - 1], nzmin = o->ulevels_nod2D[n - 1];
#define TV(k) tv[(k)]
#define TT(k) T2(ttf, k, n)
        TV(nzmin) = TT(nzmin);
        TV(nzmin + 1) = 0.5 * (TT(nzmin) + TT(nzmin + 1));
        TV(nzmax - 1) = -TT(nzmax - 2) * dmin(fsign(1.0, W2(W, nzmax - 1, n)), 0.0) + TT(nzmax - 1) * dmax(fsign(1.0, W2(W, nzmax - 1, n)), 0.0);
        TV(nzmax) = TT(nzmax - 1);
        for (int nz = nzmin + 1; nz <= nzmax - 3; nz++) {
            double dzjm1 = T2(hn, nz - 1, n), dzj = T2(hn, nz, n), dzjp1 = T2(hn, nz + 1, n), dzjp2 = T2(hn, nz + 2, n);
            double deltaj = dzj / (dzjm1 + dzj + dzjp1) *
                            ((2.0 * dzjm1 + dzj) / (dzjp1 + dzj) * (TT(nz + 1) - TT(nz)) + (dzj + 2.0 * dzjp1) / (dzjm1 + dzj) * (TT(nz) - TT(nz - 1)));
            double deltajp1 = dzjp1 / (dzj + dzjp1 + dzjp2) *
                              ((2.0 * dzj + dzjp1) / (dzjp2 + dzjp1) * (TT(nz + 2) - TT(nz + 1)) + (dzjp1 + 2.0 * dzjp2) / (dzj + dzjp1) * (TT(nz + 1) - TT(nz)));
            if ((TT(nz + 1) - TT(nz)) * (TT(nz) - TT(nz - 1)) > 0.0)
                deltaj = dmin(dmin(fabs(deltaj), 2.0 * fabs(TT(nz + 1) - TT(nz))), 2.0 * fabs(TT(nz) - TT(nz - 1))) * fsign(1.0, deltaj);
            else deltaj = 0.0;
            if ((TT(nz + 2) - TT(nz + 1)) * (TT(nz + 1) - TT(nz)) > 0.0)
                deltajp1 = dmin(dmin(fabs(deltajp1), 2.0 * fabs(TT(nz + 2) - TT(nz + 1))), 2.0 * fabs(TT(nz + 1) - TT(nz))) * fsign(1.0, deltajp1);
            else deltajp1 = 0.0;
            TV(nz + 1) = TT(nz) + dzj / (dzj + dzjp1) * (TT(nz + 1) - TT(nz)) +
                         1.0 / (dzjm1 + dzj + dzjp1 + dzjp2) *
                             ((2.0 * dzjp1 * dzj) / (dzj + dzjp1) * ((dzjm1 + dzj) / (2.0 * dzj + dzjp1) - (dzjp2 + dzjp1) / (2.0 * dzjp1 + dzj)) * (TT(nz + 1) - TT(nz)) -
                              dzj * (dzjm1 + dzj) / (2.0 * dzj + dzjp1) * deltajp1 + dzjp1 * (dzjp1 + dzjp2) / (dzj + 2.0 * dzjp1) * deltaj);
        }
        for (int k = 1; k <= nzmax; k++) tvert[k] = 0.0;
        for (int nz = nzmin; nz <= nzmax - 1; nz++) {
            if (W2(W, nz, n) <= 0.0 && W2(W, nz + 1, n) >= 0.0) continue;
            double aL = TV(nz), aR = TV(nz + 1);
            if ((aR - TT(nz)) * (TT(nz) - aL) <= 0.0) { aL = TT(nz); aR = TT(nz); }
            if ((aR - aL) * (TT(nz) - 0.5 * (aL + aR)) > (aR - aL) * (aR - aL) / 6.0) aL = 3.0 * TT(nz) - 2.0 * aR;
            if ((aR - aL) * (TT(nz) - 0.5 * (aR + aL)) < -((aR - aL) * (aR - aL)) / 6.0) aR = 3.0 * TT(nz) - 2.0 * aL;
            double dzj = T2(o->hnode, nz, n);
            double aj = 6.0 * (TT(nz) - 0.5 * (aL + aR));
            if (W2(W, nz, n) > 0.0) {
                double x = dmin(W2(W, nz, n) * dt / dzj, 1.0);
                tvert[nz] = (-aL - 0.5 * x * (aR - aL + (1.0 - 2.0 / 3.0 * x) * aj));
                tvert[nz] = tvert[nz] * W2(o->area, nz, n) * W2(W, nz, n);
            }
            if (W2(W, nz + 1, n) < 0.0) {
                double x = dmin(-W2(W, nz + 1, n) * dt / dzj, 1.0);
                tvert[nz + 1] = (-aR + 0.5 * x * (aR - aL - (1.0 - 2.0 / 3.0 * x) * aj));
                tvert[nz + 1] = tvert[nz + 1] * W2(o->area, nz + 1, n) * W2(W, nz + 1, n);
            }
        }
        tvert[nzmin] = -TV(nzmin) * W2(W, nzmin, n) * W2(o->area, nzmin, n);
        tvert[nzmax] = 0.0;
        for (int nz = nzmin; nz <= nzmax; nz++) W2(flux, nz, n) = tvert[nz] - W2(flux, nz, n);
#undef TV
#undef TT
    }
    free(tv); free(tvert);
}

/* oce_adv_tra_ver.F90:542-590 */
void o_adv_tra_ver_cdiff(oracle_t* o, const double* ttf, const double* W, double* flux, int init_zero) {
    if (init_zero) memset(flux, 0, sizeof(double) * (size_t)NL * (size_t)o->myDim_nod2D);
    double* tvert = (double*)calloc((size_t)NL + 2, sizeof(double));
    for (int n = 1; n <= o->myDim_nod2D; n++) {
        int nzmax = o->nlevels_nod2D[n - 1] - 1, nzmin = o->ulevels_nod2D[n - 1];
        tvert[nzmin] = -W2(W, nzmin, n) * T2(ttf, nzmin, n) * W2(o->area, nzmin, n);
        tvert[nzmax + 1] = 0.0;
        for (int nz = nzmin + 1; nz <= nzmax; nz++) {
            double tv = 0.5 * (T2(ttf, nz - 1, n) + T2(ttf, nz, n));
            tvert[nz] = -tv * W2(W, nz, n) * W2(o->area, nz, n);
        }
        for (int nz = nzmin; nz <= nzmax; nz++) W2(flux, nz, n) = tvert[nz] - W2(flux, nz, n);   /* note: flux(nzmax+1) untouched, as in the reference */
    }
    free(tvert);
}

/* Thomas algorithm shared by adv_tra_vert_impl and diff_ver_part_impl_ale (oce_ale_tracer.F90:828-844) */
static void thomas(int nzmin, int nzmax, const double* a, const double* b, const double* c, double* tr, double* cp, double* tp) {
    cp[nzmin] = c[nzmin] / b[nzmin];
    tp[nzmin] = tr[nzmin] / b[nzmin];
    for (int nz = nzmin + 1; nz <= nzmax - 1; nz++) {
        double m = b[nz] - cp[nz - 1] * a[nz];
        cp[nz] = c[nz] / m;
        tp[nz] = (tr[nz] - tp[nz - 1] * a[nz]) / m;
    }
    tr[nzmax - 1] = tp[nzmax - 1];
    for (int nz = nzmax - 2; nz >= nzmin; nz--) tr[nz] = tp[nz] - cp[nz] * tr[nz + 1];
}

/* oce_adv_tra_ver.F90:83-227 */
void o_adv_tra_vert_impl(oracle_t* o, double* ttf, const double* W) {
    size_t sz = (size_t)NL + 2;
    double* buf = (double*)calloc(6 * sz, sizeof(double));
    double *a = buf, *b = buf + sz, *c = buf + 2 * sz, *tr = buf + 3 * sz, *cp = buf + 4 * sz, *tp = buf + 5 * sz;
    const double dt = o->dt;
    for (int n = 1; n <= o->myDim_nod2D; n++) {
        memset(buf, 0, 6 * sz * sizeof(double));
        int nzmax = o->nlevels_nod2D[n - 1], nzmin = o->ulevels_nod2D[n - 1];
        int nz = nzmin;
        double zinv = 1.0 * dt, v_adv;
        a[nz] = 0.0;
        v_adv = zinv * W2(o->area, nz, n) / W2(o->areasvol, nz, n);
        b[nz] = T2(o->hnode_new, nz, n) + W2(W, nz, n) * v_adv;
        v_adv = zinv * W2(o->area, nz + 1, n) / W2(o->areasvol, nz, n);
        b[nz] = b[nz] - dmin(0.0, W2(W, nz + 1, n)) * v_adv;
        c[nz] = -dmax(0.0, W2(W, nz + 1, n)) * v_adv;
        for (nz = nzmin + 1; nz <= nzmax - 2; nz++) {
            v_adv = zinv * W2(o->area, nz, n) / W2(o->areasvol, nz, n);
            a[nz] = dmin(0.0, W2(W, nz, n)) * v_adv;
            b[nz] = T2(o->hnode_new, nz, n) + dmax(0.0, W2(W, nz, n)) * v_adv;
            v_adv = zinv * W2(o->area, nz + 1, n) / W2(o->areasvol, nz, n);
            b[nz] = b[nz] - dmin(0.0, W2(W, nz + 1, n)) * v_adv;
            c[nz] = -dmax(0.0, W2(W, nz + 1, n)) * v_adv;
        }
        nz = nzmax - 1;
        v_adv = zinv * W2(o->area, nz, n) / W2(o->areasvol, nz, n);
        a[nz] = dmin(0.0, W2(W, nz, n)) * v_adv;
        b[nz] = T2(o->hnode_new, nz, n) + dmax(0.0, W2(W, nz, n)) * v_adv;
        c[nz] = 0.0;
        nz = nzmin;
        double dz = T2(o->hnode_new, nz, n);
        tr[nz] = -(b[nz] - dz) * T2(ttf, nz, n) - c[nz] * T2(ttf, nz + 1, n);
        for (nz = nzmin + 1; nz <= nzmax - 2; nz++) {
            dz = T2(o->hnode_new, nz, n);
            tr[nz] = -a[nz] * T2(ttf, nz - 1, n) - (b[nz] - dz) * T2(ttf, nz, n) - c[nz] * T2(ttf, nz + 1, n);
        }
        nz = nzmax - 1;
        dz = T2(o->hnode_new, nz, n);
        tr[nz] = -a[nz] * T2(ttf, nz - 1, n) - (b[nz] - dz) * T2(ttf, nz, n);
        thomas(nzmin, nzmax, a, b, c, tr, cp, tp);
        for (nz = nzmin; nz <= nzmax - 1; nz++) T2(ttf, nz, n) = T2(ttf, nz, n) + tr[nz];
    }
    free(buf);
}

/* oce_adv_tra_fct.F90:58-276: a1, a2, a3 (vlimit=1), b1, b2 — everything before exchange_nod(fct_plus,fct_minus) */
void o_oce_tra_adv_fct_a(oracle_t* o, const double* ttf, const double* LO, double* adf_h, double* adf_v) {
    const double flux_eps = 1e-16, bignumber = 1e3;
    const double dt = o->dt;
    int N = o->myDim_nod2D + o->eDim_nod2D;
    double *fmax = o->fct_ttf_max, *fmin = o->fct_ttf_min, *fplus = o->fct_plus, *fminus = o->fct_minus, *aux = o->aux_minmax;
    double* tvert_max = (double*)calloc((size_t)NL + 2, sizeof(double));
    double* tvert_min = (double*)calloc((size_t)NL + 2, sizeof(double));
    for (int n = 1; n <= N; n++)                                                               /* a1 */
        for (int nz = o->ulevels_nod2D[n - 1]; nz <= o->nlevels_nod2D[n - 1] - 1; nz++) {
            T2(fmax, nz, n) = dmax(T2(LO, nz, n), T2(ttf, nz, n));
            T2(fmin, nz, n) = dmin(T2(LO, nz, n), T2(ttf, nz, n));
        }
    for (int elem = 1; elem <= o->myDim_elem2D; elem++) {                                      /* a2 */
        const int* en = &I2(o->elem2D_nodes, 3, 1, elem);
        int nu1 = o->ulevels[elem - 1], nl1 = o->nlevels[elem - 1];
        if (o->fct_cavity_neutral)                                                             /* see oracle.h */
            for (int nz = 1; nz <= nu1 - 1; nz++) { V3(aux, 1, nz, elem) = -bignumber; V3(aux, 2, nz, elem) = bignumber; }
        for (int nz = nu1; nz <= nl1 - 1; nz++) {
            V3(aux, 1, nz, elem) = dmax(dmax(T2(fmax, nz, en[0]), T2(fmax, nz, en[1])), T2(fmax, nz, en[2]));
            V3(aux, 2, nz, elem) = dmin(dmin(T2(fmin, nz, en[0]), T2(fmin, nz, en[1])), T2(fmin, nz, en[2]));
        }
        if (nl1 <= NL - 1)
            for (int nz = nl1; nz <= NL - 1; nz++) { V3(aux, 1, nz, elem) = -bignumber; V3(aux, 2, nz, elem) = bignumber; }
    }
    for (int n = 1; n <= o->myDim_nod2D; n++) {                                                /* a3, vlimit = 1 */
        int nu1 = o->ulevels_nod2D[n - 1], nl1 = o->nlevels_nod2D[n - 1];
        for (int nz = nu1; nz <= nl1 - 1; nz++) {
            double mx = -1e308, mn = 1e308;
            for (int k = 1; k <= o->nod_in_elem2D_num[n - 1]; k++) {
                int elem = I2(o->nod_in_elem2D, o->nod_in_elem2D_max, k, n);
                mx = dmax(mx, V3(aux, 1, nz, elem)); mn = dmin(mn, V3(aux, 2, nz, elem));
            }
            tvert_max[nz] = mx; tvert_min[nz] = mn;
        }
        T2(fmax, nu1, n) = tvert_max[nu1] - T2(LO, nu1, n);
        T2(fmin, nu1, n) = tvert_min[nu1] - T2(LO, nu1, n);
        for (int nz = nu1 + 1; nz <= nl1 - 2; nz++) {
            T2(fmax, nz, n) = dmax(dmax(tvert_max[nz - 1], tvert_max[nz]), tvert_max[nz + 1]) - T2(LO, nz, n);
            T2(fmin, nz, n) = dmin(dmin(tvert_min[nz - 1], tvert_min[nz]), tvert_min[nz + 1]) - T2(LO, nz, n);
        }
        int nz = nl1 - 1;
        T2(fmax, nz, n) = tvert_max[nz] - T2(LO, nz, n);
        T2(fmin, nz, n) = tvert_min[nz] - T2(LO, nz, n);
    }
    for (int n = 1; n <= o->myDim_nod2D; n++)                                                  /* b1 */
        for (int nz = o->ulevels_nod2D[n - 1]; nz <= o->nlevels_nod2D[n - 1] - 1; nz++) { T2(fplus, nz, n) = 0.0; T2(fminus, nz, n) = 0.0; }
    for (int n = 1; n <= o->myDim_nod2D; n++)
        for (int nz = o->ulevels_nod2D[n - 1]; nz <= o->nlevels_nod2D[n - 1] - 1; nz++) {
            T2(fplus, nz, n) = T2(fplus, nz, n) + (dmax(0.0, W2(adf_v, nz, n)) + dmax(0.0, -W2(adf_v, nz + 1, n)));
            T2(fminus, nz, n) = T2(fminus, nz, n) + (dmin(0.0, W2(adf_v, nz, n)) + dmin(0.0, -W2(adf_v, nz + 1, n)));
        }
    for (int edge = 1; edge <= o->myDim_edge2D; edge++) {
        int en1 = I2(o->edges, 2, 1, edge), en2 = I2(o->edges, 2, 2, edge);
        int el1 = I2(o->edge_tri, 2, 1, edge), el2 = I2(o->edge_tri, 2, 2, edge);
        int nl1 = o->nlevels[el1 - 1] - 1, nu1 = o->ulevels[el1 - 1], nl2 = 0, nu2 = 0;
        if (el2 > 0) { nl2 = o->nlevels[el2 - 1] - 1; nu2 = o->ulevels[el2 - 1]; }
        int nl12 = imax(nl1, nl2), nu12 = nu1;
        if (nu2 > 0) nu12 = imin(nu1, nu2);
        for (int nz = nu12; nz <= nl12; nz++) {
            T2(fplus, nz, en1) = T2(fplus, nz, en1) + dmax(0.0, T2(adf_h, nz, edge));
            T2(fminus, nz, en1) = T2(fminus, nz, en1) + dmin(0.0, T2(adf_h, nz, edge));
            T2(fplus, nz, en2) = T2(fplus, nz, en2) + dmax(0.0, -T2(adf_h, nz, edge));
            T2(fminus, nz, en2) = T2(fminus, nz, en2) + dmin(0.0, -T2(adf_h, nz, edge));
        }
    }
    for (int n = 1; n <= o->myDim_nod2D; n++)                                                  /* b2 */
        for (int nz = o->ulevels_nod2D[n - 1]; nz <= o->nlevels_nod2D[n - 1] - 1; nz++) {
            double flux = T2(fplus, nz, n) * dt / W2(o->areasvol, nz, n) + flux_eps;
            T2(fplus, nz, n) = dmin(1.0, T2(fmax, nz, n) / flux);
            flux = T2(fminus, nz, n) * dt / W2(o->areasvol, nz, n) - flux_eps;
            T2(fminus, nz, n) = dmin(1.0, T2(fmin, nz, n) / flux);
        }
    free(tvert_max); free(tvert_min);
}

/* oce_adv_tra_fct.F90:282-348: b3 — after the exchange */
void o_oce_tra_adv_fct_b(oracle_t* o, double* adf_h, double* adf_v) {
    double *fplus = o->fct_plus, *fminus = o->fct_minus;
    for (int n = 1; n <= o->myDim_nod2D; n++) {
        int nu1 = o->ulevels_nod2D[n - 1], nl1 = o->nlevels_nod2D[n - 1];
        int nz = nu1;
        double ae = 1.0, flux = W2(adf_v, nz, n);
        if (flux >= 0.0) ae = dmin(ae, T2(fplus, nz, n)); else ae = dmin(ae, T2(fminus, nz, n));
        W2(adf_v, nz, n) = ae * W2(adf_v, nz, n);
        for (nz = nu1 + 1; nz <= nl1 - 1; nz++) {
            ae = 1.0; flux = W2(adf_v, nz, n);
            if (flux >= 0.0) { ae = dmin(ae, T2(fminus, nz - 1, n)); ae = dmin(ae, T2(fplus, nz, n)); }
            else { ae = dmin(ae, T2(fplus, nz - 1, n)); ae = dmin(ae, T2(fminus, nz, n)); }
            W2(adf_v, nz, n) = ae * W2(adf_v, nz, n);
        }
    }
    for (int edge = 1; edge <= o->myDim_edge2D; edge++) {
        int en1 = I2(o->edges, 2, 1, edge), en2 = I2(o->edges, 2, 2, edge);
        int el1 = I2(o->edge_tri, 2, 1, edge), el2 = I2(o->edge_tri, 2, 2, edge);
        int nu1 = o->ulevels[el1 - 1], nl1 = o->nlevels[el1 - 1] - 1, nl2 = 0, nu2 = 0;
        if (el2 > 0) { nu2 = o->ulevels[el2 - 1]; nl2 = o->nlevels[el2 - 1] - 1; }
        int nl12 = imax(nl1, nl2), nu12 = nu1;
        if (nu2 > 0) nu12 = imin(nu1, nu2);
        for (int nz = nu12; nz <= nl12; nz++) {
            double ae = 1.0, flux = T2(adf_h, nz, edge);
            if (flux >= 0.0) { ae = dmin(ae, T2(fplus, nz, en1)); ae = dmin(ae, T2(fminus, nz, en2)); }
            else { ae = dmin(ae, T2(fminus, nz, en1)); ae = dmin(ae, T2(fplus, nz, en2)); }
            T2(adf_h, nz, edge) = ae * T2(adf_h, nz, edge);
        }
    }
}

/* oce_adv_tra_driver.F90:201-269 */
void o_oce_tra_adv_flux2dtracer(oracle_t* o, double* dttf_h, double* dttf_v, double* flux_h, double* flux_v,
                                int use_lo, const double* ttf, const double* lo) {
    const double dt = o->dt;
    if (use_lo)
        for (int n = 1; n <= o->myDim_nod2D; n++)
            for (int nz = o->ulevels_nod2D[n - 1]; nz <= o->nlevels_nod2D[n - 1] - 1; nz++)
                T2(dttf_v, nz, n) = T2(dttf_v, nz, n) - T2(ttf, nz, n) * T2(o->hnode, nz, n) + T2(lo, nz, n) * T2(o->hnode_new, nz, n);
    for (int n = 1; n <= o->myDim_nod2D; n++)
        for (int nz = o->ulevels_nod2D[n - 1]; nz <= o->nlevels_nod2D[n - 1] - 1; nz++)
            T2(dttf_v, nz, n) = T2(dttf_v, nz, n) + (W2(flux_v, nz, n) - W2(flux_v, nz + 1, n)) * dt / W2(o->areasvol, nz, n);
    for (int edge = 1; edge <= o->myDim_edge2D; edge++) {
        int en1 = I2(o->edges, 2, 1, edge), en2 = I2(o->edges, 2, 2, edge);
        int el1 = I2(o->edge_tri, 2, 1, edge), el2 = I2(o->edge_tri, 2, 2, edge);
        int nl1 = o->nlevels[el1 - 1] - 1, nu1 = o->ulevels[el1 - 1], nl2 = 0, nu2 = 0;
        if (el2 > 0) { nl2 = o->nlevels[el2 - 1] - 1; nu2 = o->ulevels[el2 - 1]; }
        int nl12 = imax(nl1, nl2), nu12 = nu1;
        if (nu2 > 0) nu12 = imin(nu1, nu2);
        for (int nz = nu12; nz <= nl12; nz++) {
            T2(dttf_h, nz, en1) = T2(dttf_h, nz, en1) + T2(flux_h, nz, edge) * dt / W2(o->areasvol, nz, en1);
            T2(dttf_h, nz, en2) = T2(dttf_h, nz, en2) - T2(flux_h, nz, edge) * dt / W2(o->areasvol, nz, en2);
        }
    }
}

/* oce_ale_tracer.F90:934-1077 */
void o_diff_part_hor_redi(oracle_t* o) {
    const double dt = o->dt;
    double isredi = o->Redi ? 1.0 : 0.0;
    double* rhs1 = (double*)calloc((size_t)NL + 2, sizeof(double));
    double* rhs2 = (double*)calloc((size_t)NL + 2, sizeof(double));
    const double *Ki = o->Ki, *tr_z = o->tr_z, *sl = o->slope_tapered, *tr_xy = o->tr_xy;
    for (int edge = 1; edge <= o->myDim_edge2D; edge++) {
        for (int k = 0; k <= NL; k++) { rhs1[k] = 0.0; rhs2[k] = 0.0; }
        double dX1 = I2(o->edge_cross_dxdy, 4, 1, edge), dY1 = I2(o->edge_cross_dxdy, 4, 2, edge), dX2 = 0.0, dY2 = 0.0;
        int el1 = I2(o->edge_tri, 2, 1, edge), el2 = I2(o->edge_tri, 2, 2, edge);
        int en1 = I2(o->edges, 2, 1, edge), en2 = I2(o->edges, 2, 2, edge);
        int nl1 = o->nlevels[el1 - 1] - 1, ul1 = o->ulevels[el1 - 1], nl2 = 0, ul2 = 0;
        if (el2 > 0) { nl2 = o->nlevels[el2 - 1] - 1; ul2 = o->ulevels[el2 - 1]; dX2 = I2(o->edge_cross_dxdy, 4, 3, edge); dY2 = I2(o->edge_cross_dxdy, 4, 4, edge); }
        int nl12 = imin(nl1, nl2), ul12 = imax(ul1, ul2);
        double Kh, dz, Tz1, Tz2, SxTz, SyTz, Tx, Ty, Fx, Fy, c;
#define COMMON(nz)                                                                                       \
        Kh = (T2(Ki, nz, en1) + T2(Ki, nz, en2)) / 2.0;                                                   \
        Tz1 = 0.5 * (W2(tr_z, nz, en1) + W2(tr_z, nz + 1, en1));                                          \
        Tz2 = 0.5 * (W2(tr_z, nz, en2) + W2(tr_z, nz + 1, en2));                                          \
        SxTz = (Tz1 * S3(sl, 1, nz, en1) + Tz2 * S3(sl, 1, nz, en2)) / 2.0;                               \
        SyTz = (Tz1 * S3(sl, 2, nz, en1) + Tz2 * S3(sl, 2, nz, en2)) / 2.0
#define FLUX()  Fx = Kh * (Tx + SxTz * isredi); Fy = Kh * (Ty + SyTz * isredi)
#define ADD(nz) rhs1[nz] = rhs1[nz] + c; rhs2[nz] = rhs2[nz] - c
        for (int nz = ul1; nz <= ul12 - 1; nz++) {                                               /* (A) */
            COMMON(nz); dz = T2(o->helem, nz, el1); Tx = V3(tr_xy, 1, nz, el1); Ty = V3(tr_xy, 2, nz, el1); FLUX();
            c = (-dX1 * Fy + dY1 * Fx) * dz; ADD(nz);
        }
        if (ul2 > 0) for (int nz = ul2; nz <= ul12 - 1; nz++) {                                  /* (B) */
            COMMON(nz); dz = T2(o->helem, nz, el2); Tx = V3(tr_xy, 1, nz, el2); Ty = V3(tr_xy, 2, nz, el2); FLUX();
            c = (dX2 * Fy - dY2 * Fx) * dz; ADD(nz);
        }
        for (int nz = ul12; nz <= nl12; nz++) {                                                  /* (C) */
            COMMON(nz); dz = (T2(o->helem, nz, el1) + T2(o->helem, nz, el2)) / 2.0;
            Tx = 0.5 * (V3(tr_xy, 1, nz, el1) + V3(tr_xy, 1, nz, el2)); Ty = 0.5 * (V3(tr_xy, 2, nz, el1) + V3(tr_xy, 2, nz, el2)); FLUX();
            c = ((dX2 - dX1) * Fy - (dY2 - dY1) * Fx) * dz; ADD(nz);
        }
        for (int nz = nl12 + 1; nz <= nl1; nz++) {                                               /* (D) */
            COMMON(nz); dz = T2(o->helem, nz, el1); Tx = V3(tr_xy, 1, nz, el1); Ty = V3(tr_xy, 2, nz, el1); FLUX();
            c = (-dX1 * Fy + dY1 * Fx) * dz; ADD(nz);
        }
        for (int nz = nl12 + 1; nz <= nl2; nz++) {                                               /* (E) */
            COMMON(nz); dz = T2(o->helem, nz, el2); Tx = V3(tr_xy, 1, nz, el2); Ty = V3(tr_xy, 2, nz, el2); FLUX();
            c = (dX2 * Fy - dY2 * Fx) * dz; ADD(nz);
        }
#undef COMMON
#undef FLUX
#undef ADD
        nl12 = imax(nl1, nl2); ul12 = ul1;
        if (ul2 > 0) ul12 = imin(ul1, ul2);
        for (int nz = ul12; nz <= nl12; nz++) {
            T2(o->del_ttf, nz, en1) = T2(o->del_ttf, nz, en1) + rhs1[nz] * dt / W2(o->areasvol, nz, en1);
            T2(o->del_ttf, nz, en2) = T2(o->del_ttf, nz, en2) + rhs2[nz] * dt / W2(o->areasvol, nz, en2);
        }
    }
    free(rhs1); free(rhs2);
}

/* diff_ver_part_redi_expl, oce_ale_tracer.F90:860-930 — "a projection of horizontal Redi diffusivity onto vertical"
 * (call site :288, under `if (Redi)`, after diff_ver_part_expl_ale and before the T* update).  tr_xy holds the gradients of
 * tr_arr (the second tracer_gradient_elements call of init_tracers_AB, oce_tracer_mod.F90:79).  tr_xynodes is a local of
 * the reference (own nodes only: "the halo is not needed", :895); zbar_n, Z_n are the module's column scratch. */
void o_diff_ver_part_redi_expl(oracle_t* o) {
    const double dt = o->dt;
    const int myN = o->myDim_nod2D;
    const double *Ki = o->Ki, *sl = o->slope_tapered, *hn = o->hnode_new, *tr_xy = o->tr_xy;
    double* tr_xynodes = (double*)calloc((size_t)2 * (size_t)LM * (size_t)myN + 1, sizeof(double));
    size_t sz = (size_t)NL + 2;
    double* buf = (double*)calloc(3 * sz, sizeof(double));
    double *vd_flux = buf, *zbar_n = buf + sz, *Z_n = buf + 2 * sz;
#define XN(c, nz, n) tr_xynodes[((size_t)((n) - 1) * (size_t)LM + (size_t)((nz) - 1)) * 2 + ((c) - 1)]
    for (int n = 1; n <= myN; n++) {                                                     /* :876-893 */
        int nl1 = o->nlevels_nod2D[n - 1] - 1, ul1 = o->ulevels_nod2D[n - 1];
        for (int nz = ul1; nz <= nl1; nz++) {
            double Tx = 0.0, Ty = 0.0;
            for (int k = 1; k <= o->nod_in_elem2D_num[n - 1]; k++) {
                int elem = I2(o->nod_in_elem2D, o->nod_in_elem2D_max, k, n);
                if (nz <= o->nlevels[elem - 1] - 1 && nz >= o->ulevels[elem - 1]) {
                    Tx = Tx + V3(tr_xy, 1, nz, elem) * o->elem_area[elem - 1];
                    Ty = Ty + V3(tr_xy, 2, nz, elem) * o->elem_area[elem - 1];
                }
            }
            XN(1, nz, n) = Tx / 3.0 / W2(o->areasvol, nz, n);
            XN(2, nz, n) = Ty / 3.0 / W2(o->areasvol, nz, n);
        }
    }
    for (int n = 1; n <= myN; n++) {                                                     /* :897-929 */
        int nl1 = o->nlevels_nod2D[n - 1] - 1, ul1 = o->ulevels_nod2D[n - 1];
        memset(buf, 0, 3 * sz * sizeof(double));
        zbar_n[nl1 + 1] = o->zbar_n_bot[n - 1];
        Z_n[nl1] = zbar_n[nl1 + 1] + T2(hn, nl1, n) / 2.0;
        for (int nz = nl1; nz >= ul1 + 1; nz--) {
            zbar_n[nz] = zbar_n[nz + 1] + T2(hn, nz, n);
            Z_n[nz - 1] = zbar_n[nz] + T2(hn, nz - 1, n) / 2.0;
        }
        zbar_n[ul1] = zbar_n[ul1 + 1] + T2(hn, ul1, n);
        for (int nz = ul1 + 1; nz <= nl1; nz++) {
            vd_flux[nz] = (Z_n[nz - 1] - zbar_n[nz]) * (S3(sl, 1, nz - 1, n) * XN(1, nz - 1, n) + S3(sl, 2, nz - 1, n) * XN(2, nz - 1, n)) * T2(Ki, nz - 1, n);
            vd_flux[nz] = vd_flux[nz] + (zbar_n[nz] - Z_n[nz]) * (S3(sl, 1, nz, n) * XN(1, nz, n) + S3(sl, 2, nz, n) * XN(2, nz, n)) * T2(Ki, nz, n);
            vd_flux[nz] = vd_flux[nz] / (Z_n[nz - 1] - Z_n[nz]) * W2(o->area, nz, n);
        }
        for (int nz = ul1; nz <= nl1; nz++)
            T2(o->del_ttf, nz, n) = T2(o->del_ttf, nz, n) + (vd_flux[nz] - vd_flux[nz + 1]) * dt / W2(o->areasvol, nz, n);
    }
#undef XN
    free(buf); free(tr_xynodes);
}

/* diff_part_bh, oce_ale_tracer.F90:1081-1148 — biharmonic diffusion of tracer ttf implemented as a filter (call site
 * :321-323, `if (smooth_bh_tra)`, after the implicit vertical diffusion).  Two rounds over the edges with two triangles
 * (myList_edge2D(ed) <= edge2D_in, passed here as edge_tri(2,ed) > 0: an edge past edge2D_in has no second triangle,
 * oce_mesh.F90:1456-1499); between them the reference exchanges temporary_ttf — the caller of a partitioned run does that
 * between o_diff_part_bh_round1 and _round2.  The reference's loop also covers eDim_edge2D edges: both nodes of those are
 * halo nodes, whose temporary_ttf the exchange overwrites and whose ttf the following exchange_nod(tr_arr) (:155)
 * overwrites, so own entries never see them.  gamma0/1/2 are gamma0_tra, gamma1_tra, gamma2_tra (oce_modules.F90:139-141). */
static void bh_edge_coef(oracle_t* o, int ed, int nz, int el1, int el2, double len, double* vi_out) {
    (void)ed;
    double u1 = V3(o->UV, 1, nz, el1) - V3(o->UV, 1, nz, el2);
    double v1 = V3(o->UV, 2, nz, el1) - V3(o->UV, 2, nz, el2);
    double vi = u1 * u1 + v1 * v1;
    vi = sqrt(dmax(o->gamma0_tra, dmax(o->gamma1_tra * sqrt(vi), o->gamma2_tra * vi)) * len);
    *vi_out = vi;
}
void o_diff_part_bh_round1(oracle_t* o, const double* ttf, double* temporary_ttf) {
    size_t cnt = (size_t)LM * (size_t)(o->myDim_nod2D + o->eDim_nod2D);
    memset(temporary_ttf, 0, cnt * sizeof(double));
    for (int ed = 1; ed <= o->myDim_edge2D; ed++) {
        int el1 = I2(o->edge_tri, 2, 1, ed), el2 = I2(o->edge_tri, 2, 2, ed);
        if (el2 <= 0) continue;
        int en1 = I2(o->edges, 2, 1, ed), en2 = I2(o->edges, 2, 2, ed);
        double len = sqrt(o->elem_area[el1 - 1] + o->elem_area[el2 - 1]);
        int ul1 = imin(o->ulevels_nod2D_max[en1 - 1], o->ulevels_nod2D_max[en2 - 1]);
        int nl1 = imax(o->nlevels_nod2D_min[en1 - 1], o->nlevels_nod2D_min[en2 - 1]) - 1;
        for (int nz = ul1; nz <= nl1; nz++) {
            double vi, tt = T2(ttf, nz, en1) - T2(ttf, nz, en2);
            bh_edge_coef(o, ed, nz, el1, el2, len, &vi);
            tt = tt * vi;
            T2(temporary_ttf, nz, en1) = T2(temporary_ttf, nz, en1) - tt;
            T2(temporary_ttf, nz, en2) = T2(temporary_ttf, nz, en2) + tt;
        }
    }
}
void o_diff_part_bh_round2(oracle_t* o, double* ttf, const double* temporary_ttf) {
    const double dt = o->dt;
    for (int ed = 1; ed <= o->myDim_edge2D; ed++) {
        int el1 = I2(o->edge_tri, 2, 1, ed), el2 = I2(o->edge_tri, 2, 2, ed);
        if (el2 <= 0) continue;
        int en1 = I2(o->edges, 2, 1, ed), en2 = I2(o->edges, 2, 2, ed);
        double len = sqrt(o->elem_area[el1 - 1] + o->elem_area[el2 - 1]);
        int ul1 = imin(o->ulevels_nod2D_max[en1 - 1], o->ulevels_nod2D_max[en2 - 1]);
        int nl1 = imax(o->nlevels_nod2D_min[en1 - 1], o->nlevels_nod2D_min[en2 - 1]) - 1;
        for (int nz = ul1; nz <= nl1; nz++) {
            double vi, tt = T2(temporary_ttf, nz, en1) - T2(temporary_ttf, nz, en2);
            bh_edge_coef(o, ed, nz, el1, el2, len, &vi);
            tt = -tt * vi * dt;
            T2(ttf, nz, en1) = T2(ttf, nz, en1) - tt / W2(o->area, nz, en1);
            T2(ttf, nz, en2) = T2(ttf, nz, en2) + tt / W2(o->area, nz, en2);
        }
    }
}
void o_diff_part_bh(oracle_t* o, double* ttf) {          /* one partition: no exchange between the rounds */
    o_diff_part_bh_round1(o, ttf, o->bh_tmp);
    o_diff_part_bh_round2(o, ttf, o->bh_tmp);
}

static double surface_flux(oracle_t* o, int n, int tr_num) {      /* the `flux` of diff_ver_part_expl_ale :354-364 */
    if (tr_num == 1) return -o->heat_flux[n - 1] / vcpw;
    if (tr_num == 2) return o->virtual_salt[n - 1] + o->relax_salt[n - 1] - o->real_salt_flux[n - 1] * o->is_nonlinfs;
    return 0.0;
}

/* oce_ale_tracer.F90:330-393 */
void o_diff_ver_part_expl_ale(oracle_t* o, int tr_num) {
    const double dt = o->dt;
    double* tr = o->tr_arr + (size_t)(tr_num - 1) * (size_t)LM * (size_t)(o->myDim_nod2D + o->eDim_nod2D);
    double* vd = (double*)calloc((size_t)NL + 2, sizeof(double));
    for (int n = 1; n <= o->myDim_nod2D; n++) {
        int nl1 = o->nlevels_nod2D[n - 1] - 1, ul1 = o->ulevels_nod2D[n - 1];
        for (int k = 0; k <= NL; k++) vd[k] = 0.0;
        vd[ul1] = surface_flux(o, n, tr_num);
        for (int nz = ul1 + 1; nz <= nl1; nz++) {
            double zinv1 = 1.0 / (T2(o->Z_3d_n, nz - 1, n) - T2(o->Z_3d_n, nz, n));
            vd[nz] = (W2(o->Kv, nz, n) + 0.0) * (T2(tr, nz - 1, n) - T2(tr, nz, n)) * zinv1 * W2(o->area, nz, n);
        }
        for (int nz = ul1; nz <= nl1 - 1; nz++)
            T2(o->del_ttf, nz, n) = T2(o->del_ttf, nz, n) + (vd[nz] - vd[nz + 1]) / (W2(o->zbar_3d_n, nz, n) - W2(o->zbar_3d_n, nz + 1, n)) * dt / W2(o->areasvol, nz, n);
        T2(o->del_ttf, nl1, n) = T2(o->del_ttf, nl1, n) + (vd[nl1] / (W2(o->zbar_3d_n, nl1, n) - W2(o->zbar_3d_n, nl1 + 1, n))) * dt / W2(o->areasvol, nl1, n);
    }
    free(vd);
}

/* bc_surface, oce_ale_tracer.F90:1154-1195 */
static double bc_surface(oracle_t* o, int n, int id) {
    const double dt = o->dt;
    int N = o->myDim_nod2D + o->eDim_nod2D; (void)N;
    if (id == 0) return -dt * (o->heat_flux[n - 1] / vcpw + T2(o->tr_arr, o->ulevels_nod2D[n - 1], n) * o->water_flux[n - 1] * o->is_nonlinfs);
    if (id == 1) return dt * (o->virtual_salt[n - 1] + o->relax_salt[n - 1] - o->real_salt_flux[n - 1] * o->is_nonlinfs);
    return 0.0;   /* ids 301..303 */
}

/* oce_ale_tracer.F90:398-856 */
void o_diff_ver_part_impl_ale(oracle_t* o, int tr_num) {
    const double dt = o->dt;
    int N = o->myDim_nod2D + o->eDim_nod2D;
    double* trc = o->tr_arr + (size_t)(tr_num - 1) * (size_t)LM * (size_t)N;
    int do_wimpl = 1;
    if (o->tra_adv_lim == O_LIM_FCT || !o->w_split) do_wimpl = 0;
    double isredi = o->Redi ? 1.0 : 0.0;
    size_t sz = (size_t)NL + 2;
    double* buf = (double*)calloc(8 * sz, sizeof(double));
    double *a = buf, *b = buf + sz, *c = buf + 2 * sz, *tr = buf + 3 * sz, *cp = buf + 4 * sz, *tp = buf + 5 * sz, *zbar_n = buf + 6 * sz, *Z_n = buf + 7 * sz;
    const double *Kv = o->Kv, *Ki = o->Ki, *sl = o->slope_tapered, *hn = o->hnode_new, *area = o->area, *asv = o->areasvol, *Wi = o->Wvel_i;
#define SL2(nz) (S3(sl, 3, nz, n) * S3(sl, 3, nz, n))
    for (int n = 1; n <= o->myDim_nod2D; n++) {
        memset(buf, 0, 8 * sz * sizeof(double));
        int nzmax = o->nlevels_nod2D[n - 1], nzmin = o->ulevels_nod2D[n - 1];
        zbar_n[nzmax] = o->zbar_n_bot[n - 1];
        Z_n[nzmax - 1] = zbar_n[nzmax] + T2(hn, nzmax - 1, n) / 2.0;
        for (int nz = nzmax - 1; nz >= nzmin + 1; nz--) {
            zbar_n[nz] = zbar_n[nz + 1] + T2(hn, nz, n);
            Z_n[nz - 1] = zbar_n[nz] + T2(hn, nz - 1, n) / 2.0;
        }
        zbar_n[nzmin] = zbar_n[nzmin + 1] + T2(hn, nzmin, n);
        int nz = nzmin;
        double zinv2 = 1.0 / (Z_n[nz] - Z_n[nz + 1]);
        double zinv = 1.0 * dt, zinv1, Ty, Ty1, v_adv;
        Ty1 = (Z_n[nz] - zbar_n[nz + 1]) * zinv2 * SL2(nz) * T2(Ki, nz, n) + (zbar_n[nz + 1] - Z_n[nz + 1]) * zinv2 * SL2(nz + 1) * T2(Ki, nz + 1, n);
        Ty1 = Ty1 * isredi;
        a[nz] = 0.0;
        c[nz] = -(W2(Kv, nz + 1, n) + Ty1) * zinv2 * zinv * W2(area, nz + 1, n) / W2(asv, nz, n);
        b[nz] = -c[nz] + T2(hn, nz, n);
        if (do_wimpl) {
            v_adv = zinv * (W2(area, nz, n) / W2(asv, nz, n));
            b[nz] = b[nz] + W2(Wi, nz, n) * v_adv;
            v_adv = zinv * W2(area, nz + 1, n) / W2(asv, nz, n);
            b[nz] = b[nz] - dmin(0.0, W2(Wi, nz + 1, n)) * v_adv;
            c[nz] = c[nz] - dmax(0.0, W2(Wi, nz + 1, n)) * v_adv;
        }
        zinv1 = zinv2;
        for (nz = nzmin + 1; nz <= nzmax - 2; nz++) {
            zinv2 = 1.0 / (Z_n[nz] - Z_n[nz + 1]);
            Ty = (Z_n[nz - 1] - zbar_n[nz]) * zinv1 * SL2(nz - 1) * T2(Ki, nz - 1, n) + (zbar_n[nz] - Z_n[nz]) * zinv1 * SL2(nz) * T2(Ki, nz, n);
            Ty1 = (Z_n[nz] - zbar_n[nz + 1]) * zinv2 * SL2(nz) * T2(Ki, nz, n) + (zbar_n[nz + 1] - Z_n[nz + 1]) * zinv2 * SL2(nz + 1) * T2(Ki, nz + 1, n);
            Ty = Ty * isredi; Ty1 = Ty1 * isredi;
            a[nz] = -(W2(Kv, nz, n) + Ty) * zinv1 * zinv * (W2(area, nz, n) / W2(asv, nz, n));
            c[nz] = -(W2(Kv, nz + 1, n) + Ty1) * zinv2 * zinv * W2(area, nz + 1, n) / W2(asv, nz, n);
            b[nz] = -a[nz] - c[nz] + T2(hn, nz, n);
            zinv1 = zinv2;
            if (do_wimpl) {
                v_adv = zinv * (W2(area, nz, n) / W2(asv, nz, n));
                a[nz] = a[nz] + dmin(0.0, W2(Wi, nz, n)) * v_adv;
                b[nz] = b[nz] + dmax(0.0, W2(Wi, nz, n)) * v_adv;
                v_adv = zinv * W2(area, nz + 1, n) / W2(asv, nz, n);
                b[nz] = b[nz] - dmin(0.0, W2(Wi, nz + 1, n)) * v_adv;
                c[nz] = c[nz] - dmax(0.0, W2(Wi, nz + 1, n)) * v_adv;
            }
        }
        nz = nzmax - 1;
        zinv = 1.0 * dt;
        Ty = (Z_n[nz - 1] - zbar_n[nz]) * zinv1 * SL2(nz - 1) * T2(Ki, nz - 1, n) + (zbar_n[nz] - Z_n[nz]) * zinv1 * SL2(nz) * T2(Ki, nz, n);
        Ty = Ty * isredi;
        a[nz] = -(W2(Kv, nz, n) + Ty) * zinv1 * zinv * (W2(area, nz, n) / W2(asv, nz, n));
        c[nz] = 0.0;
        b[nz] = -a[nz] + T2(hn, nz, n);
        if (do_wimpl) {
            v_adv = zinv * (W2(area, nz, n) / W2(asv, nz, n));
            a[nz] = a[nz] + dmin(0.0, W2(Wi, nz, n)) * v_adv;
            b[nz] = b[nz] + dmax(0.0, W2(Wi, nz, n)) * v_adv;
        }
        nz = nzmin;
        double dz = T2(hn, nz, n);
        tr[nz] = -(b[nz] - dz) * T2(trc, nz, n) - c[nz] * T2(trc, nz + 1, n);
        for (nz = nzmin + 1; nz <= nzmax - 2; nz++) {
            dz = T2(hn, nz, n);
            tr[nz] = -a[nz] * T2(trc, nz - 1, n) - (b[nz] - dz) * T2(trc, nz, n) - c[nz] * T2(trc, nz + 1, n);
        }
        nz = nzmax - 1;
        dz = T2(hn, nz, n);
        tr[nz] = -a[nz] * T2(trc, nz - 1, n) - (b[nz] - dz) * T2(trc, nz, n);
        if (o->use_kpp_nonlclflx && tr_num <= 2) {            /* KPP non-local fluxes, oce_ale_tracer.F90:688-781 (both KPP variants: same formula) */
            const double* g = tr_num == 1 ? o->kpp_gT : o->kpp_gS;
            const double* bl = tr_num == 1 ? o->kpp_b2 : o->kpp_b3;
            double rsss = o->ref_sss;
            if (tr_num == 2 && o->ref_sss_local) rsss = T2(trc, 1, n);                     /* tr_arr(1,n,2) */
#define KM(z) dmin(W2(g, z, n) * W2(bl, z, n), 1.0)
            for (nz = nzmin; nz <= nzmax - 1; nz++) {
                double x;
                if (nz == nzmin) x = (-KM(nz + 1) * (W2(area, nz + 1, n) / W2(asv, nz, n)));
                else if (nz <= nzmax - 2) x = (KM(nz) * (W2(area, nz, n) / W2(asv, nz, n)) - KM(nz + 1) * (W2(area, nz + 1, n) / W2(asv, nz, n)));
                else x = (KM(nz) * (W2(area, nz, n) / W2(asv, nz, n)));
                if (tr_num == 1) tr[nz] = tr[nz] + x * o->heat_flux[n - 1] / vcpw * dt;
                else tr[nz] = tr[nz] - x * rsss * o->water_flux[n - 1] * dt;
            }
#undef KM
        }
        if (o->use_sw_pene && tr_num == 1)
            for (nz = nzmin; nz <= nzmax - 1; nz++) {
                zinv = 1.0 * dt;
                tr[nz] = tr[nz] + (W2(o->sw_3d, nz, n) - W2(o->sw_3d, nz + 1, n) * W2(area, nz + 1, n) / W2(asv, nz, n)) * zinv;
            }
        tr[nzmin] = tr[nzmin] + bc_surface(o, n, o->tracer_id[tr_num - 1]);
        thomas(nzmin, nzmax, a, b, c, tr, cp, tp);
        for (nz = nzmin; nz <= nzmax - 1; nz++) T2(trc, nz, n) = T2(trc, nz, n) + tr[nz];
    }
#undef SL2
    free(buf);
}

/* ------------------------------------------------------------------------------------------------
 * Phases of solve_tracers_ale (oce_ale_tracer.F90:101-199), cut at the reference's halo exchanges.    */
static double* TR(oracle_t* o, int tr_num) { return o->tr_arr + (size_t)(tr_num - 1) * (size_t)LM * (size_t)(o->myDim_nod2D + o->eDim_nod2D); }
static double* TRO(oracle_t* o, int tr_num) { return o->tr_arr_old + (size_t)(tr_num - 1) * (size_t)LM * (size_t)(o->myDim_nod2D + o->eDim_nod2D); }

void o_phase1(oracle_t* o, int tr_num) {            /* init_tracers_AB, oce_tracer_mod.F90:61-68 */
    size_t cnt = (size_t)LM * (size_t)(o->myDim_nod2D + o->eDim_nod2D);
    double *t = TR(o, tr_num), *told = TRO(o, tr_num);
    memset(o->del_ttf, 0, cnt * sizeof(double));
    for (size_t q = 0; q < cnt; q++) told[q] = -(0.5 + o->epsilon) * told[q] + (1.5 + o->epsilon) * t[q];
    o_tracer_gradient_elements(o, told);
}
void o_phase2(oracle_t* o, int tr_num) {            /* oce_tracer_mod.F90:71-81 */
    o_tracer_gradient_z(o, TR(o, tr_num));
    o_fill_up_dn_grad(o);
    o_tracer_gradient_elements(o, TR(o, tr_num));
}
void o_phase3(oracle_t* o, int tr_num) {            /* adv_tracers_ale :234-235 + do_oce_adv_tra :79-134 */
    size_t cnt = (size_t)LM * (size_t)(o->myDim_nod2D + o->eDim_nod2D);
    const double dt = o->dt;
    const double* ttf = TR(o, tr_num);
    memset(o->del_ttf_advhoriz, 0, cnt * sizeof(double));
    memset(o->del_ttf_advvert, 0, cnt * sizeof(double));
    if (o->tra_adv_lim != O_LIM_FCT) return;
    o_adv_tra_hor_upw1(o, ttf, o->UV, o->adv_flux_hor, 1);
    memset(o->fct_LO, 0, cnt * sizeof(double));
    for (int e = 1; e <= o->myDim_edge2D; e++) {
        int en1 = I2(o->edges, 2, 1, e), en2 = I2(o->edges, 2, 2, e);
        int el1 = I2(o->edge_tri, 2, 1, e), el2 = I2(o->edge_tri, 2, 2, e);
        int nl1 = o->nlevels[el1 - 1] - 1, nu1 = o->ulevels[el1 - 1], nl2 = 0, nu2 = 0;
        if (el2 > 0) { nl2 = o->nlevels[el2 - 1] - 1; nu2 = o->ulevels[el2 - 1]; }
        int nl12 = imax(nl1, nl2), nu12 = nu1;
        if (nu2 > 0) nu12 = imin(nu1, nu2);
        for (int nz = nu12; nz <= nl12; nz++) {
            T2(o->fct_LO, nz, en1) = T2(o->fct_LO, nz, en1) + T2(o->adv_flux_hor, nz, e);
            T2(o->fct_LO, nz, en2) = T2(o->fct_LO, nz, en2) - T2(o->adv_flux_hor, nz, e);
        }
    }
    o_adv_tra_ver_upw1(o, ttf, o->Wvel_e, o->adv_flux_ver, 1);
    for (int n = 1; n <= o->myDim_nod2D; n++)
        for (int nz = o->ulevels_nod2D[n - 1]; nz <= o->nlevels_nod2D[n - 1] - 1; nz++)
            T2(o->fct_LO, nz, n) = (T2(ttf, nz, n) * T2(o->hnode, nz, n) +
                                    (T2(o->fct_LO, nz, n) + (W2(o->adv_flux_ver, nz, n) - W2(o->adv_flux_ver, nz + 1, n))) * dt / W2(o->areasvol, nz, n)) /
                                   T2(o->hnode_new, nz, n);
    if (o->w_split) {
        o_adv_tra_vert_impl(o, o->fct_LO, o->Wvel_i);
        o_adv_tra_ver_upw1(o, ttf, o->Wvel, o->adv_flux_ver, 1);
    }
}
void o_phase4(oracle_t* o, int tr_num) {            /* do_oce_adv_tra :137-191 up to the exchange inside oce_tra_adv_fct */
    const double* ttf = TR(o, tr_num);
    const double* ttfAB = TRO(o, tr_num);
    int zero = (o->tra_adv_lim == O_LIM_FCT) ? 0 : 1;
    if (o->tra_adv_hor == O_HOR_MUSCL) o_adv_tra_hor_muscl(o, ttfAB, o->UV, o->tra_adv_ph, o->adv_flux_hor, zero, 0);
    else if (o->tra_adv_hor == O_HOR_MFCT) o_adv_tra_hor_muscl(o, ttfAB, o->UV, o->tra_adv_ph, o->adv_flux_hor, zero, 1);
    else o_adv_tra_hor_upw1(o, ttfAB, o->UV, o->adv_flux_hor, zero);
    const double* pw = (o->tra_adv_lim == O_LIM_FCT) ? o->Wvel : o->Wvel_e;
    if (o->tra_adv_ver == O_VER_QR4C) o_adv_tra_ver_qr4c(o, ttfAB, pw, o->tra_adv_pv, o->adv_flux_ver, zero);
    else if (o->tra_adv_ver == O_VER_CDIFF) o_adv_tra_ver_cdiff(o, ttfAB, pw, o->adv_flux_ver, zero);
    else if (o->tra_adv_ver == O_VER_PPM) o_adv_tra_vert_ppm(o, ttfAB, pw, o->adv_flux_ver, zero);
    else o_adv_tra_ver_upw1(o, ttfAB, pw, o->adv_flux_ver, zero);
    if (o->tra_adv_lim == O_LIM_FCT) o_oce_tra_adv_fct_a(o, ttf, o->fct_LO, o->adv_flux_hor, o->adv_flux_ver);
}
void o_phase5(oracle_t* o, int tr_num) {
    size_t cnt = (size_t)LM * (size_t)(o->myDim_nod2D + o->eDim_nod2D);
    double* ttf = TR(o, tr_num);
    double* told = TRO(o, tr_num);
    if (o->tra_adv_lim == O_LIM_FCT) {
        o_oce_tra_adv_fct_b(o, o->adv_flux_hor, o->adv_flux_ver);
        o_oce_tra_adv_flux2dtracer(o, o->del_ttf_advhoriz, o->del_ttf_advvert, o->adv_flux_hor, o->adv_flux_ver, 1, ttf, o->fct_LO);
    } else
        o_oce_tra_adv_flux2dtracer(o, o->del_ttf_advhoriz, o->del_ttf_advvert, o->adv_flux_hor, o->adv_flux_ver, 0, NULL, NULL);
    for (size_t q = 0; q < cnt; q++) o->del_ttf[q] = o->del_ttf[q] + o->del_ttf_advhoriz[q] + o->del_ttf_advvert[q];   /* oce_ale_tracer.F90:240 */
    /* diff_tracers_ale, oce_ale_tracer.F90:253-325 */
    memcpy(told, ttf, cnt * sizeof(double));
    if (o->do_hor_diff) o_diff_part_hor_redi(o);
    if (!o->i_vert_diff) o_diff_ver_part_expl_ale(o, tr_num);
    if (o->Redi && !o->skip_redi_expl) o_diff_ver_part_redi_expl(o);                      /* :288 (skip_redi_expl: test switch) */
    for (int n = 1; n <= o->myDim_nod2D; n++) {
        int nzmax = o->nlevels_nod2D[n - 1] - 1, nzmin = o->ulevels_nod2D[n - 1];
        for (int nz = nzmin; nz <= nzmax; nz++) T2(o->del_ttf, nz, n) = T2(o->del_ttf, nz, n) + T2(ttf, nz, n) * (T2(o->hnode, nz, n) - T2(o->hnode_new, nz, n));
        for (int nz = nzmin; nz <= nzmax; nz++) T2(ttf, nz, n) = T2(ttf, nz, n) + T2(o->del_ttf, nz, n) / T2(o->hnode_new, nz, n);
    }
    if (o->i_vert_diff) o_diff_ver_part_impl_ale(o, tr_num);
    /* smooth_bh_tra (:321-323): one partition here; a partitioned caller sets bh_split and runs o_diff_part_bh_round1,
     * exchange_nod(temporary_ttf), o_diff_part_bh_round2 itself */
    if (o->smooth_bh_tra && !o->bh_split) o_diff_part_bh(o, ttf);
}
/* relax_to_clim, oce_tracer_mod.F90:87-119: own nodes, wet levels, tracers 1 (Tclim) and 2 (Sclim) */
void o_relax_to_clim(oracle_t* o, int tr_num) {
    if (!(o->clim_relax > 1.0e-8) || tr_num > 2 || !o->relax2clim) return;
    const double* clim = tr_num == 1 ? o->Tclim : o->Sclim;
    if (!clim) return;
    double* ttf = TR(o, tr_num);
    for (int n = 1; n <= o->myDim_nod2D; n++) {
        int nzmin = o->ulevels_nod2D[n - 1], nzmax = o->nlevels_nod2D[n - 1];
        for (int nz = nzmin; nz <= nzmax - 1; nz++)
            T2(ttf, nz, n) = T2(ttf, nz, n) + o->relax2clim[n - 1] * o->dt * (T2(clim, nz, n) - T2(ttf, nz, n));
    }
}
/* relax_zonal_temp, toy_channel_soufflet.F90:78-103: own + halo nodes, levels 1..nlevels_nod2D-1, ALWAYS tracer 1 (the
 * caller's tracer loop reaches it once per tracer, oce_ale_tracer.F90:150-151, so temperature is relaxed num_tracers times
 * per step); the bin index limit is the literal 100 of the reference, not nybins */
void o_relax_zonal_temp(oracle_t* o) {
    const int N = o->myDim_nod2D + o->eDim_nod2D, L = o->nl - 1;
    const double dy = o->zonal_dy;
    double* ttf = TR(o, 1);
    for (int n = 1; n <= N; n++) {
        double yy = o->coord_nod2D[2 * (n - 1) + 1] - o->zonal_lat0;
        double a = 0;
        int nn, nn1;
        if (yy < dy / 2) { nn = 1; nn1 = 1; }
        else {
            nn = (int)floor(yy / dy - 0.5) + 1;
            nn1 = nn + 1;
            if (nn1 > 100) nn1 = nn;
            a = yy / dy + 0.5 - nn;
        }
        for (int nz = 1; nz <= o->nlevels_nod2D[n - 1] - 1; nz++) {
            double Tzon = (1.0 - a) * o->zonal_ztem[(size_t)(nn - 1) * L + nz - 1] + a * o->zonal_ztem[(size_t)(nn1 - 1) * L + nz - 1];
            T2(ttf, nz, n) = T2(ttf, nz, n) + o->dt * o->zonal_tau_inv * (T2(o->Tclim, nz, n) - Tzon);
        }
    }
}
void o_phase_final(oracle_t* o) {                   /* oce_ale_tracer.F90:176-198 */
    if (!o->clamp_salt || o->num_tracers < 2) return;
    int N = o->myDim_nod2D + o->eDim_nod2D;
    double* s = TR(o, 2);
    for (int node = 1; node <= N; node++) {
        int nzmax = o->nlevels_nod2D[node - 1] - 1, nzmin = o->ulevels_nod2D[node - 1];
        for (int nz = nzmin; nz <= nzmax; nz++) if (T2(s, nz, node) > 45.0) T2(s, nz, node) = 45.0;
        for (int nz = nzmin; nz <= nzmax; nz++) if (T2(s, nz, node) < 3.0) T2(s, nz, node) = 3.0;
    }
}
void o_solve_tracers_ale(oracle_t* o) {
    for (int tr = 1; tr <= o->num_tracers; tr++) {
        o_phase1(o, tr); o_phase2(o, tr); o_phase3(o, tr); o_phase4(o, tr); o_phase5(o, tr);
        if (o->relax_zonal) o_relax_zonal_temp(o);      /* oce_ale_tracer.F90:150-154 */
        else o_relax_to_clim(o, tr);
    }
    o_phase_final(o);
}
