/* ORACLE (test infrastructure) — SSH solve: psolve.c's row scaling + the Krylov loop of pARMS'
 * bicgstab_ras.c, serial, with a pluggable preconditioner.  The reference's own RAS/ILU(k) preconditioner is
 * not restated (the product replaces it by block-Jacobi); the real pARMS build under oracle/_ref is the
 * pin for "same solution to solver tolerance".
 *   pc_kind 0: none      1: block-Jacobi, dense inverse of `block` x `block` diagonal blocks (block=1: Jacobi)
 */
#include "oracle.h"
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

void o_spmv(int n, const int* rp, const int* ci, const double* a, const double* x, double* y) {   /* parms_mat_vcsr.c MatVec */
    for (int i = 0; i < n; i++) {
        double s = 0.0;
        for (int j = rp[i]; j < rp[i + 1]; j++) s += a[j] * x[ci[j]];
        y[i] = s;
    }
}

/* dense inverse of the diagonal blocks of the (scaled) matrix, Gauss-Jordan with partial pivoting */
static int bj_setup(int n, const int* rp, const int* ci, const double* a, int bs, double* inv) {
    double* m = (double*)malloc(sizeof(double) * (size_t)bs * 2 * (size_t)bs);
    for (int b0 = 0; b0 < n; b0 += bs) {
        int m_ = (n - b0 < bs) ? n - b0 : bs;
        for (int r = 0; r < bs; r++) for (int c = 0; c < 2 * bs; c++) m[r * 2 * bs + c] = (c == r + bs) ? 1.0 : 0.0;
        for (int r = 0; r < bs; r++) if (r >= m_) m[r * 2 * bs + r] = 1.0;
        for (int r = 0; r < m_; r++)
            for (int j = rp[b0 + r]; j < rp[b0 + r + 1]; j++) { int c = ci[j] - b0; if (c >= 0 && c < m_) m[r * 2 * bs + c] += a[j]; }
        for (int k = 0; k < bs; k++) {
            int p = k; double best = fabs(m[k * 2 * bs + k]);
            for (int r = k + 1; r < bs; r++) if (fabs(m[r * 2 * bs + k]) > best) { best = fabs(m[r * 2 * bs + k]); p = r; }
            if (best == 0.0) { free(m); return 1; }
            if (p != k) for (int c = 0; c < 2 * bs; c++) { double t = m[k * 2 * bs + c]; m[k * 2 * bs + c] = m[p * 2 * bs + c]; m[p * 2 * bs + c] = t; }
            double d = 1.0 / m[k * 2 * bs + k];
            for (int c = 0; c < 2 * bs; c++) m[k * 2 * bs + c] *= d;
            for (int r = 0; r < bs; r++) if (r != k) {
                double f = m[r * 2 * bs + k];
                if (f != 0.0) for (int c = 0; c < 2 * bs; c++) m[r * 2 * bs + c] -= f * m[k * 2 * bs + c];
            }
        }
        double* out = inv + (size_t)(b0 / bs) * (size_t)bs * (size_t)bs;
        for (int r = 0; r < bs; r++) for (int c = 0; c < bs; c++) out[r * bs + c] = m[r * 2 * bs + bs + c];
    }
    free(m);
    return 0;
}
static void bj_apply(int n, int bs, const double* inv, const double* x, double* y) {
    for (int b0 = 0; b0 < n; b0 += bs) {
        const double* blk = inv + (size_t)(b0 / bs) * (size_t)bs * (size_t)bs;
        int m_ = (n - b0 < bs) ? n - b0 : bs;
        for (int r = 0; r < m_; r++) {
            double s = 0.0;
            for (int c = 0; c < m_; c++) s += blk[r * bs + c] * x[b0 + c];
            y[b0 + r] = s;
        }
    }
}

/* psolve (src/psolve.c:152-221) + parms_bicgstabras (lib/parms/src/bicgstab_ras.c:49-259), one rank.
 * rowptr0/colind0 are 0-based, x holds the initial guess on entry (warm start, psolve.c:206). */
int o_ssh_solve(int n, const int* rp, const int* ci, const double* vals, const double* rhs, double* x,
                int pc_kind, int block, double soltol, int maxits, int* its_out, double* res_out) {
    size_t nnz = (size_t)rp[n];
    double* a = (double*)malloc(sizeof(double) * nnz);
    double* scale = (double*)malloc(sizeof(double) * (size_t)n);
    for (int i = 0; i < n; i++) {                              /* psolve.c:186-193 */
        double tmp = 0.;
        for (int j = rp[i]; j < rp[i + 1]; j++) tmp += fabs(vals[j]);
        scale[i] = 1. / tmp;
        for (int j = rp[i]; j < rp[i + 1]; j++) a[j] = vals[j] * scale[i];
    }
    double* buf = (double*)calloc((size_t)8 * (size_t)n, sizeof(double));
    double *y = buf, *r0 = buf, *r = buf + n, *p = buf + 2 * (size_t)n, *pt = buf + 3 * (size_t)n, *st = buf + 4 * (size_t)n,
           *v = buf + 5 * (size_t)n, *t = buf + 6 * (size_t)n;
    for (int i = 0; i < n; i++) y[i] = rhs[i] * scale[i];      /* psolve.c:208-209; r0 = y */
    double* inv = NULL;
    if (pc_kind == 1) {
        size_t nb = (size_t)((n + block - 1) / block);
        inv = (double*)malloc(sizeof(double) * nb * (size_t)block * (size_t)block);
        if (bj_setup(n, rp, ci, a, block, inv)) { free(a); free(scale); free(buf); free(inv); return 2; }
    }
#define PCAPPLY(in, out) do { if (pc_kind == 1) bj_apply(n, block, inv, in, out); else memcpy(out, in, sizeof(double) * (size_t)n); } while (0)
    double tol = soltol * soltol;
    o_spmv(n, rp, ci, a, x, r);
    for (int i = 0; i < n; i++) r[i] = r0[i] - r[i];           /* parms_VecAYPX(r, r0, -1) */
    double omega = 0., rho = 0.;
    for (int i = 0; i < n; i++) { omega += r[i] * r[i]; rho += r0[i] * r[i]; }
    memcpy(p, r, sizeof(double) * (size_t)n);
    double alpha = 0., sigma = 1.0, rho_alt = 1.0, beta;
    int its = 0;
    if (omega > tol) {
        for (its = 0; its < maxits; its++) {
            PCAPPLY(p, pt);
            o_spmv(n, rp, ci, a, pt, v);
            alpha = 0.; for (int i = 0; i < n; i++) alpha += r0[i] * v[i];
            alpha = rho / alpha;
            for (int i = 0; i < n; i++) r[i] -= alpha * v[i];
            PCAPPLY(r, st);
            o_spmv(n, rp, ci, a, st, t);
            double t0 = 0., t1 = 0.;
            for (int i = 0; i < n; i++) { t0 += t[i] * r[i]; t1 += t[i] * t[i]; }
            sigma = t0 / t1;
            for (int i = 0; i < n; i++) r[i] -= sigma * t[i];
            t0 = 0.; t1 = 0.;
            for (int i = 0; i < n; i++) { t0 += r[i] * r[i]; t1 += r0[i] * r[i]; }
            for (int i = 0; i < n; i++) x[i] += alpha * pt[i] + sigma * st[i];
            omega = t0;
            if (omega < tol) break;
            rho = t1;
            beta = (rho * alpha) / (rho_alt * sigma);
            rho_alt = rho;
            for (int i = 0; i < n; i++) p[i] = r[i] + beta * (p[i] - sigma * v[i]);
        }
    }
    its++;
    if (its_out) *its_out = its;
    if (res_out) *res_out = sqrt(omega);
    int rc = (its == maxits + 1 && omega > tol) ? 3 : 0;
    free(a); free(scale); free(buf); free(inv);
    return rc;
}
