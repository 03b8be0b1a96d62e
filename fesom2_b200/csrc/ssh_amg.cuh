// Aggregation multigrid preconditioner of the SSH solve (replaces parms_pc_ras.c + ILU(k), lib/parms/src/parms_pc_ras.c:56-120,
// parms_ilu_vcsr.c:651-1128, whose triangular solves are sequential).  Like the reference's factor it is built ONCE
// from the initial matrix (psolve.c:200 keeps the ILU factor when values are refreshed with new=1) and applied as a
// fixed linear operator, so right-preconditioned BiCGStab (bicgstab_ras.c:49-259) is used unchanged.
//
//   setup (host, once)   strength graph |a_ij| >= theta sqrt(a_ii a_jj)  ->  greedy aggregation  ->  tentative
//                        piecewise-constant prolongator smoothed with one damped-Jacobi step on the FILTERED matrix
//                        (weak entries lumped onto the diagonal)  ->  Galerkin coarse matrix R A P, R = P^T;
//                        repeated for up to max_levels levels.
//   apply (device)       one V(1,1) cycle with damped-Jacobi smoothing, zero initial guess on every level; every
//                        matrix (A_l, P_l, R_l) is stored as sliced ELL (32-row slices) and every step of the cycle is a
//                        row-gather SpMV: no atomics, bitwise reproducible.
#pragma once
#include "common.cuh"
#include "ssh_peer.cuh"

struct HostCsr {
    int nrow = 0, ncol = 0;
    std::vector<int> rp, ci;
    std::vector<double> v;
    size_t nnz() const { return ci.size(); }
};
struct AmgHostLevel {
    HostCsr A, P, R;                 // P: nrow(A) x n_coarse, R = P^T; both empty on the coarsest level
    std::vector<double> dinv;        // 1 / diag(A)
    HostCsr M, T, Q;                 // the V(1,1) cycle's steps on this level folded into three matrices (amg_fold_cycle); empty on the coarsest level
};
struct AmgParams {
    double theta = 0.08;             // strength-of-connection threshold
    double omega_p = 2.0 / 3.0;      // prolongator smoothing weight
    double omega = 0.8;              // damped-Jacobi smoother weight
    int max_levels = 4, min_rows = 200;
};
// pure host code (no CUDA call): usable from the CPU tests through fb200_amg_host_*
void amg_host_setup(const HostCsr& A0, const AmgParams& prm, std::vector<AmgHostLevel>& out);
// With damped-Jacobi smoothing from a zero guess every step of the V(1,1) cycle is linear in the level's right-hand side b, and
// consecutive steps can be multiplied out once on the host (W = omega D^-1, S = I - W A):
//   down   b_c = R (b - A W b)                          = M b        M = R (I - A W)
//   up     x   = S (W b + P x_c) + W b                  = T b + Q x_c  T = (I + S) W,  Q = S P
//   coarsest x = W b  — folded into the Q of the level above (Q <- Q W_c).
// Same operator as the step-by-step cycle up to rounding; 2 (K-1) kernels per cycle instead of 4 (K-1), no intermediate vectors,
// and M + T + Q hold fewer entries than 2 A + P + R.
// The last level above the coarsest is small: there down + coarsest + up collapse further into ONE matrix, x = (T + Q M) b =: C b,
// kept in T (M and Q of that level stay empty).  Only below AMG_COLLAPSE_ROWS rows — the product is dense (100-200 entries per row): at 16 k rows
// the one dense kernel takes as long as the three it replaces (profiles/r02_c15_*.json).
#define AMG_COLLAPSE_ROWS 8192
void amg_fold_cycle(std::vector<AmgHostLevel>& lv, double omega);
bool amg_folded_on();               // default on; FB200_AMG_FOLD=0 keeps the step-by-step kernels

struct CycleHead { PushArgs pa; PeerCtl pc; SolverScal* sc; int mode; };

struct DevSell {                     // sliced ELL, slice = 32 rows: val[sptr[s] + k*32 + lane]
    int nrow = 0, nslice = 0;
    size_t entries = 0;
    DevBuf<int> sptr, scol;
    DevBuf<double> sval;
    void upload(const HostCsr& m);
    void release() { sptr.release(); scol.release(); sval.release(); }
};
struct AmgDevLevel {
    int n = 0;
    DevSell A, P, R;                 // step-by-step cycle
    DevSell M, T, Q;                 // folded cycle (only one of the two sets is uploaded)
    DevBuf<double> dinv, b, x, x2, r;
};
struct Amg {
    std::vector<AmgDevLevel> L;
    DevBuf<unsigned> bar;            // grid barrier of the fused coarse-level kernel
    double omega = 0.8;
    bool folded = false;
    bool ready() const { return !L.empty(); }
    int levels() const { return (int)L.size(); }
    void upload(const std::vector<AmgHostLevel>& h, double omega_);
    // out = V-cycle(b0) where the caller has already filled L[0].b (= D * in) and L[0].x (= omega dinv0 b0), see k_amg_enter
    // in ssh_solver.cu.  `done` points at the solver's device-side convergence flag: all kernels are no-ops once it is set.
    // head (folded cycle, peer-memory transport, overlapping subdomains): the first kernel of the cycle pushes this rank's b0 to the
    // ranks whose subdomains overlap its rows and waits for its own overlap rows (push_and_wait, ssh_peer.cuh)
    void cycle(double* out, const int* done, cudaStream_t st, int n_out = 0, const CycleHead* head = nullptr);
    void release();
};

#ifdef __CUDACC__
// one SELL row times x.  Column indices and values of four entries are fetched before the dependent x gathers are
// issued, so a thread keeps 8-12 loads in flight instead of one (the sum itself stays in entry order).
__device__ __forceinline__ double sell_row(const int* __restrict__ sptr, const int* __restrict__ scol, const double* __restrict__ sval,
                                           const double* __restrict__ x, int i) {
    const int s0 = sptr[i >> 5];
    const int w = (sptr[(i >> 5) + 1] - s0) >> 5;
    const int base = s0 + (i & 31);
    double s = 0.0;
    int k = 0;
    for (; k + 4 <= w; k += 4) {
        const int o = base + (k << 5);
        const int c0 = scol[o], c1 = scol[o + 32], c2 = scol[o + 64], c3 = scol[o + 96];
        const double a0 = sval[o], a1 = sval[o + 32], a2 = sval[o + 64], a3 = sval[o + 96];
        const double x0 = x[c0], x1 = x[c1], x2 = x[c2], x3 = x[c3];
        s += a0 * x0; s += a1 * x1; s += a2 * x2; s += a3 * x3;
    }
    for (; k < w; k++) s += sval[base + (k << 5)] * x[scol[base + (k << 5)]];
    return s;
}
#endif
