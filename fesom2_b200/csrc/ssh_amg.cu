// Aggregation multigrid preconditioner of the SSH solve: host setup + device V-cycle.  See ssh_amg.cuh.
#include "ssh_amg.cuh"
#include <cstdlib>
#include <cstring>
#include "fesom_b200.h"
#include <algorithm>
#include <cmath>
#include <map>

// ---------------------------------------------------------------------------------------------------
// host setup (pure C++)
static void csr_sort_rows(HostCsr& m) {
    std::vector<std::pair<int, double>> tmp;
    for (int i = 0; i < m.nrow; i++) {
        const int a = m.rp[i], b = m.rp[i + 1];
        tmp.resize(b - a);
        for (int j = a; j < b; j++) tmp[j - a] = {m.ci[j], m.v[j]};
        std::sort(tmp.begin(), tmp.end(), [](const std::pair<int, double>& x, const std::pair<int, double>& y) { return x.first < y.first; });
        for (int j = a; j < b; j++) { m.ci[j] = tmp[j - a].first; m.v[j] = tmp[j - a].second; }
    }
}

static HostCsr csr_transpose(const HostCsr& m) {
    HostCsr t;
    t.nrow = m.ncol; t.ncol = m.nrow;
    t.rp.assign(t.nrow + 1, 0);
    for (int c : m.ci) t.rp[c + 1]++;
    for (int i = 0; i < t.nrow; i++) t.rp[i + 1] += t.rp[i];
    t.ci.resize(m.nnz()); t.v.resize(m.nnz());
    std::vector<int> pos(t.rp.begin(), t.rp.end() - 1);
    for (int i = 0; i < m.nrow; i++)
        for (int j = m.rp[i]; j < m.rp[i + 1]; j++) { const int q = pos[m.ci[j]]++; t.ci[q] = i; t.v[q] = m.v[j]; }
    return t;                        // rows come out with ascending column ids
}

// C = A * B (Gustavson, dense accumulator); rows of C sorted by column; the sum for one entry runs over A's row order
static HostCsr csr_matmul(const HostCsr& A, const HostCsr& B) {
    HostCsr C;
    C.nrow = A.nrow; C.ncol = B.ncol;
    C.rp.assign(C.nrow + 1, 0);
    std::vector<double> acc(B.ncol, 0.0);
    std::vector<int> mark(B.ncol, -1), cols;
    for (int i = 0; i < A.nrow; i++) {
        cols.clear();
        for (int j = A.rp[i]; j < A.rp[i + 1]; j++) {
            const int k = A.ci[j];
            const double a = A.v[j];
            for (int q = B.rp[k]; q < B.rp[k + 1]; q++) {
                const int c = B.ci[q];
                if (mark[c] != i) { mark[c] = i; acc[c] = 0.0; cols.push_back(c); }
                acc[c] += a * B.v[q];
            }
        }
        std::sort(cols.begin(), cols.end());
        for (int c : cols) { C.ci.push_back(c); C.v.push_back(acc[c]); }
        C.rp[i + 1] = (int)C.ci.size();
    }
    return C;
}

static std::vector<double> csr_diag(const HostCsr& A) {
    std::vector<double> d(A.nrow, 0.0);
    for (int i = 0; i < A.nrow; i++)
        for (int j = A.rp[i]; j < A.rp[i + 1]; j++) if (A.ci[j] == i) d[i] += A.v[j];
    return d;
}

// greedy aggregation (Vanek, Mandel, Brezina) on the strength graph; returns the number of aggregates
static int aggregate(const HostCsr& A, const std::vector<char>& strong, std::vector<int>& agg) {
    const int n = A.nrow;
    agg.assign(n, -1);
    int na = 0;
    for (int i = 0; i < n; i++) {                                  // pass 1: a node whose strong neighbourhood is still free seeds an aggregate
        if (agg[i] >= 0) continue;
        bool any = false, free_ = true;
        for (int j = A.rp[i]; j < A.rp[i + 1]; j++) if (strong[j]) { any = true; if (agg[A.ci[j]] >= 0) { free_ = false; break; } }
        if (!any || !free_) continue;
        agg[i] = na;
        for (int j = A.rp[i]; j < A.rp[i + 1]; j++) if (strong[j]) agg[A.ci[j]] = na;
        na++;
    }
    std::vector<int> join(n, -1);                                  // pass 2: the rest joins a neighbouring pass-1 aggregate
    for (int i = 0; i < n; i++) {
        if (agg[i] >= 0) continue;
        for (int j = A.rp[i]; j < A.rp[i + 1]; j++) if (strong[j] && agg[A.ci[j]] >= 0) { join[i] = agg[A.ci[j]]; break; }
    }
    for (int i = 0; i < n; i++) if (agg[i] < 0 && join[i] >= 0) agg[i] = join[i];
    for (int i = 0; i < n; i++) {                                  // pass 3: leftovers (isolated or surrounded by pass-2 nodes)
        if (agg[i] >= 0) continue;
        agg[i] = na;
        for (int j = A.rp[i]; j < A.rp[i + 1]; j++) if (strong[j] && agg[A.ci[j]] < 0) agg[A.ci[j]] = na;
        na++;
    }
    return na;
}

void amg_host_setup(const HostCsr& A0, const AmgParams& prm, std::vector<AmgHostLevel>& out) {
    out.clear();
    HostCsr A = A0;
    csr_sort_rows(A);
    for (int lev = 0; lev < std::max(1, prm.max_levels); lev++) {
        out.emplace_back();
        AmgHostLevel& L = out.back();
        const int n = A.nrow;
        const std::vector<double> d = csr_diag(A);
        L.dinv.resize(n);
        for (int i = 0; i < n; i++) L.dinv[i] = d[i] != 0.0 ? 1.0 / d[i] : 0.0;
        L.A = std::move(A);
        if (lev == prm.max_levels - 1 || n < prm.min_rows) break;
        const HostCsr& Al = L.A;
        std::vector<char> strong(Al.nnz(), 0);
        for (int i = 0; i < n; i++)
            for (int j = Al.rp[i]; j < Al.rp[i + 1]; j++) {
                const int c = Al.ci[j];
                if (c != i && std::fabs(Al.v[j]) >= prm.theta * std::sqrt(std::fabs(d[i]) * std::fabs(d[c]))) strong[j] = 1;
            }
        std::vector<int> agg;
        const int na = aggregate(Al, strong, agg);
        if (na < 1 || na > 0.8 * n) break;                          // coarsening stalled: this level is the coarsest
        // P = (I - omega_p Df^-1 Af) P_tent,  Af = strong entries + diagonal, weak entries lumped onto the diagonal
        HostCsr P;
        P.nrow = n; P.ncol = na; P.rp.assign(n + 1, 0);
        std::vector<std::pair<int, double>> row;
        for (int i = 0; i < n; i++) {
            double df = d[i];
            for (int j = Al.rp[i]; j < Al.rp[i + 1]; j++) if (Al.ci[j] != i && !strong[j]) df += Al.v[j];
            if (!(std::fabs(df) > 0.1 * std::fabs(d[i]))) df = d[i];
            const double w = df != 0.0 ? prm.omega_p / df : 0.0;
            row.clear();
            auto add = [&](int c, double val) {
                for (auto& e : row) if (e.first == c) { e.second += val; return; }
                row.push_back({c, val});
            };
            add(agg[i], 1.0 - w * df);
            for (int j = Al.rp[i]; j < Al.rp[i + 1]; j++) if (strong[j]) add(agg[Al.ci[j]], -w * Al.v[j]);
            std::sort(row.begin(), row.end(), [](const std::pair<int, double>& x, const std::pair<int, double>& y) { return x.first < y.first; });
            for (auto& e : row) { P.ci.push_back(e.first); P.v.push_back(e.second); }
            P.rp[i + 1] = (int)P.ci.size();
        }
        L.R = csr_transpose(P);
        HostCsr AP = csr_matmul(Al, P);
        A = csr_matmul(L.R, AP);
        L.P = std::move(P);
    }
    if (amg_folded_on()) amg_fold_cycle(out, prm.omega);
}

// the cycle folded into M, T, Q per level (see ssh_amg.cuh)
static void csr_drop_zeros(HostCsr& m) {
    size_t w = 0;
    for (int i = 0; i < m.nrow; i++) {
        const int a = m.rp[i], b = m.rp[i + 1];
        m.rp[i] = (int)w;
        for (int j = a; j < b; j++) if (m.v[j] != 0.0) { m.ci[w] = m.ci[j]; m.v[w] = m.v[j]; w++; }
    }
    m.rp[m.nrow] = (int)w;
    m.ci.resize(w); m.v.resize(w);
}
// I - diag(wl) A diag(wr)   (wl or wr empty = identity)
static HostCsr csr_i_minus_scaled(const HostCsr& A, const std::vector<double>& wl, const std::vector<double>& wr) {
    HostCsr S;
    S.nrow = A.nrow; S.ncol = A.ncol; S.rp.assign(A.nrow + 1, 0);
    S.ci.reserve(A.nnz() + A.nrow); S.v.reserve(A.nnz() + A.nrow);
    for (int i = 0; i < A.nrow; i++) {
        bool diag = false;
        for (int j = A.rp[i]; j < A.rp[i + 1]; j++) {
            const int c = A.ci[j];
            if (!diag && c > i) { S.ci.push_back(i); S.v.push_back(1.0); diag = true; }      // rows are sorted: keep them so
            double v = -A.v[j];
            if (!wl.empty()) v *= wl[i];
            if (!wr.empty()) v *= wr[c];
            if (c == i) { v += 1.0; diag = true; }
            S.ci.push_back(c); S.v.push_back(v);
        }
        if (!diag) { S.ci.push_back(i); S.v.push_back(1.0); }
        S.rp[i + 1] = (int)S.ci.size();
    }
    return S;
}
// A + B, rows of both sorted by column
static HostCsr csr_add(const HostCsr& A, const HostCsr& B) {
    HostCsr C;
    C.nrow = A.nrow; C.ncol = A.ncol; C.rp.assign(A.nrow + 1, 0);
    C.ci.reserve(A.nnz() + B.nnz()); C.v.reserve(A.nnz() + B.nnz());
    for (int i = 0; i < A.nrow; i++) {
        int a = A.rp[i], b = B.rp[i];
        const int ae = A.rp[i + 1], be = B.rp[i + 1];
        while (a < ae || b < be) {
            if (b >= be || (a < ae && A.ci[a] < B.ci[b])) { C.ci.push_back(A.ci[a]); C.v.push_back(A.v[a]); a++; }
            else if (a >= ae || B.ci[b] < A.ci[a]) { C.ci.push_back(B.ci[b]); C.v.push_back(B.v[b]); b++; }
            else { C.ci.push_back(A.ci[a]); C.v.push_back(A.v[a] + B.v[b]); a++; b++; }
        }
        C.rp[i + 1] = (int)C.ci.size();
    }
    return C;
}
void amg_fold_cycle(std::vector<AmgHostLevel>& lv, double omega) {
    const std::vector<double> none;
    for (size_t l = 0; l + 1 < lv.size(); l++) {
        AmgHostLevel& L = lv[l];
        const int n = L.A.nrow;
        std::vector<double> w(n);
        for (int i = 0; i < n; i++) w[i] = omega * L.dinv[i];
        L.M = csr_matmul(L.R, csr_i_minus_scaled(L.A, none, w));
        HostCsr S = csr_i_minus_scaled(L.A, w, none);
        L.Q = csr_matmul(S, L.P);
        if (l + 2 == lv.size()) {                                   // the coarsest level's x = W_c b_c
            const std::vector<double>& dc = lv[l + 1].dinv;
            for (size_t j = 0; j < L.Q.ci.size(); j++) L.Q.v[j] *= omega * dc[L.Q.ci[j]];
        }
        L.T = std::move(S);                                         // (I + S) W
        for (int i = 0; i < n; i++)
            for (int j = L.T.rp[i]; j < L.T.rp[i + 1]; j++) { const int c = L.T.ci[j]; L.T.v[j] = (L.T.v[j] + (c == i ? 1.0 : 0.0)) * w[c]; }
        csr_drop_zeros(L.M); csr_drop_zeros(L.T); csr_drop_zeros(L.Q);
        if (l + 2 == lv.size() && n <= AMG_COLLAPSE_ROWS) {          // x = T b + Q (M b) = C b
            L.T = csr_add(L.T, csr_matmul(L.Q, L.M));
            csr_drop_zeros(L.T);
            L.M = HostCsr(); L.Q = HostCsr();
        }
    }
}
bool amg_folded_on() {
    static int on = -1;
    if (on < 0) { const char* e = getenv("FB200_AMG_FOLD"); on = (e && e[0] == '0') ? 0 : 1; }
    return on == 1;
}

// ---------------------------------------------------------------------------------------------------
// CPU-test access to the host setup (no CUDA call): build a hierarchy from a CSR matrix and read it back
static std::map<int, std::vector<AmgHostLevel>*> g_dbg;
static const HostCsr* dbg_matrix(int handle, int level, int which) {
    auto it = g_dbg.find(handle);
    if (it == g_dbg.end() || level < 0 || level >= (int)it->second->size()) return nullptr;
    const AmgHostLevel& L = (*it->second)[level];
    return which == 0 ? &L.A : which == 1 ? &L.P : which == 2 ? &L.R : which == 3 ? &L.M : which == 4 ? &L.T : &L.Q;
}
extern "C" {
int fb200_amg_host_build(int* n, int* rptr, int* cols, double* vals, double* theta, int* max_levels, int* min_rows) {
    HostCsr A;
    A.nrow = A.ncol = *n;
    A.rp.assign(rptr, rptr + *n + 1);
    A.ci.assign(cols, cols + rptr[*n]);
    A.v.assign(vals, vals + rptr[*n]);
    AmgParams prm;
    prm.theta = *theta; prm.max_levels = *max_levels; prm.min_rows = *min_rows;
    auto* h = new std::vector<AmgHostLevel>();
    amg_host_setup(A, prm, *h);
    int handle = 1;
    while (g_dbg.count(handle)) handle++;
    g_dbg[handle] = h;
    return handle;
}
int fb200_amg_host_levels(int* handle) { auto it = g_dbg.find(*handle); return it == g_dbg.end() ? 0 : (int)it->second->size(); }
// which: 0 A, 1 P, 2 R, 3 M, 4 T, 5 Q (the folded cycle's matrices, amg_fold_cycle)
void fb200_amg_host_dims(int* handle, int* level, int* which, int* nrow, int* ncol, int* nnz) {
    const HostCsr* m = dbg_matrix(*handle, *level, *which);
    *nrow = m ? m->nrow : 0; *ncol = m ? m->ncol : 0; *nnz = m ? (int)m->nnz() : 0;
}
void fb200_amg_host_fetch(int* handle, int* level, int* which, int* rptr, int* cols, double* vals) {
    const HostCsr* m = dbg_matrix(*handle, *level, *which);
    if (!m || m->rp.empty()) return;
    std::copy(m->rp.begin(), m->rp.end(), rptr);
    std::copy(m->ci.begin(), m->ci.end(), cols);
    std::copy(m->v.begin(), m->v.end(), vals);
}
void fb200_amg_host_free(int* handle) { auto it = g_dbg.find(*handle); if (it != g_dbg.end()) { delete it->second; g_dbg.erase(it); } }
}

// ---------------------------------------------------------------------------------------------------
// device side
void DevSell::upload(const HostCsr& m) {
    nrow = m.nrow;
    nslice = (nrow + 31) / 32;
    std::vector<int> sp(nslice + 1, 0);
    long long total = 0;
    for (int s = 0; s < nslice; s++) {
        int w = 0;
        for (int i = s * 32; i < std::min(nrow, s * 32 + 32); i++) w = std::max(w, m.rp[i + 1] - m.rp[i]);
        total += (long long)w * 32;
        if (total >= (1LL << 31)) { fb_set_error(48, "multigrid level too large for 32-bit sliced-ELL offsets"); return; }
        sp[s + 1] = (int)total;
    }
    entries = (size_t)sp[nslice];
    std::vector<int> col(entries, 0);
    std::vector<double> val(entries, 0.0);
    for (int i = 0; i < nrow; i++) {
        const int base = sp[i >> 5] + (i & 31), w = (sp[(i >> 5) + 1] - sp[i >> 5]) >> 5, len = m.rp[i + 1] - m.rp[i];
        const int pad = len > 0 ? m.ci[m.rp[i]] : 0;               // padding entries: value 0, a column the row reads anyway
        for (int k = 0; k < w; k++) {
            col[base + (k << 5)] = k < len ? m.ci[m.rp[i] + k] : pad;
            val[base + (k << 5)] = k < len ? m.v[m.rp[i] + k] : 0.0;
        }
    }
    sptr.upload(sp); scol.upload(col); sval.upload(val);
}

void Amg::upload(const std::vector<AmgHostLevel>& h, double omega_) {
    release();
    omega = omega_;
    folded = h.size() > 1 && h[0].T.nrow > 0;
    L.resize(h.size());
    for (size_t l = 0; l < h.size(); l++) {
        AmgDevLevel& D = L[l];
        D.n = h[l].A.nrow;
        D.dinv.upload(h[l].dinv);
        D.b.alloc(D.n); D.x.alloc(D.n);
        if (folded) {
            if (h[l].T.nrow) D.T.upload(h[l].T);
            if (h[l].M.nrow) { D.M.upload(h[l].M); D.Q.upload(h[l].Q); }          // absent on a collapsed level
            continue;
        }
        D.A.upload(h[l].A);
        if (h[l].P.nrow) { D.P.upload(h[l].P); D.R.upload(h[l].R); }
        D.x2.alloc(D.n); D.r.alloc(D.n);
    }
}

void Amg::release() {
    for (AmgDevLevel& D : L) {
        D.A.release(); D.P.release(); D.R.release(); D.M.release(); D.T.release(); D.Q.release();
        D.dinv.release(); D.b.release(); D.x.release(); D.x2.release(); D.r.release();
    }
    L.clear();
    bar.release();
}

#define ABS 256
// r = b - A x
__global__ void __launch_bounds__(ABS) k_amg_resid(int n, const int* __restrict__ sptr, const int* __restrict__ scol, const double* __restrict__ sval,
                                                   const double* __restrict__ b, const double* __restrict__ x, double* __restrict__ r, const int* done) {
    if (*done) return;
    const int i = blockIdx.x * ABS + threadIdx.x;
    if (i < n) r[i] = b[i] - sell_row(sptr, scol, sval, x, i);
}
// b_c = R r ; x_c = omega dinv_c b_c   (pre-smoothing of the coarse level from a zero guess)
__global__ void __launch_bounds__(ABS) k_amg_restrict(int nc, const int* __restrict__ sptr, const int* __restrict__ scol, const double* __restrict__ sval,
                                                      const double* __restrict__ r, const double* __restrict__ dinv, double omega,
                                                      double* __restrict__ bc, double* __restrict__ xc, const int* done) {
    if (*done) return;
    const int i = blockIdx.x * ABS + threadIdx.x;
    if (i >= nc) return;
    const double v = sell_row(sptr, scol, sval, r, i);
    bc[i] = v;
    xc[i] = omega * dinv[i] * v;
}
// x += P x_c
__global__ void __launch_bounds__(ABS) k_amg_prolong(int n, const int* __restrict__ sptr, const int* __restrict__ scol, const double* __restrict__ sval,
                                                     const double* __restrict__ xc, double* __restrict__ x, const int* done) {
    if (*done) return;
    const int i = blockIdx.x * ABS + threadIdx.x;
    if (i < n) x[i] += sell_row(sptr, scol, sval, xc, i);
}
// out = x + omega dinv (b - A x)   (post-smoothing; out != x)
__global__ void __launch_bounds__(ABS) k_amg_post(int n, const int* __restrict__ sptr, const int* __restrict__ scol, const double* __restrict__ sval,
                                                  const double* __restrict__ b, const double* __restrict__ x, const double* __restrict__ dinv, double omega,
                                                  double* __restrict__ out, const int* done) {
    if (*done) return;
    const int i = blockIdx.x * ABS + threadIdx.x;
    if (i < n) out[i] = x[i] + omega * dinv[i] * (b[i] - sell_row(sptr, scol, sval, x, i));
}

// ---- NQ threads per row.  On the coarse levels (10^5 .. 10^3 rows, 15-30 entries per row) a thread per row is a chain of
// 4-8 dependent (column -> x) load pairs and the kernel runs for ~10 us whatever its size (ncu launch list,
// profiles/r02_ncu_ssh_500k_launches.txt).  Here a block of NQ warps owns one 32-row slice: warp q takes the entries
// k = q, q + NQ, ... of every row (still one coalesced 32-entry line per load), the partial sums meet in shared memory and
// warp 0 adds them in q order — deterministic, and each thread has at most a handful of independent loads in flight.
enum { OP_RESID = 0, OP_RESTRICT = 1, OP_PROLONG = 2, OP_POST = 3 };
template <int OP, int NQ>
__global__ void __launch_bounds__(32 * NQ) k_amg_vec(int n, const int* __restrict__ sptr, const int* __restrict__ scol, const double* __restrict__ sval,
                                                    const double* __restrict__ xin, const double* __restrict__ p1, const double* __restrict__ p2, double omega,
                                                    double* __restrict__ o1, double* __restrict__ o2, const int* done) {
    if (*done) return;
    __shared__ double sh[NQ][32];
    const int lane = threadIdx.x & 31, q = threadIdx.x >> 5;
    const int s0 = sptr[blockIdx.x];
    const int w = (sptr[blockIdx.x + 1] - s0) >> 5;
    const int base = s0 + lane;
    double s = 0.0;
    int k = q;
    for (; k + 3 * NQ < w; k += 4 * NQ) {
        const int o = base + (k << 5);
        const int c0 = scol[o], c1 = scol[o + 32 * NQ], c2 = scol[o + 64 * NQ], c3 = scol[o + 96 * NQ];
        const double a0 = sval[o], a1 = sval[o + 32 * NQ], a2 = sval[o + 64 * NQ], a3 = sval[o + 96 * NQ];
        const double x0 = xin[c0], x1 = xin[c1], x2 = xin[c2], x3 = xin[c3];
        s += a0 * x0; s += a1 * x1; s += a2 * x2; s += a3 * x3;
    }
    for (; k < w; k += NQ) s += sval[base + (k << 5)] * xin[scol[base + (k << 5)]];
    sh[q][lane] = s;
    __syncthreads();
    if (q != 0) return;
    const int i = blockIdx.x * 32 + lane;
    if (i >= n) return;
    double v = sh[0][lane];
#pragma unroll
    for (int j = 1; j < NQ; j++) v += sh[j][lane];
    if (OP == OP_RESID) o1[i] = p1[i] - v;
    else if (OP == OP_RESTRICT) { o1[i] = v; o2[i] = omega * p1[i] * v; }
    else if (OP == OP_PROLONG) o1[i] += v;
    else o1[i] = xin[i] + omega * p2[i] * (p1[i] - v);
}
template <int OP>
static void amg_vec_launch(int nq, int n, const DevSell& M, const double* xin, const double* p1, const double* p2, double omega, double* o1, double* o2,
                           const int* done, cudaStream_t st) {
    const int grid = (n + 31) / 32;
    if (grid == 0) return;
    if (nq == 2) k_amg_vec<OP, 2><<<grid, 64, 0, st>>>(n, M.sptr.p, M.scol.p, M.sval.p, xin, p1, p2, omega, o1, o2, done);
    else if (nq == 4) k_amg_vec<OP, 4><<<grid, 128, 0, st>>>(n, M.sptr.p, M.scol.p, M.sval.p, xin, p1, p2, omega, o1, o2, done);
    else k_amg_vec<OP, 8><<<grid, 256, 0, st>>>(n, M.sptr.p, M.scol.p, M.sval.p, xin, p1, p2, omega, o1, o2, done);
}
// threads per row on level l: FB200_AMG_NQ0 for the fine level (default 1 = the thread-per-row kernels), FB200_AMG_NQ for the others (default 4; 2, 4 and 8 measure the same, profiles/r02_c12_threads_per_row.log)
static int amg_nq(int level) {
    static int nq0 = -1, nqc = -1;
    if (nq0 < 0) {
        auto rd = [](const char* name, int dflt) { const char* e = getenv(name); const int v = e ? atoi(e) : dflt; return (v == 1 || v == 2 || v == 4 || v == 8) ? v : dflt; };
        nq0 = rd("FB200_AMG_NQ0", 1); nqc = rd("FB200_AMG_NQ", 4);
    }
    return level == 0 ? nq0 : nqc;
}

// ---- folded cycle (amg_fold_cycle): one SpMV per level down, one two-matrix SpMV per level up.  Blocks of 8 warps; NQ warps share
// a 32-row slice (warp q takes entries q, q + NQ, ... of every row), 8 / NQ slices per block.
template <int NQ>
__device__ __forceinline__ double sell_part(const int* __restrict__ sptr, const int* __restrict__ scol, const double* __restrict__ sval,
                                            const double* __restrict__ x, int slice, int lane, int q) {
    const int s0 = sptr[slice];
    const int w = (sptr[slice + 1] - s0) >> 5;
    const int base = s0 + lane;
    double s = 0.0;
    int k = q;
    for (; k + 3 * NQ < w; k += 4 * NQ) {
        const int o = base + (k << 5);
        const int c0 = scol[o], c1 = scol[o + 32 * NQ], c2 = scol[o + 64 * NQ], c3 = scol[o + 96 * NQ];
        const double a0 = sval[o], a1 = sval[o + 32 * NQ], a2 = sval[o + 64 * NQ], a3 = sval[o + 96 * NQ];
        const double x0 = x[c0], x1 = x[c1], x2 = x[c2], x3 = x[c3];
        s += a0 * x0; s += a1 * x1; s += a2 * x2; s += a3 * x3;
    }
    for (; k < w; k += NQ) s += sval[base + (k << 5)] * x[scol[base + (k << 5)]];
    return s;
}
// out = M1 x1 (+ M2 x2)
template <int NQ, bool TWO, bool HEAD>
__global__ void __launch_bounds__(256) k_amg_fold(int n, const int* __restrict__ p1, const int* __restrict__ c1, const double* __restrict__ v1, const double* x1,
                                                  const int* __restrict__ p2, const int* __restrict__ c2, const double* __restrict__ v2, const double* __restrict__ x2,
                                                  double* __restrict__ out, const int* done, CycleHead hd) {
    if (*done) return;
    if (HEAD) push_and_wait(hd.pa, x1, hd.pc, hd.sc, 3, hd.mode);       // x1 = b0: own rows out, overlap rows in
    __shared__ double sh[8][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, q = warp % NQ;
    const int slice = blockIdx.x * (8 / NQ) + warp / NQ;
    const int nslice = (n + 31) >> 5;
    double s = 0.0;
    if (slice < nslice) {
        s = sell_part<NQ>(p1, c1, v1, x1, slice, lane, q);
        if (TWO) s += sell_part<NQ>(p2, c2, v2, x2, slice, lane, q);
    }
    if (NQ > 1) {
        sh[warp][lane] = s;
        __syncthreads();
        if (q != 0) return;
#pragma unroll
        for (int j = 1; j < NQ; j++) s += sh[warp + j][lane];
    }
    const int i = slice * 32 + lane;
    if (i < n) out[i] = s;
}
static void amg_fold_launch(int nq, int n, const DevSell& A1, const double* x1, const DevSell* A2, const double* x2, double* out, const int* done, cudaStream_t st,
                            const CycleHead* head = nullptr) {
    if (n <= 0 && !head) return;
    const int nslice = (std::max(n, 0) + 31) / 32;
    CycleHead hd; memset(&hd, 0, sizeof(hd));
    if (head) hd = *head;
#define FOLD_GO(NQ) do { const int g_ = std::max(std::max(1, hd.pa.npb), (nslice + 8 / NQ - 1) / (8 / NQ)); \
        if (A2 && head) k_amg_fold<NQ, true, true><<<g_, 256, 0, st>>>(n, A1.sptr.p, A1.scol.p, A1.sval.p, x1, A2->sptr.p, A2->scol.p, A2->sval.p, x2, out, done, hd); \
        else if (A2) k_amg_fold<NQ, true, false><<<g_, 256, 0, st>>>(n, A1.sptr.p, A1.scol.p, A1.sval.p, x1, A2->sptr.p, A2->scol.p, A2->sval.p, x2, out, done, hd); \
        else if (head) k_amg_fold<NQ, false, true><<<g_, 256, 0, st>>>(n, A1.sptr.p, A1.scol.p, A1.sval.p, x1, nullptr, nullptr, nullptr, nullptr, out, done, hd); \
        else k_amg_fold<NQ, false, false><<<g_, 256, 0, st>>>(n, A1.sptr.p, A1.scol.p, A1.sval.p, x1, nullptr, nullptr, nullptr, nullptr, out, done, hd); } while (0)
    if (nq == 8) FOLD_GO(8); else if (nq == 4) FOLD_GO(4); else if (nq == 2) FOLD_GO(2); else FOLD_GO(1);
#undef FOLD_GO
}

// ---- levels 1 .. K-1 of the V-cycle in ONE launch.  On a partition of a few hundred thousand rows the coarse levels hold
// 10^5, 10^4, 10^3 rows: each of their kernels runs for a few microseconds and the cycle is bound by launch latency (8 coarse
// launches per cycle with four levels, two cycles per BiCGStab iteration).  Here one resident grid (at most one block per SM)
// walks down and up the coarse levels with a grid-wide barrier between the steps.  Same kernels' arithmetic, row by row.
struct AmgLvDesc {
    int n;
    const int *Asp, *Asc, *Psp, *Psc, *Rsp, *Rsc;
    const double *Asv, *Psv, *Rsv, *dinv;
    double *b, *x, *x2, *r;
};
#define AMG_MAX_LEVELS 8
struct AmgFusedArgs { AmgLvDesc lv[AMG_MAX_LEVELS]; int K; double omega; unsigned* bar; };
// sense-reversing grid barrier: bar[0] = arrivals, bar[1] = generation.  All blocks of the launch are resident (grid <= SMs).
__device__ __forceinline__ void grid_barrier(unsigned* bar, unsigned nblocks) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned g = *(volatile unsigned*)(bar + 1);
        if (atomicAdd(bar, 1u) == nblocks - 1) { *(volatile unsigned*)bar = 0u; __threadfence(); atomicAdd(bar + 1, 1u); }
        else while (*(volatile unsigned*)(bar + 1) == g) { }
        __threadfence();
    }
    __syncthreads();
}
#define FBS 1024            // one block of 1024 threads per SM: 151 k threads cover level 1 of a 1 M-row partition with one row each
__global__ void __launch_bounds__(FBS, 1) k_amg_coarse_fused(AmgFusedArgs a, const int* done) {
    if (*done) return;
    const int K = a.K, stride = gridDim.x * FBS, t0 = blockIdx.x * FBS + threadIdx.x;
    for (int l = 1; l + 1 < K; l++) {
        const AmgLvDesc& F = a.lv[l];
        const AmgLvDesc& C = a.lv[l + 1];
        for (int i = t0; i < F.n; i += stride) F.r[i] = F.b[i] - sell_row(F.Asp, F.Asc, F.Asv, F.x, i);          // k_amg_resid
        grid_barrier(a.bar, gridDim.x);
        for (int i = t0; i < C.n; i += stride) {                                                                  // k_amg_restrict
            const double v = sell_row(F.Rsp, F.Rsc, F.Rsv, F.r, i);
            C.b[i] = v; C.x[i] = a.omega * C.dinv[i] * v;
        }
        grid_barrier(a.bar, gridDim.x);
    }
    const double* cur = a.lv[K - 1].x;
    for (int l = K - 2; l >= 1; l--) {
        const AmgLvDesc& F = a.lv[l];
        for (int i = t0; i < F.n; i += stride) F.x[i] += sell_row(F.Psp, F.Psc, F.Psv, cur, i);                   // k_amg_prolong
        grid_barrier(a.bar, gridDim.x);
        for (int i = t0; i < F.n; i += stride)                                                                    // k_amg_post
            F.x2[i] = F.x[i] + a.omega * F.dinv[i] * (F.b[i] - sell_row(F.Asp, F.Asc, F.Asv, F.x, i));
        if (l > 1) grid_barrier(a.bar, gridDim.x);
        cur = F.x2;
    }
}

static bool amg_fused_on() {
#ifdef FB200_EMULATION
    return false;                                   // the emulation runs the blocks of a launch one after another: no grid barrier
#else
    static int on = -1;
    // measured (profiles/r02_bench_1m_2gpu_*.json): 72-84 us per fused launch against 8 x 11 us for the separate kernels — the seven
    // grid barriers cost what the launches cost; off unless FB200_AMG_FUSED=1
    if (on < 0) { const char* e = getenv("FB200_AMG_FUSED"); on = (e && e[0] == '1') ? 1 : 0; }
    return on == 1;
#endif
}

// n_out: rows of the last (fine-level) post-smoothing step that are written to `out` — all of level 0 on one rank, the rank's own
// rows when level 0 is an overlapping subdomain (restricted additive Schwarz: the overlap part of the result is discarded)
void Amg::cycle(double* out, const int* done, cudaStream_t st, int n_out, const CycleHead* head) {
    const int K = (int)L.size();
    if (K <= 1) return;                                           // single level: the caller's Jacobi step already wrote `out`
#define AKL(name0, namec, lev, ...) do { ProfScope ps_((lev) == 0 ? name0 : namec, st); __VA_ARGS__; LAUNCH_CHECK(); } while (0)
    if (folded) {
        const bool collapsed = L[K - 2].M.nrow == 0;              // level K-2: x = C b in one step (C kept in T)
        const int ndown = collapsed ? K - 2 : K - 1;
        for (int l = 0; l < ndown; l++) {                         // b_{l+1} = M_l b_l
            AKL("k_amg_down0", "k_amg_coarse", l, amg_fold_launch(amg_nq(l + 1), L[l + 1].n, L[l].M, L[l].b.p, nullptr, nullptr, L[l + 1].b.p, done, st, l == 0 ? head : nullptr));
        }
        const double* cur = L[K - 1].b.p;                         // the coarsest level's W is inside Q_{K-2}
        for (int l = K - 2; l >= 0; l--) {                        // x_l = T_l b_l + Q_l x_{l+1}
            double* dst = l == 0 ? out : L[l].x.p;
            const int rows = (l == 0 && n_out > 0) ? std::min(n_out, L[l].n) : L[l].n;
            const bool one = collapsed && l == K - 2;
            AKL("k_amg_up0", "k_amg_coarse", l, amg_fold_launch(amg_nq(one ? 1 : l), rows, L[l].T, L[l].b.p, one ? nullptr : &L[l].Q, cur, dst, done, st,
                                                               (one && l == 0) ? head : nullptr));
            cur = dst;
        }
        return;
    }
    const bool fused = K >= 3 && K <= AMG_MAX_LEVELS && amg_fused_on();
    if (fused && !bar.p) bar.alloc(2);
    for (int l = 0; l + 1 < K; l++) {
        if (fused && l >= 1) break;
        AmgDevLevel& F = L[l];
        AmgDevLevel& C = L[l + 1];
        // the restriction's rows are the coarse level's: its threads-per-row follow the level it writes
        const int nq = amg_nq(l), nqr = amg_nq(l + 1);
        if (nq > 1) AKL("k_amg_resid0", "k_amg_coarse", l, amg_vec_launch<OP_RESID>(nq, F.n, F.A, F.x.p, F.b.p, nullptr, omega, F.r.p, nullptr, done, st));
        else AKL("k_amg_resid0", "k_amg_coarse", l, k_amg_resid<<<(F.n + ABS - 1) / ABS, ABS, 0, st>>>(F.n, F.A.sptr.p, F.A.scol.p, F.A.sval.p, F.b.p, F.x.p, F.r.p, done));
        if (nqr > 1) AKL("k_amg_restrict0", "k_amg_coarse", l, amg_vec_launch<OP_RESTRICT>(nqr, C.n, F.R, F.r.p, C.dinv.p, nullptr, omega, C.b.p, C.x.p, done, st));
        else AKL("k_amg_restrict0", "k_amg_coarse", l, k_amg_restrict<<<(C.n + ABS - 1) / ABS, ABS, 0, st>>>(C.n, F.R.sptr.p, F.R.scol.p, F.R.sval.p, F.r.p, C.dinv.p, omega, C.b.p, C.x.p, done));
    }
    if (fused) {
        AmgFusedArgs a;
        a.K = K; a.omega = omega; a.bar = bar.p;
        for (int l = 0; l < K; l++) {
            AmgDevLevel& D = L[l];
            AmgLvDesc& d = a.lv[l];
            d.n = D.n; d.Asp = D.A.sptr.p; d.Asc = D.A.scol.p; d.Asv = D.A.sval.p; d.Psp = D.P.sptr.p; d.Psc = D.P.scol.p; d.Psv = D.P.sval.p;
            d.Rsp = D.R.sptr.p; d.Rsc = D.R.scol.p; d.Rsv = D.R.sval.p; d.dinv = D.dinv.p; d.b = D.b.p; d.x = D.x.p; d.x2 = D.x2.p; d.r = D.r.p;
        }
        int sms = 148;
        { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); }
        const int grid = std::max(1, std::min(sms, (L[1].n + FBS - 1) / FBS));      // every block resident: the barrier needs it
        ProfScope ps_("k_amg_coarse_fused", st);
        k_amg_coarse_fused<<<grid, FBS, 0, st>>>(a, done);
        LAUNCH_CHECK();
    }
    const double* cur = fused ? L[1].x2.p : L[K - 1].x.p;
    for (int l = fused ? 0 : K - 2; l >= 0; l--) {
        AmgDevLevel& F = L[l];
        double* dst = l == 0 ? out : F.x2.p;
        const int nq = amg_nq(l);
        if (nq > 1) AKL("k_amg_prolong0", "k_amg_coarse", l, amg_vec_launch<OP_PROLONG>(nq, F.n, F.P, cur, nullptr, nullptr, omega, F.x.p, nullptr, done, st));
        else AKL("k_amg_prolong0", "k_amg_coarse", l, k_amg_prolong<<<(F.n + ABS - 1) / ABS, ABS, 0, st>>>(F.n, F.P.sptr.p, F.P.scol.p, F.P.sval.p, cur, F.x.p, done));
        const int npost = (l == 0 && n_out > 0) ? std::min(n_out, F.n) : F.n;
        if (nq > 1) AKL("k_amg_post0", "k_amg_coarse", l, amg_vec_launch<OP_POST>(nq, npost, F.A, F.x.p, F.b.p, F.dinv.p, omega, dst, nullptr, done, st));
        else AKL("k_amg_post0", "k_amg_coarse", l, k_amg_post<<<(npost + ABS - 1) / ABS, ABS, 0, st>>>(npost, F.A.sptr.p, F.A.scol.p, F.A.sval.p, F.b.p, F.x.p, F.dinv.p, omega, dst, done));
        cur = dst;
    }
#undef AKL
}
