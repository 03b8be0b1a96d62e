// Context of the tracer path: error status, building the device-side mesh tables from the Fortran arrays
// (fb200_init_* -> fb200_init_commit), and restaging between the reference layout and the tiled layout.
#include "common.cuh"
#include "comm.cuh"
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <mutex>

long long g_launches = 0;
static int g_status = 0;
static std::string g_status_msg;
static std::mutex g_mu;

void fb_set_error(int code, const std::string& msg) {
    std::lock_guard<std::mutex> lk(g_mu);
    fprintf(stderr, "fesom_b200: ERROR %d: %s\n", code, msg.c_str());
    if (g_status == 0) { g_status = code; g_status_msg = msg; }
    const char* ab = getenv("FB200_ABORT_ON_ERROR");
    if (ab && ab[0] == '1') abort();
}
bool fb_ok() { return g_status == 0; }
int fb_status_code() { return g_status; }
const char* fb_status_msg() { return g_status_msg.c_str(); }
void fb_clear_status() { std::lock_guard<std::mutex> lk(g_mu); g_status = 0; g_status_msg.clear(); }

TracerCtx& tctx() { static TracerCtx c; return c; }

// ---------------------------------------------------------------------------------------------------
static cudaStream_t g_stream = nullptr;
static bool g_stream_owned = false;
cudaStream_t fb_stream() {
    if (!g_stream) { CUDA_CHECK(cudaStreamCreateWithFlags(&g_stream, cudaStreamNonBlocking)); g_stream_owned = true; }
    return g_stream;
}
void fb_set_stream(cudaStream_t s) {
    if (g_stream && g_stream_owned) { cudaStreamSynchronize(g_stream); cudaStreamDestroy(g_stream); }
    g_stream = s; g_stream_owned = false;
    if (!s) g_stream = nullptr;               // next fb_stream() creates a private one again
}

// per-kernel timing ----------------------------------------------------------------------------------
#include <map>
struct ProfRec { std::string name; cudaEvent_t a, b; };
static bool g_prof = false;
static std::vector<ProfRec> g_prof_open, g_prof_done;
static std::vector<cudaEvent_t> g_prof_pool;
struct ProfAcc { long long launches = 0; double ms = 0.0; };
static std::map<std::string, ProfAcc> g_prof_acc;
void prof_enable(bool on) { g_prof = on; }
bool prof_on() { return g_prof; }
static cudaEvent_t prof_event() {
    if (!g_prof_pool.empty()) { cudaEvent_t e = g_prof_pool.back(); g_prof_pool.pop_back(); return e; }
    cudaEvent_t e; cudaEventCreate(&e); return e;
}
void prof_push(const char* name, cudaStream_t st) {
    ProfRec r; r.name = name; r.a = prof_event(); r.b = prof_event();
    cudaEventRecord(r.a, st);
    g_prof_open.push_back(r);
}
void prof_pop(cudaStream_t st) {
    if (g_prof_open.empty()) return;
    ProfRec r = g_prof_open.back(); g_prof_open.pop_back();
    cudaEventRecord(r.b, st);
    g_prof_done.push_back(r);
}
void prof_collect() {
    for (ProfRec& r : g_prof_done) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) { ProfAcc& a = g_prof_acc[r.name]; a.launches++; a.ms += ms; }
        g_prof_pool.push_back(r.a); g_prof_pool.push_back(r.b);
    }
    g_prof_done.clear();
}
void prof_reset() { prof_collect(); g_prof_acc.clear(); }
int prof_report(char* buf, int cap) {          // "name launches ms\n" per kernel
    prof_collect();
    std::string s;
    char line[256];
    for (auto& kv : g_prof_acc) { snprintf(line, sizeof(line), "%s %lld %.6f\n", kv.first.c_str(), kv.second.launches, kv.second.ms); s += line; }
    if ((int)s.size() + 1 > cap) return -1;
    memcpy(buf, s.c_str(), s.size() + 1);
    return (int)s.size();
}

static cudaStream_t g_copy_stream = nullptr;
cudaStream_t fb_copy_stream() {
    if (!g_copy_stream) CUDA_CHECK(cudaStreamCreateWithFlags(&g_copy_stream, cudaStreamNonBlocking));
    return g_copy_stream;
}
void fb_copy_sync() { if (g_copy_stream) CUDA_CHECK(cudaStreamSynchronize(g_copy_stream)); }

// ---------------------------------------------------------------------------------------------------
// restage kernels.  One block per tile of 32 entities.  src is the reference array (ncomp, lev_ref, ent).
__global__ void __launch_bounds__(256) k_ref_to_tiled(const double* __restrict__ src, double* __restrict__ dst, int count_ent,
                                                      int lev_ref, int nlp, int ncomp, int comp, int dncomp, int dcomp) {
    extern __shared__ double sm[];                 // [nlp][33]
    const int tile = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int j = warp; j < TILE; j += 8) {
        const int ent = tile * TILE + j;
        for (int k = lane; k < nlp; k += 32) {
            double v = 0.0;
            if (ent < count_ent && k < lev_ref) v = src[((size_t)ent * lev_ref + k) * ncomp + comp];
            sm[k * 33 + j] = v;
        }
    }
    __syncthreads();
    double* out = dst + (size_t)tile * nlp * TILE * dncomp + dcomp;
    for (int idx = threadIdx.x; idx < nlp * TILE; idx += 256) out[(size_t)idx * dncomp] = sm[(idx >> 5) * 33 + (idx & 31)];
}
__global__ void __launch_bounds__(256) k_tiled_to_ref(const double* __restrict__ src, double* __restrict__ dst, int count_ent,
                                                      int lev_ref, int nlp, int dncomp, int dcomp) {
    extern __shared__ double sm[];
    const int tile = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const double* in = src + (size_t)tile * nlp * TILE * dncomp + dcomp;
    for (int idx = threadIdx.x; idx < nlp * TILE; idx += 256) sm[(idx >> 5) * 33 + (idx & 31)] = in[(size_t)idx * dncomp];
    __syncthreads();
    for (int j = warp; j < TILE; j += 8) {
        const int ent = tile * TILE + j;
        if (ent >= count_ent) continue;
        for (int k = lane; k < lev_ref; k += 32) dst[(size_t)ent * lev_ref + k] = sm[k * 33 + j];
    }
}

static void ensure_stage(TracerCtx& c, size_t count) { if (c.stage.n < count) c.stage.alloc(count); }

// host (reference layout, (ncomp, lev_ref, count_ent)) -> device tiled field of component `comp`
void ctx_push_node_field(TracerCtx& c, const double* host, int lev_ref, int count_ent, double* dev_tiled, int ncomp, int comp, int dncomp, int dcomp) {
    if (!fb_ok() || count_ent == 0) return;
    const size_t cnt = (size_t)ncomp * lev_ref * count_ent;
    if (comp == 0) {                                   // components of one array are pushed back to back: upload once
        ensure_stage(c, cnt);
        CUDA_CHECK(cudaMemcpyAsync(c.stage.p, host, cnt * sizeof(double), cudaMemcpyHostToDevice, c.stream));
    }
    const int tiles = (count_ent + TILE - 1) / TILE;
    k_ref_to_tiled<<<tiles, 256, (size_t)c.nlp * 33 * sizeof(double), c.stream>>>(c.stage.p, dev_tiled, count_ent, lev_ref, c.nlp, ncomp, comp, dncomp, dcomp);
    LAUNCH_CHECK();
}
void ctx_pull_node_field(TracerCtx& c, const double* dev_tiled, int lev_ref, int count_ent, double* host, int dncomp, int dcomp, bool sync) {
    if (!fb_ok() || count_ent == 0) return;
    const size_t cnt = (size_t)lev_ref * count_ent;
    ensure_stage(c, cnt);
    const int tiles = (count_ent + TILE - 1) / TILE;
    k_tiled_to_ref<<<tiles, 256, (size_t)c.nlp * 33 * sizeof(double), c.stream>>>(dev_tiled, c.stage.p, count_ent, lev_ref, c.nlp, dncomp, dcomp);
    LAUNCH_CHECK();
    CUDA_CHECK(cudaMemcpyAsync(host, c.stage.p, cnt * sizeof(double), cudaMemcpyDeviceToHost, c.stream));
    if (sync) CUDA_CHECK(cudaStreamSynchronize(c.stream));
}

// ---------------------------------------------------------------------------------------------------
// fb200_init_commit: turn the Fortran mesh arrays into device tables.
static inline unsigned pack4(int a, int b, int c, int d) { return (unsigned)a | ((unsigned)b << 8) | ((unsigned)c << 16) | ((unsigned)d << 24); }

void ctx_commit(TracerCtx& c) {
    if (!fb_ok()) return;
    if (c.nl <= 2 || c.nl > 255) { fb_set_error(10, "nl must be in 3..255"); return; }
    if (c.h_edges.empty() || c.h_nlev.empty() || c.h_ecd.empty()) { fb_set_error(11, "fb200_init_commit before connectivity/levels/geometry"); return; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { fb_set_error(12, "no CUDA device: libfesom_b200 has no CPU fallback"); return; }
    c.stream = fb_stream();
    if (!c.ev0) { CUDA_CHECK(cudaEventCreate(&c.ev0)); CUDA_CHECK(cudaEventCreate(&c.ev1)); }
    c.nlp = c.nl;
    const int nlp = c.nlp, N = c.N, myN = c.myN, Eall = c.Eall, myE = c.myE, myEd = c.myEd;
    c.Ntiles = (N + TILE - 1) / TILE; c.myNtiles = (myN + TILE - 1) / TILE;
    c.Etiles = (Eall + TILE - 1) / TILE; c.Edtiles = (myEd + TILE - 1) / TILE;
    if ((long long)std::max(std::max(c.Ntiles, c.Etiles), c.Edtiles + 1) * nlp * TILE >= (1LL << 31)) { fb_set_error(13, "mesh too large for 32-bit tile offsets"); return; }
    const double r_earth = 6367500.0;
    // elements
    std::vector<int> el_nb((size_t)3 * Eall, 0); std::vector<unsigned> el_lev(Eall);
    for (int e = 0; e < Eall; e++) {
        for (int k = 0; k < 3; k++) { int n = c.h_elem_nodes[(size_t)e * 3 + k]; el_nb[(size_t)k * Eall + e] = n >= 1 ? tile_base(n - 1, nlp) : -1; }
        el_lev[e] = pack4(c.h_ulev[e], c.h_nlev[e] - 1, 0, 0);
    }
    c.el_nb.upload(el_nb); c.el_lev.upload(el_lev);
    std::vector<double> g((size_t)6 * myE);
    for (int e = 0; e < myE; e++) for (int k = 0; k < 6; k++) g[(size_t)k * myE + e] = c.h_gsca[(size_t)e * 6 + k];
    c.el_g.upload(g); c.el_area.upload(c.h_earea);
    // nodes
    std::vector<unsigned> n_lev(N);
    for (int n = 0; n < N; n++) n_lev[n] = pack4(c.h_ulev_n[n], c.h_nlev_n[n], c.h_nlmin[n], c.h_ulmax[n]);
    c.n_lev.upload(n_lev); c.n_nbl.upload(c.h_nbl); c.n_znb.upload(c.h_znb);
    // edges.  Internal numbering: the reference's (identity), or with FB200_EDGE_ORDER=tile an experimental one in which the
    // edges that share a FIRST node's rank k (k-th edge of that node in reference order) are consecutive over the 32 nodes of
    // a node tile — sorted by (tile of n1, k, lane of n1) — so that slot j of 32 consecutive nodes gathers (nearly)
    // consecutive edge-field entries instead of entries 3 apart (profiles/r01_gather_coalescing.txt).  Only the storage
    // order of the edge-indexed tables and fields changes: every per-node sum still runs in ascending REFERENCE edge id, so
    // results are bit-identical (tests/test_emulated.py runs both orders).
    std::vector<int>& perm = c.edge_perm;
    perm.resize(myEd);
    for (int e = 0; e < myEd; e++) perm[e] = e;
    {
        const char* eo = getenv("FB200_EDGE_ORDER");
        if (eo && !strcmp(eo, "tile")) {
            std::vector<int> rank(myEd), cnt(N, 0), order(myEd);
            for (int e = 0; e < myEd; e++) { const int n1 = c.h_edges[2 * (size_t)e] - 1; rank[e] = (n1 >= 0 && n1 < N) ? cnt[n1]++ : 0; order[e] = e; }
            std::stable_sort(order.begin(), order.end(), [&](int a, int b) {
                const int na = c.h_edges[2 * (size_t)a] - 1, nb = c.h_edges[2 * (size_t)b] - 1;
                if ((na >> 5) != (nb >> 5)) return (na >> 5) < (nb >> 5);
                if (rank[a] != rank[b]) return rank[a] < rank[b];
                return (na & 31) < (nb & 31);
            });
            for (int k = 0; k < myEd; k++) perm[order[k]] = k;
        }
    }
    std::vector<EdgeRec> er(myEd); std::vector<EdgeHo> eh(myEd);
    std::vector<int> deg(N, 0);
    for (int e = 0; e < myEd; e++) {
        const int n1 = c.h_edges[2 * (size_t)e], n2 = c.h_edges[2 * (size_t)e + 1];
        const int t1 = c.h_edge_tri[2 * (size_t)e], t2 = c.h_edge_tri[2 * (size_t)e + 1];
        if (n1 < 1 || n2 < 1 || n1 > N || n2 > N || t1 < 1 || t1 > Eall || t2 > Eall) { fb_set_error(14, "edge table out of range"); return; }
        EdgeRec& r = er[perm[e]]; memset(&r, 0, sizeof(r));
        r.bn1 = tile_base(n1 - 1, nlp); r.bn2 = tile_base(n2 - 1, nlp);
        r.be1 = tile_base(t1 - 1, nlp); r.be2 = t2 > 0 ? tile_base(t2 - 1, nlp) : -1;
        const int nu1 = c.h_ulev[t1 - 1], nl1 = c.h_nlev[t1 - 1] - 1;
        const int nu2 = t2 > 0 ? c.h_ulev[t2 - 1] : 0, nl2 = t2 > 0 ? c.h_nlev[t2 - 1] - 1 : 0;
        if (nlp > 63) { fb_set_error(16, "nl > 63 is not supported (64-bit level masks)"); return; }
        r.dX1 = c.h_ecd[4 * (size_t)e]; r.dY1 = c.h_ecd[4 * (size_t)e + 1]; r.dX2 = c.h_ecd[4 * (size_t)e + 2]; r.dY2 = c.h_ecd[4 * (size_t)e + 3];
        // the five level ranges of oce_adv_tra_hor.F90:129-209 must not overlap (they never do on a valid mesh);
        // the kernels evaluate them as a partition
        {
            const int nl12 = std::min(nl1, nl2), nu12 = std::max(nu1, nu2);
            for (int nz = 1; nz <= nlp; nz++) {
                const bool inA = nz >= nu1 && nz <= nu12 - 1, inB = nu2 > 0 && nz >= nu2 && nz <= nu12 - 1;
                // (D)/(E) also require the element to be wet at nz: on a boundary edge (nl12 = 0) the reference's loop (D) starts at
                // level 1 even when the element begins below an ice-shelf cavity; UV and helem are zero up there, so those
                // fluxes are exact zeros — dropping them changes no sum and keeps "mask bit = a wet element joins the two nodes"
                const bool inC = nz >= nu12 && nz <= nl12, inD = nz >= nl12 + 1 && nz <= nl1 && nz >= nu1, inE = nz >= nl12 + 1 && nz <= nl2 && nz >= nu2;
                if ((int)inA + inB + inC + inD + inE > 1) { fb_set_error(15, "edge with overlapping level ranges (disjoint element columns) is not supported"); return; }
                if (inA || inC || inD) r.m1 |= 1ull << nz;
                if (inB || inC || inE) r.m2 |= 1ull << nz;
            }
        }
        EdgeHo& h = eh[perm[e]]; memset(&h, 0, sizeof(h));
        double a = r_earth * c.h_ecos[t1 - 1];                       // oce_adv_tra_hor.F90:535,548
        if (t2 > 0) a = 0.5 * (a + r_earth * c.h_ecos[t2 - 1]);
        h.dxa = c.h_edxdy[2 * (size_t)e] * a; h.dyr = c.h_edxdy[2 * (size_t)e + 1] * r_earth;
        const int up = c.h_updn[2 * (size_t)e], dn = c.h_updn[2 * (size_t)e + 1];
        h.bup = up > 0 ? tile_base(up - 1, nlp) : -1; h.bdn = dn > 0 ? tile_base(dn - 1, nlp) : -1;
        if (up > 0 && dn > 0) {                                      // oce_muscl_adv.F90:306-308
            int nzmin = std::max(c.h_ulmax[n1 - 1], c.h_ulmax[n2 - 1]), nzmax = std::min(c.h_nlmin[n1 - 1], c.h_nlmin[n2 - 1]);
            h.shared = pack4(nzmin, nzmax, 1, 0);
        } else h.shared = 0;
        h.n1 = n1 - 1; h.n2 = n2 - 1;
        deg[n1 - 1]++; deg[n2 - 1]++;
    }
    c.edge.upload(er); c.edgeho.upload(eh);
    {   // levels of fill_up_dn_grad's fallback branch (oce_muscl_adv.F90:313-328, 372-440): edge without both up/downwind
        // triangles, or level outside the range they share
        std::vector<int2> fb;
        for (int e = 0; e < myEd; e++) {
            const unsigned long long many = er[e].m1 | er[e].m2;
            const int lo = eh[e].shared & 255, hi = (eh[e].shared >> 8) & 255;
            const bool good = (eh[e].shared >> 16) & 1;
            for (int nz = 1; nz <= nlp && nz < 64; nz++)
                if (((many >> nz) & 1ull) && !(good && nz >= lo && nz <= hi - 1)) fb.push_back(make_int2(e, nz));
        }
        c.n_fb = (int)fb.size();
        if (c.n_fb) c.fb_list.upload(fb);
    }
    c.maxdeg = 0; for (int n = 0; n < N; n++) c.maxdeg = std::max(c.maxdeg, deg[n]);
    {
        std::vector<int> slot((size_t)c.Ntiles * c.maxdeg * TILE, -1), fill(N, 0);
        for (int e = 0; e < myEd; e++)            // ascending edge id == the reference's scatter order
            for (int s = 0; s < 2; s++) {
                int n = c.h_edges[2 * (size_t)e + s] - 1;
                slot[((size_t)(n >> 5) * c.maxdeg + fill[n]) * TILE + (n & 31)] = (perm[e] << 1) | s;
                fill[n]++;
            }
        c.ne_slot.upload(slot);
        const size_t ns = slot.size();
        std::vector<int4> ni(ns, make_int4(0, 0, -1, 0));
        std::vector<ulonglong2> mk(ns, make_ulonglong2(0ull, 0ull));
        std::vector<double2> d1(ns, make_double2(0., 0.)), d2(ns, make_double2(0., 0.));
        for (size_t k = 0; k < ns; k++) {
            if (slot[k] < 0) continue;
            const int e = slot[k] >> 1, second = slot[k] & 1;
            const EdgeRec& r = er[e];
            ni[k] = make_int4(second ? r.bn1 : r.bn2, r.be1, r.be2, tile_base(e, nlp)); mk[k] = make_ulonglong2(r.m1, r.m2);
            d1[k] = make_double2(r.dX1, r.dY1); d2[k] = make_double2(r.dX2, r.dY2);
        }
        c.ns_i.upload(ni); c.ns_m.upload(mk);
        {
            std::vector<SlotRec> sp(ns);
            const int zero_edge = c.Edtiles * nlp * TILE;      // first entry of the never-written extra tile of the edge fields
            for (size_t k = 0; k < ns; k++) {
                sp[k].bo = ni[k].x; sp[k].bed = ni[k].w;
                sp[k].mk = slot[k] < 0 ? 0ull : ((mk[k].x | mk[k].y) & ~1ull) | (unsigned long long)(slot[k] & 1);
                if (slot[k] < 0) {                              // empty slot: the node's own column and a zero edge entry (tracer_kernels2.cu, SlotBatch)
                    const int node = (int)(k / ((size_t)c.maxdeg * TILE)) * TILE + (int)(k % TILE);
                    sp[k].bo = tile_base(node, nlp); sp[k].bed = zero_edge;
                }
            }
            c.ns_p.upload(sp);
        }
        std::vector<int> tdeg(c.Ntiles, 0);
        for (int n = 0; n < N; n++) tdeg[n >> 5] = std::max(tdeg[n >> 5], deg[n]);
        c.tile_deg.upload(tdeg);
        c.ns_d1.upload(d1); c.ns_d2.upload(d2);
    }
    c.maxnie = 0; for (int n = 0; n < N; n++) c.maxnie = std::max(c.maxnie, c.h_nie_num[n]);
    {
        std::vector<int> slot((size_t)c.Ntiles * c.maxnie * TILE, -1);
        for (int n = 0; n < N; n++)
            for (int k = 0; k < c.h_nie_num[n]; k++) {
                int e = c.h_nie[(size_t)n * c.nie_ld + k];
                slot[((size_t)(n >> 5) * c.maxnie + k) * TILE + (n & 31)] = e >= 1 ? e - 1 : -1;
            }
        c.nie_slot.upload(slot);
    }
    // fields
    const size_t nf = (size_t)c.Ntiles * nlp * TILE, ef = (size_t)c.Etiles * nlp * TILE, df = (size_t)(c.Edtiles + 1) * nlp * TILE;   // edge fields: + 1 never-written tile (tracer_kernels2.cu)
    for (DevBuf<double>* b : {&c.area, &c.areasvol, &c.W, &c.We, &c.Wi, &c.hnode, &c.hnode_new, &c.Z3, &c.zbar3, &c.Kv, &c.Ki,
                              &c.sl1, &c.sl2, &c.sl3, &c.sw3d}) b->alloc(nf);
    // work fields: one set per tracer of a pass (the column kernels take two tracers per pass)
    c.nwork = c.ntr >= 2 ? 2 : 1;
    const size_t nw = (size_t)c.nwork;
    for (DevBuf<double>* b : {&c.trz, &c.LO, &c.adf_v, &c.tstar}) b->alloc(nf * nw);
    c.fpm.alloc(nf * nw);
    c.helem.alloc(ef); c.UV.alloc(ef); c.txy.alloc(ef * nw);
    c.adf_h.alloc(df * nw); c.eflx.alloc(df * nw);
    c.T.alloc(nf * c.ntr); c.Told.alloc(nf * c.ntr);
    for (DevBuf<double>* b : {&c.heat_flux, &c.water_flux, &c.virtual_salt, &c.relax_salt, &c.real_salt_flux}) b->alloc(N);
    ctx_push_node_field(c, c.h_area.data(), c.nl, N, c.area.p);
    ctx_push_node_field(c, c.h_areasvol.data(), c.nl, N, c.areasvol.p);
    CUDA_CHECK(cudaStreamSynchronize(c.stream));
    if (fb_ok()) c.committed = true;
}

void ctx_release(TracerCtx& c) {
    if (c.stream) { cudaStreamSynchronize(c.stream); }
    c.which_ale = -1;                                   // the device-side ALE producers need a new fb200_ale_init
    for (DevBuf<double>* b : {&c.area, &c.areasvol, &c.W, &c.We, &c.Wi, &c.hnode, &c.hnode_new, &c.Z3, &c.zbar3, &c.Kv, &c.Ki,
                              &c.sl1, &c.sl2, &c.sl3, &c.sw3d, &c.trz, &c.LO, &c.adf_v, &c.tstar,
                              &c.helem, &c.adf_h, &c.eflx, &c.T, &c.Told, &c.heat_flux, &c.water_flux, &c.virtual_salt,
                              &c.relax_salt, &c.real_salt_flux, &c.relax2clim, &c.Tclim, &c.Sclim, &c.zon_ztem, &c.zon_a, &c.kgT, &c.kgS, &c.kb2, &c.kb3, &c.hbar, &c.hbar_old, &c.ssh_rhs_old, &c.dhe, &c.esum, &c.CFL, &c.stage, &c.el_g, &c.el_area, &c.n_znb, &c.redi_v, &c.bh_tmp}) b->release();
    c.tile_deg.release(); c.fb_list.release(); c.n_fb = 0; c.ns_p.release(); c.ns_i.release(); c.ns_m.release(); c.ns_d1.release(); c.ns_d2.release();
    c.UV.release(); c.txy.release(); c.fpm.release(); c.extrema_part.release(); c.zon_nn.release();
    for (int k = 0; k < 8; k++) { c.mean8[k].release(); c.mean4[k].release(); }
    c.sa_rowptr.release(); c.sa_colind.release(); c.sa_values.release(); c.sa_rhs.release(); c.sa_eta.release(); c.UVr.release();
    c.edge.release(); c.edgeho.release(); c.ne_slot.release(); c.nie_slot.release(); c.el_nb.release(); c.el_lev.release();
    c.n_lev.release(); c.n_nbl.release();
    if (c.halo_nod) { c.halo_nod->release(); delete c.halo_nod; }
    if (c.halo_elem) { c.halo_elem->release(); delete c.halo_elem; }
    if (c.ev0) cudaEventDestroy(c.ev0);
    if (c.ev1) cudaEventDestroy(c.ev1);
    c = TracerCtx();
}
