// Shared declarations of libfesom_b200.so: device data layout, context, error plumbing.
//
// DATA LAYOUT IN HBM ("level-blocked"): every (level, entity) field — entity = node, element or edge — is
// stored in tiles of TILE = 32 consecutive entities; inside a tile the level index is the slow one and the
// entity the fast one:   offset(ent, k) = (ent / 32) * (NLP * 32) + k * 32 + (ent % 32),  k = level-1, NLP = nl.
// A warp that owns 32 consecutive entities therefore reads one fully coalesced 256-byte line per level and
// walks through ONE contiguous NLP*256-byte region per field, and the reference's (level, entity) column-major
// block of the same 32 entities is contiguous too, so restaging is a tile-local transpose.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

#define TILE 32

struct __align__(16) EdgeRec {      // one record per local edge, read by both endpoint nodes (64 B)
    int bn1, bn2;                   // tile bases of the two nodes          (see tile_base)
    int be1, be2;                   // tile bases of the two elements, be2 = -1 on a boundary edge
    unsigned long long m1, m2;      // level masks: bit nz set where the segment in element 1 / 2 carries a flux at level nz
    double dX1, dY1, dX2, dY2;      // edge_cross_dxdy(1:4)
};
// per (node, slot) record of the node kernels: tile bases of the other node and of the edge, and the level mask
// m1|m2 of the edge (bit nz set where the edge carries a flux at level nz; levels start at 1) with bit 0 = "this
// node is the edge's second node" (the flux enters with a minus sign)
struct __align__(16) SlotRec { int bo, bed; unsigned long long mk; };
struct __align__(16) EdgeHo {       // extra per-edge data of the high-order reconstruction (48 B)
    double dxa, dyr;                // edge_dxdy(1)*a, edge_dxdy(2)*r_earth  (a = r_earth * mean elem_cos)
    int bup, bdn;                   // tile bases of the up/downwind triangles, -1 if absent
    unsigned shared;                // nzmin_s | nzmax_s<<8 : levels nzmin_s..nzmax_s-1 use the up/downwind triangles
    int n1, n2;                     // node ids (for the node-mean gradient fallback)
    int pad;
    double pad2;
};

__host__ __device__ inline int tile_base(int ent, int nlp) { return (ent >> 5) * (nlp * TILE) + (ent & 31); }

struct Options {
    int tra_adv_hor = 1, tra_adv_ver = 0, tra_adv_lim = 0;
    double ph = 1.0, pv = 1.0;
    int w_split = 0, i_vert_diff = 1, Redi = 0, use_sw_pene = 0, do_hor_diff = 1, clamp_salt = 1;
    double is_nonlinfs = 1.0, dt = 0.0, epsilon = 0.1;
    int use_kpp_nonlclflx = 0, ref_sss_local = 0; double ref_sss = 34.7;   // fb200_push_kpp_nonlocal (oce_modules.F90:51-52)
    double clim_relax = 0.0;        // > 1e-8: relax_to_clim is applied to tracers 1 and 2 (fb200_set_clim)
    int smooth_bh_tra = 0; double gamma0_tra = 0.0005, gamma1_tra = 0.0125, gamma2_tra = 0.0;   // diff_part_bh (fb200_set_bh_filter; oce_modules.F90:138-141)
    int relax_zonal = 0;            // relax_zonal_temp of the soufflet channel instead of relax_to_clim (fb200_set_zonal_relax)
    int tracer_id[8] = {0, 1, 301, 301, 301, 301, 301, 301};
};

// error plumbing -------------------------------------------------------------------------------------
void fb_set_error(int code, const std::string& msg);
bool fb_ok();
extern long long g_launches;
#define CUDA_CHECK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fb_set_error(100 + (int)e_, std::string(#x) + ": " + cudaGetErrorString(e_)); } } while (0)
#define LAUNCH_CHECK() do { g_launches++; cudaError_t e_ = cudaGetLastError(); if (e_ != cudaSuccess) fb_set_error(100 + (int)e_, std::string("kernel launch: ") + cudaGetErrorString(e_)); } while (0)

// per-kernel device timing (bench.py's roofline leg): when enabled every ProfScope records a CUDA event pair on the
// launching stream; prof_collect() (after a stream sync) folds them into per-name totals
void prof_enable(bool on);
bool prof_on();
void prof_push(const char* name, cudaStream_t st);
void prof_pop(cudaStream_t st);
void prof_collect();
struct ProfScope {
    cudaStream_t st; bool on;
    ProfScope(const char* name, cudaStream_t s) : st(s), on(prof_on()) { if (on) prof_push(name, st); }
    ~ProfScope() { if (on) prof_pop(st); }
};

template <class T> struct DevBuf {
    T* p = nullptr; size_t n = 0;
    void alloc(size_t count) { release(); n = count; if (count) { CUDA_CHECK(cudaMalloc(&p, count * sizeof(T))); if (p) { CUDA_CHECK(cudaMemset(p, 0, count * sizeof(T))); CUDA_CHECK(cudaDeviceSynchronize()); } } }
    void upload(const std::vector<T>& h) { alloc(h.size()); if (p && n) { CUDA_CHECK(cudaMemcpy(p, h.data(), n * sizeof(T), cudaMemcpyHostToDevice)); CUDA_CHECK(cudaDeviceSynchronize()); } }
    void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
};

struct HaloPlan;
struct TracerCtx {
    // dimensions
    int nl = 0, nlp = 0, myN = 0, N = 0, myE = 0, E = 0, Eall = 0, myEd = 0, ntr = 0;
    int Ntiles = 0, myNtiles = 0, Etiles = 0, Edtiles = 0, maxdeg = 0, maxnie = 0;
    int nwork = 1;                  // sets of per-tracer work fields (= tracers per pass)
    Options opt;
    bool committed = false;
    // host staging of the init_* calls
    std::vector<int> h_edges, h_edge_tri, h_elem_nodes, h_nie_num, h_nie, h_updn, h_nlev, h_ulev, h_nlev_n, h_ulev_n, h_nlmin, h_ulmax, h_nbl;
    std::vector<double> h_ecd, h_edxdy, h_ecos, h_earea, h_gsca, h_area, h_areasvol, h_znb;
    int nie_ld = 0;
    // static device data
    DevBuf<EdgeRec> edge; DevBuf<EdgeHo> edgeho;
    std::vector<int> edge_perm;     // internal edge id of reference edge e (identity unless FB200_EDGE_ORDER=tile, see ctx_commit)
    DevBuf<int> ne_slot;            // [Ntile][maxdeg][32]   (internal edge<<1)|second, -1 = none; in ascending REFERENCE edge id
    // per (node, incident-edge slot) copies of the edge data, [Ntile][maxdeg][32]: a warp of 32 consecutive nodes reads
    // every item fully coalesced (the edge-indexed records would be a scattered 64-byte gather per lane)
    DevBuf<int> tile_deg;                        // [Ntile] max degree within the tile
    DevBuf<int2> fb_list; int n_fb = 0;          // (edge, level) pairs of k_ho_edge_fb: gradients from the node-mean fallback
    DevBuf<SlotRec> ns_p;                        // packed (other node, edge, mask|sign), see SlotRec
    DevBuf<int4> ns_i;                           // tile bases: other node, element 1, element 2 (-1), edge (for adf_h)
    DevBuf<ulonglong2> ns_m;                     // EdgeRec::m1, m2
    DevBuf<double2> ns_d1, ns_d2;                // (dX1,dY1), (dX2,dY2)
    DevBuf<int> nie_slot;           // [Ntile][maxnie][32]   element id in nod_in_elem2D order, -1 = none
    DevBuf<int> el_nb;              // [3][Eall] node tile bases of element corners
    DevBuf<unsigned> el_lev;        // nu | nl<<8 per element (nl = nlevels-1)
    DevBuf<double> el_g;            // [6][myE] gradient_sca
    DevBuf<double> el_area;         // [Eall]
    DevBuf<unsigned> n_lev;         // nu | nlev<<8 | nlmin<<16 | ulmax<<24 per node (nlev = nlevels_nod2D)
    DevBuf<int> n_nbl;              // nboundary_lay
    DevBuf<double> n_znb;           // zbar_n_bot
    DevBuf<double> area, areasvol;  // tiled (nlp levels)
    // per-step state, tiled
    DevBuf<double2> UV;             // (u, v) per element-level, tiled
    DevBuf<double> helem, W, We, Wi, hnode, hnode_new, Z3, zbar3, Kv, Ki, sl1, sl2, sl3, sw3d;
    DevBuf<double> heat_flux, water_flux, virtual_salt, relax_salt, real_salt_flux;   // (N)
    // device-side ALE producers (fb200_ale_*): surface elevation, its right-hand side, element thickness change, edge sums
    DevBuf<double> hbar, hbar_old, ssh_rhs_old, dhe, esum, CFL;
    int which_ale = -1; double w_max_cfl = 1.0;
    // SSH assembly on the device (fb200_ssh_assembly_init, ..._update_stiff_mat, ..._compute_ssh_rhs, ..._update_vel)
    DevBuf<int> sa_rowptr, sa_colind; DevBuf<double> sa_values, sa_rhs, sa_eta; DevBuf<double2> UVr;
    double sa_g = 9.81, sa_alpha = 1.0, sa_theta = 1.0;
    DevBuf<double> kgT, kgS, kb2, kb3;          // KPP non-local transport: ghats / kpp_nonlcltransp{T,S}, blmc(:,:,2:3) (nl, N) tiled
    DevBuf<double> relax2clim, Tclim, Sclim;   // climatology relaxation: (N), tiled, tiled (fb200_set_clim)
    // relax_zonal_temp (fb200_set_zonal_relax): zonal-mean temperature (nl-1, nybins), per node the two bins and the weight
    DevBuf<double> zon_ztem, zon_a; DevBuf<int2> zon_nn; double zon_fac = 0.0;   // zon_fac = tau_inv
    // output means of the resident tracers (fb200_mean_*; io_meandata.F90 update_means): one accumulator per tracer in the
    // stream's precision, allocated on first use
    DevBuf<double> mean8[8]; DevBuf<float> mean4[8]; int mean_count[8] = {0, 0, 0, 0, 0, 0, 0, 0}, mean_prec[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    // tracers, tiled: [ntr] fields each
    DevBuf<double> T, Told;
    // work fields, tiled
    DevBuf<double2> txy, fpm;       // (tr_xy x, y) per element-level; (fct_plus, fct_minus) per node-level
    DevBuf<double> trz, LO, adf_h, adf_v, tstar, eflx;
    DevBuf<double2> extrema_part;                // per-block partials of fb200_tracer_extrema, kept between calls
    DevBuf<double> redi_v, bh_tmp;   // diff_ver_part_redi_expl's del_ttf term (Redi), temporary_ttf of diff_part_bh; allocated on first use
    // staging for host<->device transfers (reference layout)
    DevBuf<double> stage;
    // halo plans of com_nod2D / com_elem2D_full (fb200_init_halo); null on a single partition
    HaloPlan* halo_nod = nullptr; HaloPlan* halo_elem = nullptr;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    float last_ms = 0.f;
};
// stream every entry point works on: the caller's (fb200_set_stream) or one created by the library
cudaStream_t fb_stream();
void fb_set_stream(cudaStream_t s);
// second stream for host<->device copies that overlap kernels (fb200_push_state_early, fb200_solve_tracers_ale);
// fb_copy_sync waits for everything enqueued on it (no-op when it was never used)
cudaStream_t fb_copy_stream();
void fb_copy_sync();

// context.cu
TracerCtx& tctx();
void ctx_commit(TracerCtx& c);
// host (ncomp, lev_ref, ent) component comp  ->  tiled field with dncomp interleaved components, component dcomp
void ctx_push_node_field(TracerCtx& c, const double* host, int lev_ref, int count_ent, double* dev_tiled, int ncomp = 1, int comp = 0, int dncomp = 1, int dcomp = 0);
void ctx_pull_node_field(TracerCtx& c, const double* dev_tiled, int lev_ref, int count_ent, double* host, int dncomp = 1, int dcomp = 0, bool sync = true);
void ctx_release(TracerCtx& c);
// tracer_kernels.cu
void tracer_step_all(TracerCtx& c);
bool tracer_step_begin(TracerCtx& c);
void tracer_step_tracer(TracerCtx& c, int tr);
int tracer_group_size(const TracerCtx& c, int t0);
void tracer_step_group(TracerCtx& c, int t0, int nt);
void tracer_step_end(TracerCtx& c);
void tracer_extrema(TracerCtx& c, int tr, double* tmin, double* tmax);
void tracer_extrema_noalloc(TracerCtx& c, int tr, int nblocks, double* tmin, double* tmax);
void tracer_mean_accumulate(TracerCtx& c, int tr, int accuracy);
void tracer_mean_fetch(TracerCtx& c, int tr, void* mean_out);
