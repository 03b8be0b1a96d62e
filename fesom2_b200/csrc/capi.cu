// extern "C" entry points of the tracer path (include/fesom_b200.h).  Plain pointers and sizes only.
#include "common.cuh"
#include <cmath>
#include "comm.cuh"
#include "fesom_b200.h"
#include <cstring>

int fb_status_code(); const char* fb_status_msg(); void fb_clear_status();
const double* tracer_debug_field(TracerCtx& c, const char* name, int* kind, int* ncomp, int* comp);
double ssh_last_ms();
void prof_reset(); int prof_report(char* buf, int cap);

extern "C" {

int fb200_status(void) { return fb_status_code(); }
const char* fb200_status_message(void) { return fb_status_msg(); }
void fb200_clear_status(void) { fb_clear_status(); }
long long fb200_kernel_launches(void) { return g_launches; }
double fb200_last_ms(const char* what) {
    if (what && !strcmp(what, "psolve")) return ssh_last_ms();
    return (double)tctx().last_ms;
}

void fb200_set_stream(void* cuda_stream) { fb_set_stream((cudaStream_t)cuda_stream); }
void fb200_profile(int* enable) { prof_enable(*enable != 0); if (*enable == 2) prof_reset(); }
int fb200_profile_report(char* buf, int* cap) { return prof_report(buf, *cap); }

void fb200_init_dims(int* nl, int* myDim_nod2D, int* eDim_nod2D, int* myDim_elem2D, int* eDim_elem2D, int* eXDim_elem2D,
                     int* myDim_edge2D, int* num_tracers, int* fcomm) {
    (void)fcomm;
    TracerCtx& c = tctx();
    if (c.committed) ctx_release(c);
    c.nl = *nl; c.myN = *myDim_nod2D; c.N = *myDim_nod2D + *eDim_nod2D;
    c.myE = *myDim_elem2D; c.E = *myDim_elem2D + *eDim_elem2D; c.Eall = c.E + *eXDim_elem2D;
    c.myEd = *myDim_edge2D; c.ntr = *num_tracers;
    if (c.nl < 3 || c.myN < 1 || c.myE < 1 || c.myEd < 1 || c.ntr < 1 || c.ntr > 8) fb_set_error(30, "fb200_init_dims: bad dimensions");
}

void fb200_init_connectivity(int* edges, int* edge_tri, int* elem2D_nodes, int* nod_in_elem2D_num, int* nod_in_elem2D,
                             int* nod_in_elem2D_ld, int* edge_up_dn_tri) {
    TracerCtx& c = tctx();
    c.h_edges.assign(edges, edges + 2 * (size_t)c.myEd);
    c.h_edge_tri.assign(edge_tri, edge_tri + 2 * (size_t)c.myEd);
    c.h_elem_nodes.assign(elem2D_nodes, elem2D_nodes + 3 * (size_t)c.Eall);
    c.h_nie_num.assign(nod_in_elem2D_num, nod_in_elem2D_num + c.N);
    c.nie_ld = *nod_in_elem2D_ld;
    c.h_nie.assign(nod_in_elem2D, nod_in_elem2D + (size_t)c.nie_ld * c.N);
    c.h_updn.assign(edge_up_dn_tri, edge_up_dn_tri + 2 * (size_t)c.myEd);
}

void fb200_init_levels(int* nlevels, int* ulevels, int* nlevels_nod2D, int* ulevels_nod2D, int* nlevels_nod2D_min,
                       int* ulevels_nod2D_max, int* nboundary_lay) {
    TracerCtx& c = tctx();
    c.h_nlev.assign(nlevels, nlevels + c.Eall); c.h_ulev.assign(ulevels, ulevels + c.Eall);
    c.h_nlev_n.assign(nlevels_nod2D, nlevels_nod2D + c.N); c.h_ulev_n.assign(ulevels_nod2D, ulevels_nod2D + c.N);
    c.h_nlmin.assign(nlevels_nod2D_min, nlevels_nod2D_min + c.N); c.h_ulmax.assign(ulevels_nod2D_max, ulevels_nod2D_max + c.N);
    c.h_nbl.assign(nboundary_lay, nboundary_lay + c.N);
    for (int n = 0; n < c.myN; n++)
        if (c.h_nlev_n[n] - c.h_ulev_n[n] < 3) { fb_set_error(31, "columns with fewer than 3 layers are not supported (QR4C boundary statements would overlap)"); break; }
}

void fb200_init_geometry(double* edge_cross_dxdy, double* edge_dxdy, double* elem_cos, double* elem_area, double* gradient_sca,
                         double* area, double* areasvol, double* zbar_n_bot) {
    TracerCtx& c = tctx();
    c.h_ecd.assign(edge_cross_dxdy, edge_cross_dxdy + 4 * (size_t)c.myEd);
    c.h_edxdy.assign(edge_dxdy, edge_dxdy + 2 * (size_t)c.myEd);
    c.h_ecos.assign(elem_cos, elem_cos + c.Eall); c.h_earea.assign(elem_area, elem_area + c.Eall);
    c.h_gsca.assign(gradient_sca, gradient_sca + 6 * (size_t)c.myE);
    c.h_area.assign(area, area + (size_t)c.nl * c.N); c.h_areasvol.assign(areasvol, areasvol + (size_t)c.nl * c.N);
    c.h_znb.assign(zbar_n_bot, zbar_n_bot + c.N);
}

void fb200_set_options(int* tra_adv_hor, int* tra_adv_ver, int* tra_adv_lim, double* tra_adv_ph, double* tra_adv_pv, int* w_split,
                       int* i_vert_diff, int* Redi, int* use_sw_pene, int* do_hor_diff, int* clamp_salt, double* is_nonlinfs,
                       double* dt, double* epsilon, int* tracer_id) {
    TracerCtx& c = tctx();
    Options& o = c.opt;
    o.tra_adv_hor = *tra_adv_hor; o.tra_adv_ver = *tra_adv_ver; o.tra_adv_lim = *tra_adv_lim; o.ph = *tra_adv_ph; o.pv = *tra_adv_pv;
    o.w_split = *w_split; o.i_vert_diff = *i_vert_diff; o.Redi = *Redi; o.use_sw_pene = *use_sw_pene; o.do_hor_diff = *do_hor_diff;
    o.clamp_salt = *clamp_salt; o.is_nonlinfs = *is_nonlinfs; o.dt = *dt; o.epsilon = *epsilon;
    for (int k = 0; k < c.ntr && k < 8; k++) o.tracer_id[k] = tracer_id[k];
    if (o.tra_adv_hor < 0 || o.tra_adv_hor > 2 || o.tra_adv_ver < 0 || o.tra_adv_ver > 3 || o.tra_adv_lim < 0 || o.tra_adv_lim > 1) fb_set_error(32, "unknown advection scheme");
    for (int k = 0; k < c.ntr && k < 8; k++) {             // bc_surface, oce_ale_tracer.F90:1169-1193: any other id stops the reference too
        const int id = o.tracer_id[k];
        if (id == 101) fb_set_error(32, "tracer_id 101 (bc_surface: prec_rain) is not supported: the ABI carries no prec_rain array");
        else if (!(id == 0 || id == 1 || id == 301 || id == 302 || id == 303)) fb_set_error(32, "invalid tracer_id (bc_surface knows 0, 1, 101, 301, 302, 303)");
    }
}

void fb200_set_bh_filter(int* smooth_bh_tra, double* gamma0_tra, double* gamma1_tra, double* gamma2_tra) {
    Options& o = tctx().opt;
    o.smooth_bh_tra = *smooth_bh_tra != 0; o.gamma0_tra = *gamma0_tra; o.gamma1_tra = *gamma1_tra; o.gamma2_tra = *gamma2_tra;
}

void fb200_init_halo(int* nod_rPEnum, int* nod_rPE, int* nod_rptr, int* nod_rlist, int* nod_sPEnum, int* nod_sPE, int* nod_sptr, int* nod_slist,
                     int* elem_rPEnum, int* elem_rPE, int* elem_rptr, int* elem_rlist, int* elem_sPEnum, int* elem_sPE, int* elem_sptr,
                     int* elem_slist) {
    TracerCtx& c = tctx();
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { fb_set_error(12, "no CUDA device: libfesom_b200 has no CPU fallback"); return; }
    if (!c.halo_nod) { c.halo_nod = new HaloPlan(); c.halo_elem = new HaloPlan(); }
    halo_plan_build(*c.halo_nod, *nod_rPEnum, nod_rPE, nod_rptr, nod_rlist, *nod_sPEnum, nod_sPE, nod_sptr, nod_slist);
    halo_plan_build(*c.halo_elem, *elem_rPEnum, elem_rPE, elem_rptr, elem_rlist, *elem_sPEnum, elem_sPE, elem_sptr, elem_slist);
}

void fb200_init_commit(void) { ctx_commit(tctx()); }

// ---- per-step state.  Two halves, cut where the reference's time step cuts them (oce_timestep_ale, oce_ale.F90:2521-2790):
// "early" = what exists before solve_ssh_ale (old thickness and its level depths, mixing coefficients, surface fluxes),
// "late"  = what the SSH solution produces (update_vel -> UV, compute_hbar_ale/vert_vel_ale -> Wvel*, hnode_new, helem).
// fb200_push_state_early only ENQUEUES its copies on the library's copy stream and returns, so they overlap psolve;
// fb200_push_state_late waits for them, pushes the rest and returns when everything is resident.
static cudaStream_t copy_stream() { return fb_copy_stream(); }
static void push_early_on(TracerCtx& c, cudaStream_t s, double* hnode, double* Z_3d_n, double* zbar_3d_n, double* Kv, double* Ki,
                          double* slope_tapered, double* heat_flux, double* water_flux, double* virtual_salt, double* relax_salt,
                          double* real_salt_flux, double* sw_3d) {
    cudaStream_t keep = c.stream; c.stream = s;
    const int L = c.nl - 1, nl = c.nl, N = c.N;
    // NULL = "unchanged on the host side since the library last saw it": hnode then becomes the resident hnode_new of the
    // previous step (update_thickness_ale assigns exactly that, oce_ale.F90:924,936,967), Ki keeps its resident copy (it is
    // set once in oce_setup_step.F90:330 and only rewritten per step under Fer_GM / Redi, oce_fer_gm.F90:243-333)
    if (hnode) ctx_push_node_field(c, hnode, L, N, c.hnode.p);
    else if (c.which_ale < 0) CUDA_CHECK(cudaMemcpyAsync(c.hnode.p, c.hnode_new.p, c.hnode.n * sizeof(double), cudaMemcpyDeviceToDevice, s));
    // (with the device-side ALE producers, fb200_ale_step_end has already updated hnode)
    if (Z_3d_n) ctx_push_node_field(c, Z_3d_n, L, N, c.Z3.p);          // NULL: resident (kept current by fb200_ale_step_end)
    if (zbar_3d_n) ctx_push_node_field(c, zbar_3d_n, nl, N, c.zbar3.p);
    ctx_push_node_field(c, Kv, nl, N, c.Kv.p);
    if (Ki) ctx_push_node_field(c, Ki, L, N, c.Ki.p);
    if (c.opt.Redi) {
        ctx_push_node_field(c, slope_tapered, L, N, c.sl1.p, 3, 0);
        ctx_push_node_field(c, slope_tapered, L, N, c.sl2.p, 3, 1);
        ctx_push_node_field(c, slope_tapered, L, N, c.sl3.p, 3, 2);
    }
    if (c.opt.use_sw_pene) ctx_push_node_field(c, sw_3d, nl, N, c.sw3d.p);
    const size_t nb = (size_t)N * sizeof(double);
    CUDA_CHECK(cudaMemcpyAsync(c.heat_flux.p, heat_flux, nb, cudaMemcpyHostToDevice, s));
    CUDA_CHECK(cudaMemcpyAsync(c.water_flux.p, water_flux, nb, cudaMemcpyHostToDevice, s));
    CUDA_CHECK(cudaMemcpyAsync(c.virtual_salt.p, virtual_salt, nb, cudaMemcpyHostToDevice, s));
    CUDA_CHECK(cudaMemcpyAsync(c.relax_salt.p, relax_salt, nb, cudaMemcpyHostToDevice, s));
    CUDA_CHECK(cudaMemcpyAsync(c.real_salt_flux.p, real_salt_flux, nb, cudaMemcpyHostToDevice, s));
    c.stream = keep;
}
static void push_late_on(TracerCtx& c, cudaStream_t s, double* UV, double* Wvel, double* Wvel_e, double* Wvel_i, double* hnode_new, double* helem) {
    cudaStream_t keep = c.stream; c.stream = s;
    const int L = c.nl - 1, nl = c.nl, N = c.N;
    ctx_push_node_field(c, UV, L, c.E, (double*)c.UV.p, 2, 0, 2, 0);
    ctx_push_node_field(c, UV, L, c.E, (double*)c.UV.p, 2, 1, 2, 1);
    ctx_push_node_field(c, helem, L, c.myE, c.helem.p);
    ctx_push_node_field(c, Wvel, nl, N, c.W.p);
    if (c.opt.w_split) {                               // without w_split Wvel_e == Wvel and Wvel_i == 0 (oce_ale.F90:2193-2201): not read
        ctx_push_node_field(c, Wvel_e, nl, N, c.We.p);
        ctx_push_node_field(c, Wvel_i, nl, N, c.Wi.p);
    }
    ctx_push_node_field(c, hnode_new, L, N, c.hnode_new.p);
    c.stream = keep;
}

void fb200_push_state_early(double* hnode, double* Z_3d_n, double* zbar_3d_n, double* Kv, double* Ki, double* slope_tapered,
                            double* heat_flux, double* water_flux, double* virtual_salt, double* relax_salt, double* real_salt_flux,
                            double* sw_3d) {
    TracerCtx& c = tctx();
    if (!c.committed) { fb_set_error(33, "fb200_push_state_early before fb200_init_commit"); return; }
    c.stream = fb_stream();
    push_early_on(c, copy_stream(), hnode, Z_3d_n, zbar_3d_n, Kv, Ki, slope_tapered, heat_flux, water_flux, virtual_salt, relax_salt,
                  real_salt_flux, sw_3d);                 // no synchronisation: overlaps whatever the host calls next (psolve)
}

void fb200_push_state_late(double* UV, double* Wvel, double* Wvel_e, double* Wvel_i, double* hnode_new, double* helem) {
    TracerCtx& c = tctx();
    if (!c.committed) { fb_set_error(33, "fb200_push_state_late before fb200_init_commit"); return; }
    c.stream = fb_stream();
    push_late_on(c, copy_stream(), UV, Wvel, Wvel_e, Wvel_i, hnode_new, helem);    // same stream as the early half: ordered after it
    CUDA_CHECK(cudaStreamSynchronize(copy_stream()));
}

void fb200_push_step_state(double* UV, double* Wvel, double* Wvel_e, double* Wvel_i, double* hnode, double* hnode_new, double* helem,
                           double* Z_3d_n, double* zbar_3d_n, double* Kv, double* Ki, double* slope_tapered, double* heat_flux,
                           double* water_flux, double* virtual_salt, double* relax_salt, double* real_salt_flux, double* sw_3d) {
    fb200_push_state_early(hnode, Z_3d_n, zbar_3d_n, Kv, Ki, slope_tapered, heat_flux, water_flux, virtual_salt, relax_salt,
                           real_salt_flux, sw_3d);
    if (fb_ok()) fb200_push_state_late(UV, Wvel, Wvel_e, Wvel_i, hnode_new, helem);
}

// relax_to_clim (oce_tracer_mod.F90:87-119): the climatology is static (set with the initial conditions), so it is pushed
// once; *clim_relax <= 1e-8 switches the relaxation off like the reference's test does
void fb200_set_clim(double* clim_relax, double* relax2clim, double* Tclim, double* Sclim) {
    TracerCtx& c = tctx();
    if (!c.committed) { fb_set_error(33, "fb200_set_clim before fb200_init_commit"); return; }
    c.opt.clim_relax = *clim_relax;
    if (!(c.opt.clim_relax > 1.0e-8)) return;
    fb_copy_sync();
    c.stream = fb_stream();
    const size_t nf = (size_t)c.Ntiles * c.nlp * TILE;
    if (!c.relax2clim.p) { c.relax2clim.alloc(c.N); c.Tclim.alloc(nf); c.Sclim.alloc(nf); }
    CUDA_CHECK(cudaMemcpyAsync(c.relax2clim.p, relax2clim, (size_t)c.N * sizeof(double), cudaMemcpyHostToDevice, c.stream));
    ctx_push_node_field(c, Tclim, c.nl - 1, c.N, c.Tclim.p);
    ctx_push_node_field(c, Sclim, c.nl - 1, c.N, c.Sclim.p);
    CUDA_CHECK(cudaStreamSynchronize(c.stream));
}

// relax_zonal_temp of the soufflet channel (toy_channel_soufflet.F90:78-103), which solve_tracers_ale calls INSTEAD of
// relax_to_clim when toy_ocean and which_toy == "soufflet" (oce_ale_tracer.F90:150-151).  ztem(nl-1, nybins) is the zonal-mean
// temperature of compute_zonal_mean (:143-200; call again whenever the host refreshes it), coord_nod2D(2, N) the node
// coordinates in radians, lat0 / dy / tau_inv the module's parameters (:19-24, :207), Tclim(nl-1, N) the restoring target.
// The bin of a node and its interpolation weight (:86-96) are evaluated here on the host, once per call.
void fb200_set_zonal_relax(int* enable, int* nybins, double* lat0, double* dy, double* tau_inv, double* coord_nod2D, double* ztem, double* Tclim) {
    TracerCtx& c = tctx();
    if (!c.committed) { fb_set_error(33, "fb200_set_zonal_relax before fb200_init_commit"); return; }
    c.opt.relax_zonal = *enable != 0;
    if (!c.opt.relax_zonal) return;
    if (*nybins < 1 || !(*dy > 0.0)) { fb_set_error(33, "fb200_set_zonal_relax: nybins >= 1 and dy > 0 required"); c.opt.relax_zonal = 0; return; }
    fb_copy_sync();
    c.stream = fb_stream();
    const int L = c.nl - 1;
    std::vector<int2> bins(c.N);
    std::vector<double> wgt(c.N);
    for (int n = 0; n < c.N; n++) {
        const double yy = coord_nod2D[2 * (size_t)n + 1] - *lat0;
        double a = 0;
        int nn = 1, nn1 = 1;
        if (!(yy < *dy / 2)) {
            nn = (int)floor(yy / *dy - 0.5) + 1;
            nn1 = nn + 1;
            if (nn1 > 100) nn1 = nn;                       // the reference's literal 100, not nybins
            a = yy / *dy + 0.5 - nn;
        }
        if (nn < 1 || nn1 > *nybins) { fb_set_error(33, "fb200_set_zonal_relax: a node lies outside the nybins latitude bins"); c.opt.relax_zonal = 0; return; }
        bins[n] = make_int2(nn, nn1);
        wgt[n] = a;
    }
    const size_t nf = (size_t)c.Ntiles * c.nlp * TILE;
    if (!c.zon_nn.p) { c.zon_nn.alloc(c.N); c.zon_a.alloc(c.N); }
    if (c.zon_ztem.n < (size_t)L * *nybins) c.zon_ztem.alloc((size_t)L * *nybins);
    if (!c.Tclim.p) c.Tclim.alloc(nf);
    CUDA_CHECK(cudaMemcpyAsync(c.zon_nn.p, bins.data(), (size_t)c.N * sizeof(int2), cudaMemcpyHostToDevice, c.stream));
    CUDA_CHECK(cudaMemcpyAsync(c.zon_a.p, wgt.data(), (size_t)c.N * sizeof(double), cudaMemcpyHostToDevice, c.stream));
    CUDA_CHECK(cudaMemcpyAsync(c.zon_ztem.p, ztem, (size_t)L * *nybins * sizeof(double), cudaMemcpyHostToDevice, c.stream));
    ctx_push_node_field(c, Tclim, L, c.N, c.Tclim.p);
    c.zon_fac = *tau_inv;
    CUDA_CHECK(cudaStreamSynchronize(c.stream));            // bins / wgt are locals
}

// KPP non-local transport (use_kpp_nonlclflx, oce_ale_tracer.F90:688-781): per-step arrays of the mixing scheme.
// fesom KPP (mix_scheme_nmb 1, 17): gT = gS = ghats, b2/b3 = blmc(:,:,2)/blmc(:,:,3); cvmix KPP (3, 37): gT/gS =
// kpp_nonlcltranspT/S, b2/b3 = kpp_oblmixc(:,:,2)/(:,:,3).  *enable = 0 switches the term off.
void fb200_push_kpp_nonlocal(int* enable, double* gT, double* gS, double* b2, double* b3, double* ref_sss, int* ref_sss_local) {
    TracerCtx& c = tctx();
    if (!c.committed) { fb_set_error(33, "fb200_push_kpp_nonlocal before fb200_init_commit"); return; }
    c.opt.use_kpp_nonlclflx = *enable != 0;
    if (!c.opt.use_kpp_nonlclflx) return;
    c.opt.ref_sss = *ref_sss; c.opt.ref_sss_local = *ref_sss_local != 0;
    fb_copy_sync();
    c.stream = fb_stream();
    const size_t nf = (size_t)c.Ntiles * c.nlp * TILE;
    if (!c.kgT.p) { c.kgT.alloc(nf); c.kgS.alloc(nf); c.kb2.alloc(nf); c.kb3.alloc(nf); }
    ctx_push_node_field(c, gT, c.nl, c.N, c.kgT.p);
    ctx_push_node_field(c, gS, c.nl, c.N, c.kgS.p);
    ctx_push_node_field(c, b2, c.nl, c.N, c.kb2.p);
    ctx_push_node_field(c, b3, c.nl, c.N, c.kb3.p);
    CUDA_CHECK(cudaStreamSynchronize(c.stream));
}

void fb200_push_tracers(double* tr_arr, double* tr_arr_old) {
    TracerCtx& c = tctx();
    if (!c.committed) { fb_set_error(34, "fb200_push_tracers before fb200_init_commit"); return; }
    fb_copy_sync();
    c.stream = fb_stream();
    const int L = c.nl - 1;
    const size_t nf = (size_t)c.Ntiles * c.nlp * TILE, rf = (size_t)L * c.N;
    for (int tr = 0; tr < c.ntr; tr++) {
        ctx_push_node_field(c, tr_arr + rf * tr, L, c.N, c.T.p + nf * tr);
        ctx_push_node_field(c, tr_arr_old + rf * tr, L, c.N, c.Told.p + nf * tr);
    }
    CUDA_CHECK(cudaStreamSynchronize(c.stream));
}

void fb200_solve_tracers(void) { tracer_step_all(tctx()); }

void fb200_pull_tracers(double* tr_arr, double* tr_arr_old) {
    TracerCtx& c = tctx();
    if (!c.committed) { fb_set_error(35, "fb200_pull_tracers before fb200_init_commit"); return; }
    fb_copy_sync();
    c.stream = fb_stream();
    const int L = c.nl - 1;
    const size_t nf = (size_t)c.Ntiles * c.nlp * TILE, rf = (size_t)L * c.N;
    for (int tr = 0; tr < c.ntr; tr++) {
        ctx_pull_node_field(c, c.T.p + nf * tr, L, c.N, tr_arr + rf * tr);
        ctx_pull_node_field(c, c.Told.p + nf * tr, L, c.N, tr_arr_old + rf * tr);
    }
}

// solve_tracers_ale with HOST arrays in and out (the call the Fortran shim makes).  The tracers are
// independent of each other, so the transfers of one overlap the kernels of the other: a copy stream carries
// push(0), push(1), ..., pull(0), pull(1), ... in that order, the library stream the kernels; events tie kernels(tr) to
// push(tr) and pull(tr) to kernels(tr).  With pinned host arrays (cudaHostRegister by the caller) the copies are
// asynchronous; with pageable arrays the same code runs, just without the overlap.
#define MAX_TR_EVENTS 64
static cudaEvent_t g_ev_push[MAX_TR_EVENTS], g_ev_done[MAX_TR_EVENTS];
static int g_tr_events = 0;
void fb200_solve_tracers_ale(double* tr_arr, double* tr_arr_old) {
    TracerCtx& c = tctx();
    if (!c.committed) { fb_set_error(34, "fb200_solve_tracers_ale before fb200_init_commit"); return; }
    if (c.ntr > MAX_TR_EVENTS) {                       // no overlap bookkeeping for that many tracers: plain sequence
        fb200_push_tracers(tr_arr, tr_arr_old); fb200_solve_tracers(); fb200_pull_tracers(tr_arr, tr_arr_old);
        return;
    }
    copy_stream();
    for (; g_tr_events < c.ntr; g_tr_events++) {
        CUDA_CHECK(cudaEventCreateWithFlags(&g_ev_push[g_tr_events], cudaEventDisableTiming));
        CUDA_CHECK(cudaEventCreateWithFlags(&g_ev_done[g_tr_events], cudaEventDisableTiming));
    }
    if (!tracer_step_begin(c)) return;
    cudaStream_t comp = c.stream, copy = copy_stream();
    const int L = c.nl - 1;
    const size_t nf = (size_t)c.Ntiles * c.nlp * TILE, rf = (size_t)L * c.N;
    for (int t0 = 0; t0 < c.ntr && fb_ok();) {             // the tracers of a pass (tracer_group_size: 1 or 2) go together
        const int nt = tracer_group_size(c, t0);
        c.stream = copy;
        for (int tr = t0; tr < t0 + nt; tr++) {
            ctx_push_node_field(c, tr_arr + rf * tr, L, c.N, c.T.p + nf * tr);
            ctx_push_node_field(c, tr_arr_old + rf * tr, L, c.N, c.Told.p + nf * tr);
        }
        CUDA_CHECK(cudaEventRecord(g_ev_push[t0], copy));
        c.stream = comp;
        CUDA_CHECK(cudaStreamWaitEvent(comp, g_ev_push[t0], 0));
        tracer_step_group(c, t0, nt);
        for (int tr = t0; tr < t0 + nt; tr++) CUDA_CHECK(cudaEventRecord(g_ev_done[tr], comp));
        t0 += nt;
    }
    const bool late = c.opt.relax_zonal != 0;             // relax_zonal_temp touches tracer 1 on every later pass
    for (int tr = 0; tr < c.ntr && fb_ok(); tr++) {
        c.stream = copy;
        CUDA_CHECK(cudaStreamWaitEvent(copy, g_ev_done[late ? c.ntr - 1 : tr], 0));
        ctx_pull_node_field(c, c.T.p + nf * tr, L, c.N, tr_arr + rf * tr, 1, 0, false);
        ctx_pull_node_field(c, c.Told.p + nf * tr, L, c.N, tr_arr_old + rf * tr, 1, 0, false);
    }
    c.stream = comp;
    tracer_step_end(c);
    CUDA_CHECK(cudaStreamSynchronize(copy));
}

// Time stepping with device-resident tracers: the step consumes the tracers the previous step left in HBM (first step:
// fb200_push_tracers) and the host receives tr_arr only — tr_arr_old is read by nothing but init_tracers_AB
// (oce_tracer_mod.F90:64), i.e. by this library.  The copy of tracer k overlaps the kernels of tracer k+1.
void fb200_solve_tracers_ale_resident(double* tr_arr_out) {
    TracerCtx& c = tctx();
    if (!c.committed) { fb_set_error(34, "fb200_solve_tracers_ale_resident before fb200_init_commit"); return; }
    if (c.ntr > MAX_TR_EVENTS) { fb_set_error(38, "fb200_solve_tracers_ale_resident: too many tracers"); return; }
    copy_stream();
    for (; g_tr_events < c.ntr; g_tr_events++) {
        CUDA_CHECK(cudaEventCreateWithFlags(&g_ev_push[g_tr_events], cudaEventDisableTiming));
        CUDA_CHECK(cudaEventCreateWithFlags(&g_ev_done[g_tr_events], cudaEventDisableTiming));
    }
    if (!tracer_step_begin(c)) return;
    cudaStream_t comp = c.stream, copy = copy_stream();
    const int L = c.nl - 1;
    const size_t nf = (size_t)c.Ntiles * c.nlp * TILE, rf = (size_t)L * c.N;
    const bool late = c.opt.relax_zonal != 0;             // relax_zonal_temp touches tracer 1 on every later pass: copy after the loop
    for (int t0 = 0; t0 < c.ntr && fb_ok();) {
        const int nt = tracer_group_size(c, t0);
        tracer_step_group(c, t0, nt);
        const int last = t0 + nt - 1;
        CUDA_CHECK(cudaEventRecord(g_ev_done[last], comp));
        t0 += nt;
        if (late && t0 < c.ntr) continue;
        c.stream = copy;
        CUDA_CHECK(cudaStreamWaitEvent(copy, g_ev_done[last], 0));
        for (int k = late ? 0 : last - nt + 1; k <= last; k++) ctx_pull_node_field(c, c.T.p + nf * k, L, c.N, tr_arr_out + rf * k, 1, 0, false);
        c.stream = comp;
    }
    tracer_step_end(c);
    CUDA_CHECK(cudaStreamSynchronize(copy));
}

// extrema of the resident tracer *tr_num (1-based) over this partition's own wet cells (the host combines the ranks with
// its own MPI_Allreduce, as write_step_info.F90 does)
void fb200_tracer_extrema(int* tr_num, double* tmin, double* tmax) { tracer_extrema(tctx(), *tr_num - 1, tmin, tmax); }
// the same scan without any allocation (static device scratch, see tracer_kernels.cu); *nblocks <= 0: the default grid
void fb200_tracer_extrema_noalloc(int* tr_num, int* nblocks, double* tmin, double* tmax) { tracer_extrema_noalloc(tctx(), *tr_num - 1, *nblocks, tmin, tmax); }

// output means of the resident tracer *tr_num (1-based): fb200_mean_accumulate = update_means for that stream
// (io_meandata.F90:768-794; *accuracy 8 = real64, 4 = real32), fb200_mean_fetch = compute_means + clean_meanarrays (:894-900)
void fb200_mean_accumulate(int* tr_num, int* accuracy) { tracer_mean_accumulate(tctx(), *tr_num - 1, *accuracy); }
void fb200_mean_fetch(int* tr_num, void* mean_out) { tracer_mean_fetch(tctx(), *tr_num - 1, mean_out); }

void fb200_debug_fetch(const char* name, double* out) {
    TracerCtx& c = tctx();
    if (!c.committed) { fb_set_error(36, "fb200_debug_fetch before fb200_init_commit"); return; }
    fb_copy_sync();
    c.stream = fb_stream();
    int kind = 0, ncomp = 1, comp = 0;
    std::string s(name);
    const double* f = nullptr;
    int lev = c.nl - 1, cnt = c.N;
    if (s == "edge_perm") { for (int e = 0; e < c.myEd; e++) out[e] = (double)c.edge_perm[e]; return; }   // internal id of reference edge e
    if (s == "ttfAB") { f = c.Told.p; }
    else f = tracer_debug_field(c, name, &kind, &ncomp, &comp);
    if (!f) { fb_set_error(37, "fb200_debug_fetch: unknown field " + s); return; }
    if (s == "adf_v" || kind == 3) lev = c.nl;
    if (kind == 1) cnt = c.myEd;
    if (kind == 2) cnt = c.Eall;
    if (kind == 1) {                                   // edge field: stored in the internal edge order, returned in the reference's
        std::vector<double> tmp((size_t)lev * cnt);
        ctx_pull_node_field(c, f, lev, cnt, tmp.data(), ncomp, comp);
        if (!fb_ok()) return;
        for (int e = 0; e < cnt; e++) memcpy(out + (size_t)e * lev, tmp.data() + (size_t)c.edge_perm[e] * lev, (size_t)lev * sizeof(double));
        return;
    }
    ctx_pull_node_field(c, f, lev, cnt, out, ncomp, comp);
}

void fb200_finalize(void) { fb_copy_sync(); ctx_release(tctx()); }

}  // extern "C"
