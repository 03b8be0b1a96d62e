/* ORACLE (test infrastructure) — context plumbing: string-keyed setters so the Python side never has to
 * mirror the struct layout. */
#include "oracle.h"
#include <stdlib.h>
#include <string.h>

#define INT_FIELDS(X) X(nl) X(myDim_nod2D) X(eDim_nod2D) X(myDim_elem2D) X(eDim_elem2D) X(eXDim_elem2D) X(myDim_edge2D) \
    X(nod_in_elem2D_max) X(num_tracers) X(tra_adv_hor) X(tra_adv_ver) X(tra_adv_lim) X(w_split) X(i_vert_diff) X(Redi)    \
    X(use_sw_pene) X(do_hor_diff) X(clamp_salt) X(use_kpp_nonlclflx) X(ref_sss_local) X(fct_cavity_neutral) X(relax_zonal) X(smooth_bh_tra) X(bh_split) X(skip_redi_expl)
#define DBL_FIELDS(X) X(dt) X(epsilon) X(tra_adv_ph) X(tra_adv_pv) X(is_nonlinfs) X(clim_relax) X(ref_sss) X(zonal_lat0) X(zonal_dy) X(zonal_tau_inv) X(gamma0_tra) X(gamma1_tra) X(gamma2_tra)
#define CIPTR_FIELDS(X) X(edges) X(edge_tri) X(elem2D_nodes) X(nlevels) X(ulevels) X(nlevels_nod2D) X(ulevels_nod2D) \
    X(nlevels_nod2D_min) X(ulevels_nod2D_max) X(nod_in_elem2D) X(nod_in_elem2D_num) X(edge_up_dn_tri) X(nboundary_lay)
#define CDPTR_FIELDS(X) X(edge_cross_dxdy) X(edge_dxdy) X(elem_cos) X(elem_area) X(gradient_sca) X(area) X(areasvol) \
    X(zbar_n_bot) X(UV) X(Wvel) X(Wvel_e) X(Wvel_i) X(hnode) X(hnode_new) X(helem) X(Z_3d_n) X(zbar_3d_n) X(Kv) X(Ki)  \
    X(slope_tapered) X(heat_flux) X(water_flux) X(virtual_salt) X(relax_salt) X(real_salt_flux) X(sw_3d) X(relax2clim) X(Tclim) X(Sclim) X(kpp_gT) X(kpp_gS) X(kpp_b2) X(kpp_b3) X(zonal_ztem) X(coord_nod2D)
#define DPTR_FIELDS(X) X(tr_arr) X(tr_arr_old)
#define WORK_FIELDS(X) X(del_ttf) X(del_ttf_advhoriz) X(del_ttf_advvert) X(tr_xy) X(tr_z) X(edge_up_dn_grad) X(fct_LO) \
    X(adv_flux_hor) X(adv_flux_ver) X(fct_ttf_max) X(fct_ttf_min) X(fct_plus) X(fct_minus) X(aux_minmax) X(bh_tmp)

oracle_t* oracle_create(void) {
    oracle_t* o = (oracle_t*)calloc(1, sizeof(oracle_t));
    o->epsilon = 0.1;          /* oce_modules.F90:82 */
    o->tra_adv_ph = 1.0; o->tra_adv_pv = 1.0;
    o->tra_adv_hor = O_HOR_MFCT; o->tra_adv_ver = O_VER_QR4C; o->tra_adv_lim = O_LIM_FCT;   /* config/namelist.oce:74-78 */
    o->i_vert_diff = 1; o->do_hor_diff = 1; o->clamp_salt = 1; o->is_nonlinfs = 1.0;
    o->num_tracers = 2; o->tracer_id[0] = 0; o->tracer_id[1] = 1;
    o->ref_sss = 34.7;          /* oce_modules.F90:52 */
    o->gamma0_tra = 0.0005; o->gamma1_tra = 0.0125; o->gamma2_tra = 0.0;   /* oce_modules.F90:139-141 */
    return o;
}
void oracle_free(oracle_t* o) {
    if (!o) return;
#define X(f) free(o->f);
    WORK_FIELDS(X)
#undef X
    free(o);
}
int oracle_set_int(oracle_t* o, const char* name, int v) {
#define X(f) if (!strcmp(name, #f)) { o->f = v; return 0; }
    INT_FIELDS(X)
#undef X
    if (!strncmp(name, "tracer_id", 9)) { int k = atoi(name + 9); if (k >= 0 && k < 8) { o->tracer_id[k] = v; return 0; } }
    return 1;
}
int oracle_set_dbl(oracle_t* o, const char* name, double v) {
#define X(f) if (!strcmp(name, #f)) { o->f = v; return 0; }
    DBL_FIELDS(X)
#undef X
    return 1;
}
int oracle_set_ptr(oracle_t* o, const char* name, void* p) {
#define X(f) if (!strcmp(name, #f)) { o->f = (const int*)p; return 0; }
    CIPTR_FIELDS(X)
#undef X
#define X(f) if (!strcmp(name, #f)) { o->f = (const double*)p; return 0; }
    CDPTR_FIELDS(X)
#undef X
#define X(f) if (!strcmp(name, #f)) { o->f = (double*)p; return 0; }
    DPTR_FIELDS(X)
#undef X
    return 1;
}
void* oracle_get_ptr(oracle_t* o, const char* name) {
#define X(f) if (!strcmp(name, #f)) return (void*)o->f;
    WORK_FIELDS(X) DPTR_FIELDS(X)
#undef X
    return NULL;
}
int oracle_alloc_work(oracle_t* o) {     /* array_setup + oce_adv_tra_fct_init + find_up_downwind_triangles allocations */
    size_t N = (size_t)(o->myDim_nod2D + o->eDim_nod2D), L = (size_t)(o->nl - 1), nl = (size_t)o->nl;
    size_t Eall = (size_t)(o->myDim_elem2D + o->eDim_elem2D + o->eXDim_elem2D), Ed = (size_t)o->myDim_edge2D;
#define AL(f, cnt) do { free(o->f); o->f = (double*)calloc((cnt) ? (cnt) : 1, sizeof(double)); if (!o->f) return 1; } while (0)
    AL(del_ttf, L * N); AL(del_ttf_advhoriz, L * N); AL(del_ttf_advvert, L * N);
    AL(tr_xy, 2 * L * Eall); AL(tr_z, nl * N); AL(edge_up_dn_grad, 4 * L * Ed);
    AL(fct_LO, L * N); AL(adv_flux_hor, L * Ed); AL(adv_flux_ver, nl * (size_t)o->myDim_nod2D);
    AL(fct_ttf_max, L * N); AL(fct_ttf_min, L * N); AL(fct_plus, L * N); AL(fct_minus, L * N);
    AL(aux_minmax, 2 * L * (size_t)(o->myDim_elem2D + o->eDim_elem2D));
    AL(bh_tmp, L * N);
#undef AL
    return 0;
}

/* Output means, io_meandata.F90: update_means (:768-794) adds the stream's field to local_values in the stream's precision
 * (accuracy 8 = real64; 4 = real32 with the field rounded to real32 first) — the caller counts addcounter; compute_means +
 * clean_meanarrays (:894-900) divide by real(addcounter) in that precision and clear. */
void o_update_means(int accuracy, void* local_values, const double* ptr3, long long n) {
    if (accuracy == 8) { double* a = (double*)local_values; for (long long i = 0; i < n; i++) a[i] = a[i] + ptr3[i]; }
    else { float* a = (float*)local_values; for (long long i = 0; i < n; i++) a[i] = a[i] + (float)ptr3[i]; }
}
void o_compute_means(int accuracy, void* local_values, void* copy, long long n, int addcounter) {
    if (accuracy == 8) { double* a = (double*)local_values; double* c = (double*)copy; const double d = (double)addcounter; for (long long i = 0; i < n; i++) { c[i] = a[i] / d; a[i] = 0.0; } }
    else { float* a = (float*)local_values; float* c = (float*)copy; const float d = (float)addcounter; for (long long i = 0; i < n; i++) { c[i] = a[i] / d; a[i] = 0.0f; } }
}
