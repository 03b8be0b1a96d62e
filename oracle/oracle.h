/* ORACLE — TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement, in plain C, of the reference's (cnissenawi/fesom2) per-timestep tracer transport and
 * SSH solve.  It exists to CHECK the CUDA path (tests/, __graft_entry__.smoke(), bench.py's cpu_baseline
 * and --impl reference legs); the product (fesom2_b200/csrc) never links, imports or calls it.
 *
 * Every function cites the reference file:line it follows and keeps the reference's loop order and
 * floating-point operation order (compile with -ffp-contract=off).  Arrays are in the reference's own
 * layout: Fortran column-major, level fastest, index arrays hold 1-based values.
 *
 * Pinning status: the Fortran part of the reference cannot be built here (no Fortran compiler, MPI or
 * NetCDF) and the reference ships no kernel-level golden vectors => for the tracer routines this oracle
 * is "parity unpinned" against the reference binary; it is pinned instead by the invariants the
 * reference's code implies (tests/test_oracle_*.py).  The SSH solve IS pinned: oracle/_ref builds the
 * reference's own psolve.c + pARMS from /root/reference and tests compare against it.
 */
#ifndef FESOM_ORACLE_H
#define FESOM_ORACLE_H
#include <stddef.h>

enum { O_HOR_MUSCL = 0, O_HOR_MFCT = 1, O_HOR_UPW1 = 2 };
enum { O_VER_QR4C = 0, O_VER_CDIFF = 1, O_VER_PPM = 2, O_VER_UPW1 = 3 };
enum { O_LIM_FCT = 0, O_LIM_NONE = 1 };

typedef struct oracle {
    /* dimensions (g_PARSUP / t_mesh) */
    int nl, myDim_nod2D, eDim_nod2D, myDim_elem2D, eDim_elem2D, eXDim_elem2D, myDim_edge2D, nod_in_elem2D_max;
    int num_tracers;
    /* mesh connectivity + geometry (MOD_MESH.F90:19-95) */
    const int *edges, *edge_tri, *elem2D_nodes, *nlevels, *ulevels, *nlevels_nod2D, *ulevels_nod2D,
              *nlevels_nod2D_min, *ulevels_nod2D_max, *nod_in_elem2D, *nod_in_elem2D_num,
              *edge_up_dn_tri, *nboundary_lay;
    const double *edge_cross_dxdy, *edge_dxdy, *elem_cos, *elem_area, *gradient_sca, *area, *areasvol, *zbar_n_bot;
    /* per-step inputs (o_ARRAYS) */
    const double *UV, *Wvel, *Wvel_e, *Wvel_i, *hnode, *hnode_new, *helem, *Z_3d_n, *zbar_3d_n, *Kv, *Ki,
                 *slope_tapered, *heat_flux, *water_flux, *virtual_salt, *relax_salt, *real_salt_flux, *sw_3d;
    /* climatology relaxation (o_ARRAYS: relax2clim(N), Tclim/Sclim(nl-1,N); namelist oce_tra: clim_relax) */
    const double *relax2clim, *Tclim, *Sclim;
    /* relax_zonal_temp of the soufflet channel (toy_channel_soufflet.F90:78-103), taken instead of relax_to_clim when
     * relax_zonal != 0: zonal-mean temperature zonal_ztem(nl-1, nybins) from compute_zonal_mean (:143-200), node latitudes
     * coord_nod2D(2,N) [rad], module parameters lat0, dy = ysize/nybins/r_earth, tau_inv (:19-24, :207) */
    const double *zonal_ztem, *coord_nod2D;
    double zonal_lat0, zonal_dy, zonal_tau_inv;
    int relax_zonal;
    /* KPP non-local transport (use_kpp_nonlclflx): ghats (fesom KPP) or kpp_nonlcltranspT/S (cvmix) as kpp_gT/kpp_gS (nl,N),
     * blmc(:,:,2:3) or kpp_oblmixc(:,:,2:3) as kpp_b2/kpp_b3 (nl,N); ref_sss, ref_sss_local (oce_modules.F90:51-52) */
    const double *kpp_gT, *kpp_gS, *kpp_b2, *kpp_b3;
    /* tracers, (nl-1, N, num_tracers) */
    double *tr_arr, *tr_arr_old;
    /* work arrays owned by the oracle (o_ARRAYS / o_MESH work arrays) */
    double *del_ttf, *del_ttf_advhoriz, *del_ttf_advvert, *tr_xy, *tr_z, *edge_up_dn_grad, *fct_LO,
           *adv_flux_hor, *adv_flux_ver, *fct_ttf_max, *fct_ttf_min, *fct_plus, *fct_minus, *aux_minmax, *bh_tmp;
    /* parameters (namelists oce_tra / oce_dyn, g_config) */
    double dt, epsilon, tra_adv_ph, tra_adv_pv, is_nonlinfs, clim_relax, ref_sss;
    int tra_adv_hor, tra_adv_ver, tra_adv_lim, w_split, i_vert_diff, Redi, use_sw_pene, do_hor_diff, clamp_salt, use_kpp_nonlclflx, ref_sss_local;
    /* oce_tra_adv_fct a2/a3 under cavities: a3 (oce_adv_tra_fct.F90:139-140) reads UV_rhs(1:2,nz,elem) of elements whose
     * cavity top lies below nz, entries a2 (:108-121) never wrote — in the reference they hold whatever the momentum step
     * left in UV_rhs.  0: the scratch as allocated (zeros, then the previous call's values); 1: those entries are neutral
     * (-bignumber / +bignumber, the rule a2 applies below the bottom), which is what the CUDA path computes. */
    int fct_cavity_neutral;
    /* diff_part_bh (oce_ale_tracer.F90:1081-1148): smooth_bh_tra, gamma0_tra..gamma2_tra (namelist oce_tra, oce_modules.F90:138-141);
     * bh_split: the caller runs the two rounds itself with the reference's exchange_nod(temporary_ttf) in between */
    int smooth_bh_tra, bh_split;
    int skip_redi_expl;   /* tests only: leave the term out to show that it matters */
    double gamma0_tra, gamma1_tra, gamma2_tra;
    int tracer_id[8];
} oracle_t;

oracle_t* oracle_create(void);
void oracle_free(oracle_t* o);
int oracle_set_int(oracle_t* o, const char* name, int v);
int oracle_set_dbl(oracle_t* o, const char* name, double v);
int oracle_set_ptr(oracle_t* o, const char* name, void* p);
void* oracle_get_ptr(oracle_t* o, const char* name);
int oracle_alloc_work(oracle_t* o);

/* routines, one per reference subroutine */
void o_tracer_gradient_elements(oracle_t* o, const double* ttf);
void o_tracer_gradient_z(oracle_t* o, const double* ttf);
void o_fill_up_dn_grad(oracle_t* o);
void o_adv_tra_hor_upw1(oracle_t* o, const double* ttf, const double* vel, double* flux, int init_zero);
void o_adv_tra_hor_muscl(oracle_t* o, const double* ttf, const double* vel, double num_ord, double* flux, int init_zero, int mfct);
void o_adv_tra_ver_upw1(oracle_t* o, const double* ttf, const double* W, double* flux, int init_zero);
void o_adv_tra_ver_qr4c(oracle_t* o, const double* ttf, const double* W, double num_ord, double* flux, int init_zero);
void o_adv_tra_vert_ppm(oracle_t* o, const double* ttf, const double* W, double* flux, int init_zero);
void o_adv_tra_ver_cdiff(oracle_t* o, const double* ttf, const double* W, double* flux, int init_zero);
void o_adv_tra_vert_impl(oracle_t* o, double* ttf, const double* W);
void o_oce_tra_adv_fct_a(oracle_t* o, const double* ttf, const double* lo, double* adf_h, double* adf_v);
void o_oce_tra_adv_fct_b(oracle_t* o, double* adf_h, double* adf_v);
void o_oce_tra_adv_flux2dtracer(oracle_t* o, double* dttf_h, double* dttf_v, double* flux_h, double* flux_v,
                                int use_lo, const double* ttf, const double* lo);
void o_diff_part_hor_redi(oracle_t* o);
void o_diff_ver_part_expl_ale(oracle_t* o, int tr_num);
void o_diff_ver_part_redi_expl(oracle_t* o);
void o_diff_part_bh(oracle_t* o, double* ttf);
void o_diff_part_bh_round1(oracle_t* o, const double* ttf, double* temporary_ttf);
void o_diff_part_bh_round2(oracle_t* o, double* ttf, const double* temporary_ttf);
void o_diff_ver_part_impl_ale(oracle_t* o, int tr_num);
void o_relax_to_clim(oracle_t* o, int tr_num);
void o_relax_zonal_temp(oracle_t* o);
void o_update_means(int accuracy, void* local_values, const double* ptr3, long long n);
void o_compute_means(int accuracy, void* local_values, void* copy, long long n, int addcounter);

/* phases of solve_tracers_ale between halo exchanges (see tracer.c) */
void o_phase1(oracle_t* o, int tr_num);   /* AB2 + tracer_gradient_elements(AB)            -> exchange_elem(tr_xy)            */
void o_phase2(oracle_t* o, int tr_num);   /* tracer_gradient_z, fill_up_dn_grad, grad(T)   -> exchange_nod(tr_z), elem(tr_xy) */
void o_phase3(oracle_t* o, int tr_num);   /* LO solution                                    -> exchange_nod(fct_LO)            */
void o_phase4(oracle_t* o, int tr_num);   /* HO fluxes, FCT a1..b2                          -> exchange_nod(fct_plus,fct_minus)*/
void o_phase5(oracle_t* o, int tr_num);   /* FCT b3, update, diffusion                      -> exchange_nod(tr_arr)            */
void o_phase_final(oracle_t* o);          /* salinity clamp, oce_ale_tracer.F90:176-198 */
void o_solve_tracers_ale(oracle_t* o);    /* one partition, no exchanges */

/* SSH: psolve.c row scaling + bicgstab_ras.c Krylov loop with a pluggable preconditioner */
int o_ssh_solve(int n, const int* rowptr0, const int* colind0, const double* values, const double* rhs, double* x,
                int pc_kind, int block, double tol, int maxits, int* its_out, double* res_out);
void o_spmv(int n, const int* rowptr0, const int* colind0, const double* a, const double* x, double* y);

#endif
