/* fesom_b200.h — C ABI of libfesom_b200.so: FESOM2's per-timestep tracer transport (FCT advection + implicit
 * vertical diffusion) and SSH elliptic solve as hand-written FP64 CUDA kernels for sm_100a.
 *
 * The library is called from FESOM2's own Fortran time loop through ISO_C_BINDING (INTEGRATION.md shows the
 * interface blocks).  Conventions are the reference's own (src/psolve.c, src/oce_ale.F90:2272-2291):
 *   - every argument is passed BY REFERENCE (Fortran default), integer(C_INT) / real(C_DOUBLE);
 *   - arrays are the caller's Fortran arrays as they are: column-major, level index fastest,
 *     index arrays hold 1-based values (psolver_init receives the 0-based shifted temporaries the
 *     reference passes, oce_ale.F90:2332-2333);
 *   - the caller owns all arrays; the library copies what it keeps;
 *   - no return codes: an error prints "fesom_b200: ERROR ..." to stderr and sets a sticky status that
 *     fb200_status() reports; with FB200_ABORT_ON_ERROR=1 in the environment it calls abort()
 *     (the analogue of par_ex(1), src/gen_modules_partitioning.F90:113-157);
 *   - there is no CPU fallback: without a CUDA device every compute entry point raises that error.
 */
#ifndef FESOM_B200_H
#define FESOM_B200_H
#ifdef __cplusplus
extern "C" {
#endif

/* ---------------------------------------------------------------------------------------------------
 * SSH solve: symbol-for-symbol replacement of src/psolve.c (psolver_init :16, psolve :152, psolver_final :117).
 * stype/pctype/pcilutype/ilulevel/fillin/droptol/restart are accepted and ignored: the Krylov method is
 * right-preconditioned BiCGStab (lib/parms/src/bicgstab_ras.c:49-259); the preconditioner (RAS + ILU(k) in the
 * reference, lib/parms/src/parms_pc_ras.c:56-120) is block-Jacobi or an aggregation-multigrid V-cycle, see
 * fb200_psolver_set_pc.
 *   part  : npes+1 row offsets of the PE-contiguous numbering (0-based)
 *   rptr  : N_own+1 CSR row pointers (0-based)         cols : nnz global PE-contiguous column ids (0-based)
 *   vals  : nnz matrix values                          reuse: 1 = values will be replaced by later psolve calls
 *   fcomm : Fortran MPI communicator handle (kept for ABI identity; the ranks are introduced by fb200_comm_init, which must come
 *           first when there is more than one.  A process that a launcher started as one of several — OMPI_COMM_WORLD_SIZE,
 *           PMI_SIZE, MV2_COMM_WORLD_SIZE, SLURM_NTASKS or WORLD_SIZE > 1 — and that did not call fb200_comm_init is refused with
 *           error 47 instead of solving part[0..1] on every rank; FB200_SINGLE_RANK=1 states that this is intended.)
 * Errors found while the ranks negotiate the interface lists (44, 45) are agreed on before any rank returns: all fail together.
 */
void psolver_init(int* id, int* stype, int* pctype, int* pcilutype, int* ilulevel, int* fillin, double* droptol,
                  int* maxits, int* restart, double* soltol, int* part, int* rptr, int* cols, double* vals,
                  int* reuse, int* fcomm);
/* rhs: N_own (+halo) right-hand side; sol: in = initial guess (warm start, psolve.c:206), out = solution;
 * newvalues: 0 keep matrix, 1 replace values (keep preconditioner), 2 replace values and rebuild preconditioner.
 * rhs, vals and sol may be HOST pointers (the reference's call) or DEVICE pointers (unified addressing decides per pointer): a
 * host that assembles the system on the device solves without crossing PCIe.
 * On several ranks the multigrid preconditioner works on overlapping subdomains (restricted additive Schwarz, FB200_SSH_OVERLAP
 * layers of neighbour rows, default 4; the reference's RAS extends by one layer) and the small messages of the iteration travel over peer
 * memory (CUDA IPC over NVLink) when every rank can map its peers' buffers, else over NCCL (FB200_SSH_COMM=nccl forces it). */
void psolve(int* id, double* rhs, double* vals, double* sol, int* newvalues);
void psolver_final(void);
/* iteration count and final scaled residual norm of the last psolve(id) (the reference keeps them in the
 * pARMS solver object, parms_SolverGetIts / bicgstab_ras.c:257) */
void fb200_psolve_info(int* id, int* its, double* resnorm);
/* block size of the block-Jacobi preconditioner for solvers created afterwards (1 = point Jacobi) */
void fb200_psolver_set_block(int* block);
/* preconditioner of solvers created afterwards: -1 = default, 0 = block-Jacobi, 1 = aggregation multigrid (one V(1,1) cycle with
 * damped-Jacobi smoothing on this rank's diagonal block; hierarchy built once at psolver_init like the reference's ILU
 * factor, rebuilt by psolve(newvalues=2)).  Without a call the environment variable FB200_SSH_PC ("amg" | "bj") decides,
 * else the library default (multigrid). */
void fb200_psolver_set_pc(int* kind);
/* multigrid parameters: strength-of-connection threshold, smoother weight, number of levels, rows below which
 * coarsening stops (defaults 0.08, 0.8, 4, 200) */
void fb200_psolver_set_amg(double* theta, double* omega, int* max_levels, int* min_rows);
/* rows and stored sliced-ELL entries of A_l, P_l, R_l of the hierarchy of solver id (rows = 0 past the last level) */
void fb200_psolve_amg_info(int* id, int* level, int* rows, long long* a_entries, long long* p_entries, long long* r_entries);
/* 1 when solver *id applies the FOLDED cycle (default; FB200_AMG_FOLD=0 keeps the step-by-step kernels): per level one SpMV down
 * (M = R (I - A W)) and one two-matrix SpMV up (T = (I + S) W, Q = S P; W = omega D^-1, S = I - W A) instead of residual /
 * restriction / prolongation / smoothing; fb200_psolve_amg_info then reports the stored entries of T, Q, M in place of A, P, R. */
int fb200_psolve_amg_folded(int* id);
/* host-only access to the multigrid setup for CPU tests (no CUDA call): build a hierarchy from a 0-based CSR matrix,
 * read level `level`'s matrix `which` (0 A, 1 P, 2 R) back as CSR, free it */
int fb200_amg_host_build(int* n, int* rptr, int* cols, double* vals, double* theta, int* max_levels, int* min_rows);
int fb200_amg_host_levels(int* handle);
void fb200_amg_host_dims(int* handle, int* level, int* which, int* nrow, int* ncol, int* nnz);
void fb200_amg_host_fetch(int* handle, int* level, int* which, int* rptr, int* cols, double* vals);
void fb200_amg_host_free(int* handle);
/* 1 (default): replay the BiCGStab iterations from a captured CUDA graph; 0: plain launches */
void fb200_psolver_use_graph(int* on);

/* ---------------------------------------------------------------------------------------------------
 * Tracer path.  Replaces the body of solve_tracers_ale (src/oce_ale_tracer.F90:101-199): init_tracers_AB,
 * adv_tracers_ale -> do_oce_adv_tra (src/oce_adv_tra_driver.F90:41), diff_tracers_ale, salinity clamp.
 * Call order: fb200_init_dims, fb200_init_connectivity, fb200_init_levels, fb200_init_geometry,
 * [fb200_init_halo], fb200_set_options, fb200_init_commit — once, after mesh_setup/ocean_setup; then per step
 * fb200_push_step_state + fb200_solve_tracers_ale (or the push/solve/pull triple to keep tracers resident).
 */
void fb200_init_dims(int* nl, int* myDim_nod2D, int* eDim_nod2D, int* myDim_elem2D, int* eDim_elem2D,
                     int* eXDim_elem2D, int* myDim_edge2D, int* num_tracers, int* fcomm);
/* t_mesh connectivity (src/MOD_MESH.F90:19-95), local 1-based numbering:
 * edges(2,myDim_edge2D) edge_tri(2,myDim_edge2D) elem2D_nodes(3,myDim+eDim+eXDim elems)
 * nod_in_elem2D_num(N) nod_in_elem2D(ld,N)  edge_up_dn_tri(2,myDim_edge2D) (src/oce_muscl_adv.F90:141) */
void fb200_init_connectivity(int* edges, int* edge_tri, int* elem2D_nodes, int* nod_in_elem2D_num,
                             int* nod_in_elem2D, int* nod_in_elem2D_ld, int* edge_up_dn_tri);
/* nlevels/ulevels (all local elems), nlevels_nod2D/ulevels_nod2D/nlevels_nod2D_min/ulevels_nod2D_max (N),
 * nboundary_lay(N) (src/oce_muscl_adv.F90:74-104) */
void fb200_init_levels(int* nlevels, int* ulevels, int* nlevels_nod2D, int* ulevels_nod2D, int* nlevels_nod2D_min,
                       int* ulevels_nod2D_max, int* nboundary_lay);
/* edge_cross_dxdy(4,E) edge_dxdy(2,E) elem_cos(elems) elem_area(elems) gradient_sca(6,myDim_elem2D)
 * area(nl,N) areasvol(nl,N) zbar_n_bot(N) */
void fb200_init_geometry(double* edge_cross_dxdy, double* edge_dxdy, double* elem_cos, double* elem_area,
                         double* gradient_sca, double* area, double* areasvol, double* zbar_n_bot);
/* namelist switches the Fortran routines read from modules (src/oce_modules.F90:82-99, config/namelist.oce):
 * tra_adv_hor 0 MUSCL 1 MFCT 2 UPW1; tra_adv_ver 0 QR4C 1 CDIFF 2 PPM 3 UPW1; tra_adv_lim 0 FCT 1 NONE */
void fb200_set_options(int* tra_adv_hor, int* tra_adv_ver, int* tra_adv_lim, double* tra_adv_ph, double* tra_adv_pv,
                       int* w_split, int* i_vert_diff, int* Redi, int* use_sw_pene, int* do_hor_diff, int* clamp_salt,
                       double* is_nonlinfs, double* dt, double* epsilon, int* tracer_id /* num_tracers */);
/* Limits the reference does not have (each raises an error instead of computing something else): nl <= 63 (64-bit level
 * masks per edge); every own column has at least 3 layers (QR4C's boundary statements would overlap); the two triangles
 * of an edge must share at least one level (error 15); num_tracers <= 8; tracer_id in {0, 1, 301, 302, 303} — id 101
 * (bc_surface's prec_rain case, src/oce_ale_tracer.F90:1178) needs an array this ABI does not carry and is refused.
 * Redi = 1 runs diff_part_hor_redi with the slope terms (:934-1077), diff_ver_part_redi_expl (:860-930, call site :288)
 * and the Ki*slope^2 terms of diff_ver_part_impl_ale; it needs slope_tapered in fb200_push_step_state. */
/* biharmonic tracer filter diff_part_bh (src/oce_ale_tracer.F90:1081-1148, call site :321-323): namelist oce_tra
 * smooth_bh_tra, gamma0_tra, gamma1_tra, gamma2_tra (src/oce_modules.F90:138-141).  Off unless called with
 * *smooth_bh_tra != 0.  Applied to every tracer after the vertical diffusion, with the reference's
 * exchange_nod(temporary_ttf) between its two rounds on more than one partition. */
void fb200_set_bh_filter(int* smooth_bh_tra, double* gamma0_tra, double* gamma1_tra, double* gamma2_tra);
/* halo lists of this partition, the reference's com_nod2D and com_elem2D_full (src/gen_modules_partitioning.F90:17-29,
 * read from dist_N/com_infoNNNNN.out, src/oce_mesh.F90:568-640): PE ids 0-based, rptr/sptr 1-based offsets, rlist/slist
 * 1-based local entity numbers.  Only needed on more than one partition; call after fb200_comm_init. */
void fb200_init_halo(int* nod_rPEnum, int* nod_rPE, int* nod_rptr, int* nod_rlist, int* nod_sPEnum, int* nod_sPE, int* nod_sptr,
                     int* nod_slist, int* elem_rPEnum, int* elem_rPE, int* elem_rptr, int* elem_rlist, int* elem_sPEnum,
                     int* elem_sPE, int* elem_sptr, int* elem_slist);
void fb200_init_commit(void);
/* per-step inputs produced by the host (shapes: src/oce_setup_step.F90:205-323, src/oce_ale.F90:94-104):
 * UV(2,nl-1,elems) Wvel/Wvel_e/Wvel_i(nl,N) hnode/hnode_new(nl-1,N) helem(nl-1,myDim_elem2D) Z_3d_n(nl-1,N)
 * zbar_3d_n(nl,N) Kv(nl,N) Ki(nl-1,N) slope_tapered(3,nl-1,N) sw_3d(nl,N); surface fluxes (N).
 * Only what the selected options read crosses PCIe: Wvel_e/Wvel_i when w_split (otherwise Wvel_e == Wvel and Wvel_i == 0,
 * src/oce_ale.F90:2193-2201), slope_tapered when Redi, sw_3d when use_sw_pene. */
void fb200_push_step_state(double* UV, double* Wvel, double* Wvel_e, double* Wvel_i, double* hnode, double* hnode_new,
                           double* helem, double* Z_3d_n, double* zbar_3d_n, double* Kv, double* Ki,
                           double* slope_tapered, double* heat_flux, double* water_flux, double* virtual_salt,
                           double* relax_salt, double* real_salt_flux, double* sw_3d);
/* The same state in the two halves the reference's time step produces it in (oce_timestep_ale, src/oce_ale.F90:2521-2790):
 * early = known before solve_ssh_ale (old thickness hnode and its level depths, mixing coefficients from the mixing
 * step, surface fluxes); late = what the SSH solution yields (update_vel -> UV, compute_hbar_ale / vert_vel_ale ->
 * Wvel*, hnode_new, helem).  fb200_push_state_early only enqueues its copies and returns, so they overlap psolve; the
 * host must leave those arrays untouched until fb200_push_state_late has returned (it waits for both halves). */
/* Two of the early arrays may be passed as NULL (Fortran: c_null_ptr through a type(c_ptr), value dummy) = "the host has
 * not changed it since the library last saw it": hnode = NULL makes the resident hnode_new of the previous step the old
 * thickness (what update_thickness_ale assigns, src/oce_ale.F90:924,936,967); Ki = NULL keeps the resident copy (Ki is set
 * once in src/oce_setup_step.F90:330 and rewritten per step only under Fer_GM/Redi, src/oce_fer_gm.F90:243-333).
 * Z_3d_n / zbar_3d_n = NULL keep the resident copies, which fb200_ale_step_end maintains (with fb200_ale_init, hnode = NULL
 * likewise means the copy fb200_ale_step_end has just updated). */
void fb200_push_state_early(double* hnode, double* Z_3d_n, double* zbar_3d_n, double* Kv, double* Ki, double* slope_tapered,
                            double* heat_flux, double* water_flux, double* virtual_salt, double* relax_salt,
                            double* real_salt_flux, double* sw_3d);
void fb200_push_state_late(double* UV, double* Wvel, double* Wvel_e, double* Wvel_i, double* hnode_new, double* helem);
/* KPP non-local transport in the implicit vertical diffusion (use_kpp_nonlclflx, src/oce_ale_tracer.F90:688-781), per step
 * when the mixing scheme has run: (nl,N) arrays gT/gS = ghats (fesom KPP, mix_scheme_nmb 1/17) or kpp_nonlcltranspT/S
 * (cvmix KPP, 3/37), b2/b3 = blmc(:,:,2:3) or kpp_oblmixc(:,:,2:3); ref_sss, ref_sss_local (src/oce_modules.F90:51-52).
 * *enable = 0 switches the term off (the arrays are then not read). */
void fb200_push_kpp_nonlocal(int* enable, double* gT, double* gS, double* b2, double* b3, double* ref_sss, int* ref_sss_local);
/* climatology relaxation inside the step, relax_to_clim (src/oce_tracer_mod.F90:87-119, called at
 * src/oce_ale_tracer.F90:150-154 between diff_tracers_ale and exchange_nod(tr_arr)): clim_relax (namelist oce_tra),
 * relax2clim(N), Tclim/Sclim(nl-1,N).  Static data: call once after fb200_init_commit; *clim_relax <= 1e-8 = off. */
void fb200_set_clim(double* clim_relax, double* relax2clim, double* Tclim, double* Sclim);
/* relax_zonal_temp of the soufflet channel (src/toy_channel_soufflet.F90:78-103), which solve_tracers_ale calls instead of
 * relax_to_clim when toy_ocean and which_toy == "soufflet" (src/oce_ale_tracer.F90:150-151) — once per pass of the tracer
 * loop and always on tracer 1, as the reference does.  nybins, lat0, dy, tau_inv: the module's parameters (:19-24, :207);
 * coord_nod2D(2,N) in radians; ztem(nl-1,nybins) from compute_zonal_mean (:143-200) — call again whenever the host
 * refreshes it; Tclim(nl-1,N).  *enable = 0 switches back to fb200_set_clim's relaxation. */
void fb200_set_zonal_relax(int* enable, int* nybins, double* lat0, double* dy, double* tau_inv, double* coord_nod2D, double* ztem,
                           double* Tclim);
/* tr_arr / tr_arr_old (nl-1, N, num_tracers) */
void fb200_push_tracers(double* tr_arr, double* tr_arr_old);
void fb200_solve_tracers(void);                         /* all tracers, device resident */
void fb200_pull_tracers(double* tr_arr, double* tr_arr_old);
void fb200_solve_tracers_ale(double* tr_arr, double* tr_arr_old);   /* push + solve + pull */
/* time stepping with device-resident tracers: the step starts from the tracers the previous step left in HBM (first step:
 * fb200_push_tracers) and returns tr_arr to the host; tr_arr_old stays on the device (its only reader is
 * init_tracers_AB, src/oce_tracer_mod.F90:64).  The copy of tracer k overlaps the kernels of tracer k+1. */
void fb200_solve_tracers_ale_resident(double* tr_arr_out);
/* Device-side producers of the tracer step's inputs (any number of partitions: with more than one the four exchanges
 * of oce_ale.F90:1654,1663,2135-2136 run over NCCL, fb200_init_halo required): compute_hbar_ale (src/oce_ale.F90:1585-1676),
 * vert_vel_ale (:1692-2204; linfs, plain zlevel, zstar) and update_thickness_ale (:800-993).  With them the thicknesses,
 * level depths and vertical velocities stay resident; per step the host sends UV (after update_vel; NULL = the resident velocity fb200_update_vel produced) and water_flux and
 * receives hbar(N), ssh_rhs_old(N), dhe(myDim_elem2D).  Call order per step: fb200_ale_step_begin, [fb200_push_state_early
 * for Kv and the surface fluxes], fb200_solve_tracers_ale_resident, fb200_ale_step_end.  Once: fb200_push_step_state with
 * the initial thicknesses, then fb200_ale_init(which_ale 0 linfs | 1 zlevel | 2 zstar, w_max_cfl, hbar). */
void fb200_ale_init(int* which_ale, double* w_max_cfl, double* hbar);
void fb200_ale_step_begin(double* UV, double* water_flux, double* hbar_out, double* ssh_rhs_old_out, double* dhe_out);
void fb200_ale_step_end(void);
/* SSH assembly on the device, for hosts that run the ALE producers above (fb200_ale_init required):
 *   fb200_ssh_assembly_init  once: ssh_stiff as init_stiff_mat_ale leaves it (src/oce_ale.F90:1088-1354) — rowptr(myDim_nod2D+1)
 *                            1-based, colind_loc(nnz) LOCAL 1-based node numbers, values(nnz) —, the host's current
 *                            dhe(myDim_elem2D) and ssh_rhs_old(N) of its last compute_hbar_ale, and g, alpha, theta;
 *   fb200_update_stiff_mat   update_stiff_mat_ale (:1371-1470) from the resident dhe; values_out(nnz) = updated matrix for psolve;
 *   fb200_compute_ssh_rhs    compute_ssh_rhs_ale (:1478-1572): H2D UV_rhs(2,nl-1,elems) of the momentum step, resident UV, helem,
 *                            water_flux, ssh_rhs_old; ssh_rhs_out(N), halo entries exchanged (:1570);
 *   fb200_update_vel         update_vel (src/oce_dyn.F90:101-127) with that UV_rhs and the solution d_eta(N): the resident UV
 *                            becomes the new velocity (then pass UV = NULL to fb200_ale_step_begin), UV_out(2,nl-1,myDim+eDim+eXDim elements)
 *                            returns it to the host's momentum code; eta_n += d_eta stays on the host. */
void fb200_ssh_assembly_init(int* rowptr, int* colind_loc, double* values, double* dhe, double* ssh_rhs_old, double* g, double* alpha,
                             double* theta);
void fb200_update_stiff_mat(double* values_out);
void fb200_compute_ssh_rhs(double* UV_rhs, double* ssh_rhs_out);
void fb200_update_vel(double* d_eta, double* UV_out);
/* min and max of the device-resident tracer *tr_num (1-based) over this partition's own wet cells: what check_blowup and
 * the extrema lines of write_step_info (src/write_step_info.F90:220-504) test; lets a host that keeps the tracers resident
 * monitor them without pulling the fields (combine the ranks with MPI_Allreduce as the reference does) */
void fb200_tracer_extrema(int* tr_num, double* tmin, double* tmax);
/* The same scan WITHOUT ANY ALLOCATION.  It has no counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid allocation).  fb200_tracer_extrema allocates and frees its per-block partials on every call; this variant keeps
 * them in a file-scope `static __device__` array of FB200_EXTREMA_MAX_BLOCKS = 2048 partials + 1 result slot (32 784 bytes
 * per GPU, held from library load until the process exits) and calls no cudaMalloc, cudaFree or allocating library
 * routine.  *nblocks = number of thread blocks (= partials) of the scan, <= 0 for the default (min(1184, myDim_nod2D/256));
 * *nblocks > 2048 is refused with error 28 and the outputs are left at 0.  Every call shares the array: calls that overlap
 * on the GPU (two streams, two host threads) race on it; calls on the library stream are ordered.  Each call overwrites
 * what it reads, so nothing has to be cleared between calls. */
void fb200_tracer_extrema_noalloc(int* tr_num, int* nblocks, double* tmin, double* tmax);
/* Output means of the device-resident tracers (src/io_meandata.F90): fb200_mean_accumulate is update_means (:768-794) for
 * the stream of tracer *tr_num (1-based) — local_values += tr_arr(:,:,tr_num) in the stream's precision, *accuracy = 8
 * (real64) or 4 (real32: the field is rounded to real32 before it is added, as the reference does), addcounter += 1;
 * fb200_mean_fetch is compute_means + clean_meanarrays at output time (:894-900): mean_out(nl-1, myDim_nod2D) =
 * local_values / real(addcounter) in that precision (real64 or real32 array), accumulator and counter cleared. */
void fb200_mean_accumulate(int* tr_num, int* accuracy);
void fb200_mean_fetch(int* tr_num, void* mean_out);
/* debugging / parity: copy an internal (level, node|edge|elem) work field back in the reference layout.
 * name: "fct_LO","adf_h","adf_v","fct_plus","fct_minus","tr_xy_x","tr_xy_y","ttfAB","tstar" ; returns via status */
void fb200_debug_fetch(const char* name, double* out);
void fb200_finalize(void);

/* ---------------------------------------------------------------------------------------------------
 * Multi-GPU: one process (MPI rank) per GPU, NCCL over NVLink replaces MPI_Isend/Irecv of gen_halo_exchange.F90 and
 * the solver's MPI_Allreduce.  Rank 0 obtains an id, the host broadcasts its 128 bytes (MPI_Bcast over
 * MPI_COMM_FESOM), every rank calls fb200_comm_init (collective) after selecting its device (cudaSetDevice). */
void fb200_comm_unique_id(char* id128);
void fb200_comm_init(int* nranks, int* rank, char* id128);
void fb200_comm_finalize(void);

/* self-test: compares the reciprocal-based exact division used inside the kernels with IEEE '/' on *n operand pairs */
void fb200_selftest_div(long long* n, long long* mismatches);

/* 0 = ok; otherwise the code of the first error since the last fb200_clear_status() */
int fb200_status(void);
const char* fb200_status_message(void);
void fb200_clear_status(void);
/* CUDA stream (cudaStream_t passed as void*) every later call works on; NULL = a private stream of the library.
 * Lets the host driver order library work with its own copies/kernels and time it with its own events. */
void fb200_set_stream(void* cuda_stream);
/* per-kernel device timing: *enable = 1 on, 0 off, 2 on + reset totals.  fb200_profile_report writes one line
 * "kernel launches total_ms" per kernel into buf (capacity *cap) and returns the length, -1 if buf is too small */
void fb200_profile(int* enable);
int fb200_profile_report(char* buf, int* cap);
/* number of kernels this library has launched since load (bench.py's gpu_launches) */
long long fb200_kernel_launches(void);
/* device timing of the last fb200_solve_tracers / psolve in milliseconds (CUDA events on the library's stream) */
double fb200_last_ms(const char* what);

#ifdef __cplusplus
}
#endif
#endif
