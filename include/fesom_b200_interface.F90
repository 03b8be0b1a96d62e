! fesom_b200_interface.F90 — Fortran bind(C) interfaces of libfesom_b200.so (include/fesom_b200.h), ready to add to FESOM2's src/.
! The same text as the module in INTEGRATION.md section 2, which also lists the call sites.  psolver_init / psolve are not
! declared here: src/oce_ale.F90:2272-2291 already does.  Not compiled in the authoring image (no Fortran compiler there).
module fesom_b200_interface
  use iso_c_binding
  implicit none
  interface
     subroutine fb200_init_dims(nl, myDim_nod2D, eDim_nod2D, myDim_elem2D, eDim_elem2D, eXDim_elem2D, myDim_edge2D, &
                                num_tracers, fcomm) bind(C, name='fb200_init_dims')
       import :: c_int
       integer(c_int) :: nl, myDim_nod2D, eDim_nod2D, myDim_elem2D, eDim_elem2D, eXDim_elem2D, myDim_edge2D, num_tracers, fcomm
     end subroutine
     subroutine fb200_init_connectivity(edges, edge_tri, elem2D_nodes, nod_in_elem2D_num, nod_in_elem2D, ld, &
                                        edge_up_dn_tri) bind(C, name='fb200_init_connectivity')
       import :: c_int
       integer(c_int) :: edges(*), edge_tri(*), elem2D_nodes(*), nod_in_elem2D_num(*), nod_in_elem2D(*), ld, edge_up_dn_tri(*)
     end subroutine
     subroutine fb200_init_levels(nlevels, ulevels, nlevels_nod2D, ulevels_nod2D, nlevels_nod2D_min, ulevels_nod2D_max, &
                                  nboundary_lay) bind(C, name='fb200_init_levels')
       import :: c_int
       integer(c_int) :: nlevels(*), ulevels(*), nlevels_nod2D(*), ulevels_nod2D(*), nlevels_nod2D_min(*), &
                         ulevels_nod2D_max(*), nboundary_lay(*)
     end subroutine
     subroutine fb200_init_geometry(edge_cross_dxdy, edge_dxdy, elem_cos, elem_area, gradient_sca, area, areasvol, &
                                    zbar_n_bot) bind(C, name='fb200_init_geometry')
       import :: c_double
       real(c_double) :: edge_cross_dxdy(*), edge_dxdy(*), elem_cos(*), elem_area(*), gradient_sca(*), area(*), areasvol(*), zbar_n_bot(*)
     end subroutine
     subroutine fb200_init_halo(n_rn, n_rPE, n_rptr, n_rlist, n_sn, n_sPE, n_sptr, n_slist, &
                                e_rn, e_rPE, e_rptr, e_rlist, e_sn, e_sPE, e_sptr, e_slist) bind(C, name='fb200_init_halo')
       import :: c_int
       integer(c_int) :: n_rn, n_rPE(*), n_rptr(*), n_rlist(*), n_sn, n_sPE(*), n_sptr(*), n_slist(*)
       integer(c_int) :: e_rn, e_rPE(*), e_rptr(*), e_rlist(*), e_sn, e_sPE(*), e_sptr(*), e_slist(*)
     end subroutine
     subroutine fb200_set_options(hor, ver, lim, ph, pv, w_split, i_vert_diff, Redi, use_sw_pene, do_hor_diff, clamp_salt, &
                                  is_nonlinfs, dt, epsilon, tracer_id) bind(C, name='fb200_set_options')
       import :: c_int, c_double
       integer(c_int) :: hor, ver, lim, w_split, i_vert_diff, Redi, use_sw_pene, do_hor_diff, clamp_salt, tracer_id(*)
       real(c_double) :: ph, pv, is_nonlinfs, dt, epsilon
     end subroutine
     subroutine fb200_init_commit() bind(C, name='fb200_init_commit')
     end subroutine
     subroutine fb200_push_step_state(UV, Wvel, Wvel_e, Wvel_i, hnode, hnode_new, helem, Z_3d_n, zbar_3d_n, Kv, Ki, &
          slope_tapered, heat_flux, water_flux, virtual_salt, relax_salt, real_salt_flux, sw_3d) bind(C, name='fb200_push_step_state')
       import :: c_double
       real(c_double) :: UV(*), Wvel(*), Wvel_e(*), Wvel_i(*), hnode(*), hnode_new(*), helem(*), Z_3d_n(*), zbar_3d_n(*), &
                         Kv(*), Ki(*), slope_tapered(*), heat_flux(*), water_flux(*), virtual_salt(*), relax_salt(*), &
                         real_salt_flux(*), sw_3d(*)
     end subroutine
     subroutine fb200_push_state_early(hnode, Z_3d_n, zbar_3d_n, Kv, Ki, slope_tapered, heat_flux, water_flux, virtual_salt, &
          relax_salt, real_salt_flux, sw_3d) bind(C, name='fb200_push_state_early')
       import :: c_double
       real(c_double) :: hnode(*), Z_3d_n(*), zbar_3d_n(*), Kv(*), Ki(*), slope_tapered(*), heat_flux(*), water_flux(*), &
                         virtual_salt(*), relax_salt(*), real_salt_flux(*), sw_3d(*)
     end subroutine
     subroutine fb200_push_state_late(UV, Wvel, Wvel_e, Wvel_i, hnode_new, helem) bind(C, name='fb200_push_state_late')
       import :: c_double
       real(c_double) :: UV(*), Wvel(*), Wvel_e(*), Wvel_i(*), hnode_new(*), helem(*)
     end subroutine
     subroutine fb200_solve_tracers_ale(tr_arr, tr_arr_old) bind(C, name='fb200_solve_tracers_ale')
       import :: c_double
       real(c_double) :: tr_arr(*), tr_arr_old(*)
     end subroutine
     ! resident time stepping: only tr_arr returns; tr_arr_old stays in HBM (first step: fb200_push_tracers)
     subroutine fb200_push_tracers(tr_arr, tr_arr_old) bind(C, name='fb200_push_tracers')
       import :: c_double
       real(c_double) :: tr_arr(*), tr_arr_old(*)
     end subroutine
     subroutine fb200_solve_tracers_ale_resident(tr_arr) bind(C, name='fb200_solve_tracers_ale_resident')
       import :: c_double
       real(c_double) :: tr_arr(*)
     end subroutine
     ! the same entry as fb200_push_state_early with hnode and Ki as C pointers, so that c_null_ptr ("unchanged") can be passed
     subroutine fb200_push_state_early_p(hnode, Z_3d_n, zbar_3d_n, Kv, Ki, slope_tapered, heat_flux, water_flux, virtual_salt, &
          relax_salt, real_salt_flux, sw_3d) bind(C, name='fb200_push_state_early')
       import :: c_double, c_ptr
       type(c_ptr), value :: hnode, Ki
       real(c_double) :: Z_3d_n(*), zbar_3d_n(*), Kv(*), slope_tapered(*), heat_flux(*), water_flux(*), &
                         virtual_salt(*), relax_salt(*), real_salt_flux(*), sw_3d(*)
     end subroutine
     subroutine fb200_comm_unique_id(id) bind(C, name='fb200_comm_unique_id')
       import :: c_char
       character(kind=c_char) :: id(128)
     end subroutine
     subroutine fb200_comm_init(nranks, rank, id) bind(C, name='fb200_comm_init')
       import :: c_int, c_char
       integer(c_int) :: nranks, rank
       character(kind=c_char) :: id(128)
     end subroutine
     ! optional: the ALE producers, the SSH assembly, the soufflet restoring and the output means on the device
     subroutine fb200_ale_init(which_ale, w_max_cfl, hbar) bind(C, name='fb200_ale_init')
       import :: c_int, c_double
       integer(c_int) :: which_ale                  ! 0 linfs, 1 zlevel, 2 zstar
       real(c_double) :: w_max_cfl, hbar(*)
     end subroutine
     subroutine fb200_ale_step_begin(UV, water_flux, hbar, ssh_rhs_old, dhe) bind(C, name='fb200_ale_step_begin')
       import :: c_double, c_ptr
       type(c_ptr), value :: UV                     ! c_loc(UV), or c_null_ptr after fb200_update_vel
       real(c_double) :: water_flux(*), hbar(*), ssh_rhs_old(*), dhe(*)
     end subroutine
     subroutine fb200_ale_step_end() bind(C, name='fb200_ale_step_end')
     end subroutine
     subroutine fb200_ssh_assembly_init(rowptr, colind_loc, values, dhe, ssh_rhs_old, g, alpha, theta) bind(C, name='fb200_ssh_assembly_init')
       import :: c_int, c_double
       integer(c_int) :: rowptr(*), colind_loc(*)
       real(c_double) :: values(*), dhe(*), ssh_rhs_old(*), g, alpha, theta
     end subroutine
     subroutine fb200_update_stiff_mat(values) bind(C, name='fb200_update_stiff_mat')
       import :: c_double
       real(c_double) :: values(*)
     end subroutine
     subroutine fb200_compute_ssh_rhs(UV_rhs, ssh_rhs) bind(C, name='fb200_compute_ssh_rhs')
       import :: c_double
       real(c_double) :: UV_rhs(*), ssh_rhs(*)
     end subroutine
     subroutine fb200_update_vel(d_eta, UV) bind(C, name='fb200_update_vel')
       import :: c_double
       real(c_double) :: d_eta(*), UV(*)
     end subroutine
     subroutine fb200_set_zonal_relax(enable, nybins, lat0, dy, tau_inv, coord_nod2D, ztem, Tclim) bind(C, name='fb200_set_zonal_relax')
       import :: c_int, c_double
       integer(c_int) :: enable, nybins
       real(c_double) :: lat0, dy, tau_inv, coord_nod2D(*), ztem(*), Tclim(*)
     end subroutine
     subroutine fb200_mean_accumulate(tr_num, accuracy) bind(C, name='fb200_mean_accumulate')
       import :: c_int
       integer(c_int) :: tr_num, accuracy           ! accuracy 8 = i_real8, 4 = i_real4 of the stream
     end subroutine
     subroutine fb200_mean_fetch(tr_num, local_values_copy) bind(C, name='fb200_mean_fetch')
       import :: c_int, c_ptr
       integer(c_int) :: tr_num
       type(c_ptr), value :: local_values_copy      ! c_loc of entry%local_values_r8_copy or ..._r4_copy
     end subroutine
     function fb200_status() bind(C, name='fb200_status') result(code)
       import :: c_int
       integer(c_int) :: code
     end function
     ! ---- entry points used less often (declared in include/fesom_b200.h beside the ones above)
     subroutine fb200_set_bh_filter(smooth_bh_tra, gamma0_tra, gamma1_tra, gamma2_tra) bind(C, name='fb200_set_bh_filter')
       import :: c_int, c_double
       integer(c_int) :: smooth_bh_tra
       real(c_double) :: gamma0_tra, gamma1_tra, gamma2_tra
     end subroutine
     subroutine fb200_set_clim(clim_relax, relax2clim, Tclim, Sclim) bind(C, name='fb200_set_clim')
       import :: c_double
       real(c_double) :: clim_relax, relax2clim(*), Tclim(*), Sclim(*)
     end subroutine
     subroutine fb200_push_kpp_nonlocal(enable, gT, gS, b2, b3, ref_sss, ref_sss_local) bind(C, name='fb200_push_kpp_nonlocal')
       import :: c_int, c_double
       integer(c_int) :: enable, ref_sss_local
       real(c_double) :: gT(*), gS(*), b2(*), b3(*), ref_sss
     end subroutine
     subroutine fb200_push_tracers(tr_arr, tr_arr_old) bind(C, name='fb200_push_tracers')
       import :: c_double
       real(c_double) :: tr_arr(*), tr_arr_old(*)
     end subroutine
     subroutine fb200_solve_tracers() bind(C, name='fb200_solve_tracers')
     end subroutine
     subroutine fb200_pull_tracers(tr_arr, tr_arr_old) bind(C, name='fb200_pull_tracers')
       import :: c_double
       real(c_double) :: tr_arr(*), tr_arr_old(*)
     end subroutine
     subroutine fb200_tracer_extrema(tr_num, tmin, tmax) bind(C, name='fb200_tracer_extrema')
       import :: c_int, c_double
       integer(c_int) :: tr_num
       real(c_double) :: tmin, tmax
     end subroutine
     subroutine fb200_tracer_extrema_noalloc(tr_num, nblocks, tmin, tmax) bind(C, name='fb200_tracer_extrema_noalloc')
       import :: c_int, c_double
       integer(c_int) :: tr_num, nblocks
       real(c_double) :: tmin, tmax
     end subroutine
     subroutine fb200_psolver_set_pc(kind) bind(C, name='fb200_psolver_set_pc')
       import :: c_int
       integer(c_int) :: kind
     end subroutine
     subroutine fb200_psolver_set_block(block) bind(C, name='fb200_psolver_set_block')
       import :: c_int
       integer(c_int) :: block
     end subroutine
     subroutine fb200_psolver_set_amg(theta, omega, max_levels, min_rows) bind(C, name='fb200_psolver_set_amg')
       import :: c_int, c_double
       real(c_double) :: theta, omega
       integer(c_int) :: max_levels, min_rows
     end subroutine
     subroutine fb200_psolve_info(id, its, resnorm) bind(C, name='fb200_psolve_info')
       import :: c_int, c_double
       integer(c_int) :: id, its
       real(c_double) :: resnorm
     end subroutine
     subroutine fb200_set_stream(cuda_stream) bind(C, name='fb200_set_stream')
       import :: c_ptr
       type(c_ptr), value :: cuda_stream
     end subroutine
     subroutine fb200_clear_status() bind(C, name='fb200_clear_status')
     end subroutine
     subroutine fb200_comm_finalize() bind(C, name='fb200_comm_finalize')
     end subroutine
     subroutine fb200_finalize() bind(C, name='fb200_finalize')
     end subroutine
  end interface
end module
