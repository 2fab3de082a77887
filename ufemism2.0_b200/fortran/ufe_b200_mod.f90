module ufe_b200_mod
  ! ISO_C_BINDING layer between UFEMISM's Fortran host code and libufe_b200.so (include/ufe_b200.h).
  !
  ! This is the reference-side binding a maintainer adds to src/UPSY/basic/ (see INTEGRATION.md).  It is written
  ! against the reference's own types (type_sparse_matrix_CSR_dp, type_mesh, C%...) and keeps the reference's
  ! init_routine / finalise_routine / crash conventions, so resource_tracking.nc and checksum_logfile.txt stay
  ! comparable.  NOTE: no Fortran compiler exists in the build image (SURVEY F3); this file is syntax-reviewed only,
  ! the runnable stand-in that exercises the same C ABI is ufemism2.0_b200/ufe_b200.py + tests/.
  !
  ! Selected at run time by the new namelist key
  !     choice_linear_algebra_backend_config = 'PETSc' (default) | 'B200'
  ! (+ B200_KSP_type_config, B200_PC_type_config, B200_use_initial_guess_config, B200_stiffness_format_config).

  use, intrinsic :: iso_c_binding
  use precisions, only: dp
  use mpi_basic, only: par
  use mpi_f08, only: MPI_BCAST, MPI_BYTE, MPI_COMM_WORLD
  use call_stack_and_comp_time_tracking, only: init_routine, finalise_routine, crash
  use model_configuration, only: C
  use CSR_sparse_matrix_type, only: type_sparse_matrix_CSR_dp
  use mesh_types, only: type_mesh
  use ice_model_types, only: type_ice_model, type_ice_pc, type_ice_velocity_solver_DIVA, type_ice_velocity_solver_SSA
  use bed_roughness_model_types, only: type_bed_roughness_model

  implicit none

  private

  public :: ufe_b200_initialise, ufe_b200_finalise
  public :: solve_matrix_equation_CSR_B200
  public :: ufe_b200_upload_mesh, ufe_b200_destroy_mesh, remap_velocity_solver_B200
  public :: ufe_b200_fill_config, ufe_b200_config
  public :: solve_DIVA_infinite_slab_B200, solve_SSA_B200, calc_dHi_dt_B200, set_ice_velocities_B200, calc_pc_truncation_error_B200

  type(c_ptr), save :: handle = c_null_ptr

  ! ufe_b200_config of include/ufe_b200.h (same member order)
  type, bind(C) :: ufe_b200_config
    real(c_double) :: visc_it_norm_dUV_tol, visc_it_relax, visc_eff_min, vel_max
    real(c_double) :: stress_balance_PETSc_rtol, stress_balance_PETSc_abstol
    real(c_double) :: Glens_flow_law_exponent, Glens_flow_law_epsilon_sq_0
    real(c_double) :: slid_delta_v, slid_Weertman_m, slid_Budd_q_plastic, slid_Budd_u_threshold, slid_ZI_ut, slid_ZI_p, slid_beta_max
    real(c_double) :: subgrid_friction_exponent_on_B_grid, Hi_min, uniform_Glens_flow_factor
    real(c_double) :: refgeo_idealised_SSA_icestream_Hi, refgeo_idealised_SSA_icestream_dhdx
    real(c_double) :: refgeo_idealised_SSA_icestream_L, refgeo_idealised_SSA_icestream_m
    real(c_double) :: refgeo_idealised_ISMIP_HOM_L
    real(c_double) :: dHi_semiimplicit_fs, dHi_PETSc_rtol, dHi_PETSc_abstol, dt_ice_min, dt_ice_max
    real(c_double) :: ice_density, seawater_density, grav
    integer(c_int32_t) :: visc_it_nit, do_include_SSADIVA_crossterms
    integer(c_int32_t) :: BC_u(4), BC_v(4)
    integer(c_int32_t) :: choice_sliding_law, do_GL_subgrid_friction, choice_ice_integration_method
    integer(c_int32_t) :: BC_H(4)
    integer(c_int32_t) :: ksp_type, pc_type, ksp_maxit, write_diagnostics, ksp_initial_guess_nonzero, stiffness_format
    integer(c_int32_t) :: reserved(4)
  end type ufe_b200_config

  ! ufe_b200_mesh_desc of include/ufe_b200.h
  type, bind(C) :: ufe_b200_mesh_desc
    integer(c_int32_t) :: nV, nTri, nE, nC_mem, nz
    real(c_double)     :: xmin, xmax, ymin, ymax
    type(c_ptr) :: V, nC, C, niTri, iTri, VBI, Tri, TriC, TriBI, TriGC, A, Cw, D_x, D_y, D, VE, ETri, zeta, Tricc
  end type ufe_b200_mesh_desc

  interface

    function c_ufe_b200_get_unique_id( id128) bind(C, name='ufe_b200_get_unique_id') result( ierr)
      import :: c_int, c_char
      character(kind=c_char), intent(inout) :: id128(128)
      integer(c_int) :: ierr
    end function

    function c_ufe_b200_init( device, rank, nranks, id128, h) bind(C, name='ufe_b200_init') result( ierr)
      import :: c_int, c_char, c_ptr
      integer(c_int), value :: device, rank, nranks
      character(kind=c_char), intent(in) :: id128(128)
      type(c_ptr), intent(out) :: h
      integer(c_int) :: ierr
    end function

    function c_ufe_b200_finalize( h) bind(C, name='ufe_b200_finalize') result( ierr)
      import :: c_int, c_ptr
      type(c_ptr), value :: h
      integer(c_int) :: ierr
    end function

    function c_ufe_b200_last_error() bind(C, name='ufe_b200_last_error') result( msg)
      import :: c_ptr
      type(c_ptr) :: msg
    end function

    function c_ufe_b200_ksp_solve( h, m_glob, m_loc, i1, ptr, ind, val, bb, xx, rtol, abstol, maxit, ksp_type, pc_type, n_its) &
      bind(C, name='ufe_b200_ksp_solve') result( ierr)
      import :: c_int, c_int32_t, c_double, c_ptr
      type(c_ptr), value :: h
      integer(c_int), value :: m_glob, m_loc, i1, maxit, ksp_type, pc_type
      integer(c_int32_t), intent(in) :: ptr(*), ind(*)
      real(c_double), intent(in) :: val(*), bb(*)
      real(c_double), intent(inout) :: xx(*)
      real(c_double), value :: rtol, abstol
      integer(c_int), intent(out) :: n_its
      integer(c_int) :: ierr
    end function

    function c_ufe_b200_mesh_create( h, mesh) bind(C, name='ufe_b200_mesh_create') result( ierr)
      import :: c_int, c_ptr, ufe_b200_mesh_desc
      type(c_ptr), value :: h
      type(ufe_b200_mesh_desc), intent(in) :: mesh
      integer(c_int) :: ierr
    end function

    function c_ufe_b200_mesh_destroy( h) bind(C, name='ufe_b200_mesh_destroy') result( ierr)
      import :: c_int, c_ptr
      type(c_ptr), value :: h
      integer(c_int) :: ierr
    end function

    function c_ufe_b200_calc_matrix_operators( h) bind(C, name='ufe_b200_calc_matrix_operators') result( ierr)
      import :: c_int, c_ptr
      type(c_ptr), value :: h
      integer(c_int) :: ierr
    end function

    function c_ufe_b200_solve_SSA_DIVA_linearised( h, cfg, u_b, v_b, N_b, dN_dx_b, dN_dy_b, beta_b, tau_dx_b, tau_dy_b, &
      u_b_prev, v_b_prev, rtol, abstol, n_Axb_its, BC_prescr_mask_b, BC_prescr_u_b, BC_prescr_v_b) &
      bind(C, name='ufe_b200_solve_SSA_DIVA_linearised') result( ierr)
      import :: c_int, c_int32_t, c_double, c_ptr, ufe_b200_config
      type(c_ptr), value :: h
      type(ufe_b200_config), intent(in) :: cfg
      real(c_double), intent(inout) :: u_b(*), v_b(*), u_b_prev(*), v_b_prev(*)
      real(c_double), intent(in) :: N_b(*), dN_dx_b(*), dN_dy_b(*), beta_b(*), tau_dx_b(*), tau_dy_b(*)
      real(c_double), value :: rtol, abstol
      integer(c_int), intent(out) :: n_Axb_its
      integer(c_int32_t), intent(in) :: BC_prescr_mask_b(*)
      real(c_double), intent(in) :: BC_prescr_u_b(*), BC_prescr_v_b(*)
      integer(c_int) :: ierr
    end function

    ! solve_SIA (SIA_main.f90:65-158)
    function c_ufe_b200_solve_SIA( h, cfg, SIA_maximum_diffusivity, Hi, Hs, A_flow, u_3D_b, v_3D_b, D_3D_b, du_dz_3D, dv_dz_3D) &
      bind(C, name='ufe_b200_solve_SIA') result( ierr)
      import :: c_int, c_double, c_ptr, ufe_b200_config
      type(c_ptr), value :: h
      type(ufe_b200_config), intent(in) :: cfg
      real(c_double), value :: SIA_maximum_diffusivity
      real(c_double), intent(in) :: Hi(*), Hs(*), A_flow(*)
      real(c_double), intent(out) :: u_3D_b(*), v_3D_b(*), D_3D_b(*), du_dz_3D(*), dv_dz_3D(*)
      integer(c_int) :: ierr
    end function

    ! multigrid preconditioner of the SSA/DIVA solve (C%B200_PC_type_config = 'multigrid' -> cfg%pc_type = 4)
    function c_ufe_b200_set_mg_params( h, nu_pre, nu_post, omega, overcorr, w_levels, fp32, dense_max) &
      bind(C, name='ufe_b200_set_mg_params') result( ierr)
      import :: c_int, c_double, c_ptr
      type(c_ptr), value :: h
      integer(c_int), value :: nu_pre, nu_post, w_levels, fp32, dense_max
      real(c_double), value :: omega, overcorr
      integer(c_int) :: ierr
    end function

    ! further multigrid knobs by name (C%B200_MG_<name>_config); name = a null-terminated c_char string
    function c_ufe_b200_set_mg_option( h, name, value) bind(C, name='ufe_b200_set_mg_option') result( ierr)
      import :: c_int, c_double, c_ptr, c_char
      type(c_ptr), value :: h
      character(kind=c_char), dimension(*), intent(in) :: name
      real(c_double), value :: value
      integer(c_int) :: ierr
    end function

    ! solve_DIVA_infinite_slab / solve_SSA on the model's arrays (DIVA_solver_infinite_slab.f90:52, SSA_main.f90:88); ice, state = c_loc of
    ! ufe_b200_ice_inputs / ufe_b200_velocity_state filled with c_loc of the model arrays; BC_prescr_* = c_null_ptr when absent
    function c_ufe_b200_solve_DIVA( h, cfg, ice, state, n_visc_its, n_Axb_its, BC_prescr_mask_b, BC_prescr_u_b, BC_prescr_v_b) &
      bind(C, name='ufe_b200_solve_DIVA') result( ierr)
      import :: c_int, c_ptr, ufe_b200_config
      type(c_ptr), value :: h
      type(ufe_b200_config), intent(in) :: cfg
      type(c_ptr), value :: ice, state
      integer(c_int), intent(out) :: n_visc_its, n_Axb_its
      type(c_ptr), value :: BC_prescr_mask_b, BC_prescr_u_b, BC_prescr_v_b
      integer(c_int) :: ierr
    end function
    function c_ufe_b200_solve_SSA( h, cfg, ice, state, n_visc_its, n_Axb_its, BC_prescr_mask_b, BC_prescr_u_b, BC_prescr_v_b) &
      bind(C, name='ufe_b200_solve_SSA') result( ierr)
      import :: c_int, c_ptr, ufe_b200_config
      type(c_ptr), value :: h
      type(ufe_b200_config), intent(in) :: cfg
      type(c_ptr), value :: ice, state
      integer(c_int), intent(out) :: n_visc_its, n_Axb_its
      type(c_ptr), value :: BC_prescr_mask_b, BC_prescr_u_b, BC_prescr_v_b
      integer(c_int) :: ierr
    end function

    ! the same solve on device-resident fields: upload once, solve, download what the host needs
    function c_ufe_b200_velocity_upload( h, ice, state) bind(C, name='ufe_b200_velocity_upload') result( ierr)
      import :: c_int, c_ptr
      type(c_ptr), value :: h, ice, state
      integer(c_int) :: ierr
    end function
    function c_ufe_b200_velocity_solve_resident( h, cfg, approximation, n_visc_its, n_Axb_its) bind(C, name='ufe_b200_velocity_solve_resident') result( ierr)
      import :: c_int, c_ptr, ufe_b200_config
      type(c_ptr), value :: h
      type(ufe_b200_config), intent(in) :: cfg
      integer(c_int), value :: approximation          ! 0 SSA, 1 DIVA
      integer(c_int), intent(out) :: n_visc_its, n_Axb_its
      integer(c_int) :: ierr
    end function
    function c_ufe_b200_velocity_download( h, state) bind(C, name='ufe_b200_velocity_download') result( ierr)
      import :: c_int, c_ptr
      type(c_ptr), value :: h, state
      integer(c_int) :: ierr
    end function

    ! calc_dHi_dt (conservation_of_mass_main.f90:27); io = c_loc of a ufe_b200_thickness_io; finish = 0 leaves the spill-over / final update to
    ! ufe_b200_dHi_dt_finish after the host ran calc_and_apply_spill_over_flux itself
    function c_ufe_b200_calc_dHi_dt( h, cfg, io, dt, finish, n_Axb_its) bind(C, name='ufe_b200_calc_dHi_dt') result( ierr)
      import :: c_int, c_double, c_ptr, ufe_b200_config
      type(c_ptr), value :: h
      type(ufe_b200_config), intent(in) :: cfg
      type(c_ptr), value :: io
      real(c_double), intent(inout) :: dt
      integer(c_int), value :: finish
      integer(c_int), intent(out) :: n_Axb_its
      integer(c_int) :: ierr
    end function
    function c_ufe_b200_dHi_dt_finish( h, Hi, Hi_tplusdt, dt, AMB, dHi_dt) bind(C, name='ufe_b200_dHi_dt_finish') result( ierr)
      import :: c_int, c_double, c_ptr
      type(c_ptr), value :: h
      real(c_double), intent(in) :: Hi(*)
      real(c_double), intent(inout) :: Hi_tplusdt(*)
      real(c_double), value :: dt
      real(c_double), intent(inout) :: AMB(*), dHi_dt(*)
      integer(c_int) :: ierr
    end function
    function c_ufe_b200_thickness_upload( h, io) bind(C, name='ufe_b200_thickness_upload') result( ierr)
      import :: c_int, c_ptr
      type(c_ptr), value :: h, io
      integer(c_int) :: ierr
    end function
    function c_ufe_b200_thickness_download( h, Hi, dHi_dt, divQ) bind(C, name='ufe_b200_thickness_download') result( ierr)
      import :: c_int, c_ptr
      type(c_ptr), value :: h, Hi, dHi_dt, divQ       ! c_loc of the (vi1:vi2) arrays, c_null_ptr = skipped
      integer(c_int) :: ierr
    end function
    function c_ufe_b200_pc_step_resident( h, cfg, approximation, dt, n_visc_its, n_Axb_its_velocity, n_Axb_its_thickness) &
      bind(C, name='ufe_b200_pc_step_resident') result( ierr)
      import :: c_int, c_double, c_ptr, ufe_b200_config
      type(c_ptr), value :: h
      type(ufe_b200_config), intent(in) :: cfg
      integer(c_int), value :: approximation
      real(c_double), value :: dt
      integer(c_int), intent(out) :: n_visc_its, n_Axb_its_velocity, n_Axb_its_thickness
      integer(c_int) :: ierr
    end function

    ! SpMV seam: multiply_CSR_matrix_with_vector_{1D,2D}_wrapper (CSR_matrix_vector_multiplication.f90:24,112) and the 16 operator wrappers
    ! (mesh_disc_apply_operators.f90:18-431)
    function c_ufe_b200_csr_register( h, id, m_glob, n_glob, m_loc, i1, n_loc, j1, ptr, ind, val) bind(C, name='ufe_b200_csr_register') result( ierr)
      import :: c_int, c_int32_t, c_double, c_ptr
      type(c_ptr), value :: h
      integer(c_int), value :: id, m_glob, n_glob, m_loc, i1, n_loc, j1
      integer(c_int32_t), intent(in) :: ptr(*), ind(*)
      real(c_double), intent(in) :: val(*)
      integer(c_int) :: ierr
    end function
    function c_ufe_b200_csr_update_values( h, id, val) bind(C, name='ufe_b200_csr_update_values') result( ierr)
      import :: c_int, c_double, c_ptr
      type(c_ptr), value :: h
      integer(c_int), value :: id
      real(c_double), intent(in) :: val(*)
      integer(c_int) :: ierr
    end function
    function c_ufe_b200_spmv( h, id, xx, yy, ncols) bind(C, name='ufe_b200_spmv') result( ierr)
      import :: c_int, c_double, c_ptr
      type(c_ptr), value :: h
      integer(c_int), value :: id
      real(c_double), intent(in) :: xx(*)
      real(c_double), intent(inout) :: yy(*)
      integer(c_int), value :: ncols
      integer(c_int) :: ierr
    end function
    function c_ufe_b200_csr_release( h, id) bind(C, name='ufe_b200_csr_release') result( ierr)
      import :: c_int, c_ptr
      type(c_ptr), value :: h
      integer(c_int), value :: id
      integer(c_int) :: ierr
    end function
    function c_ufe_b200_apply_operator( h, op, d_in, d_out, ncols) bind(C, name='ufe_b200_apply_operator') result( ierr)
      import :: c_int, c_double, c_ptr
      type(c_ptr), value :: h
      integer(c_int), value :: op                     ! UFE_OP_* in the order of calc_all_matrix_operators_mesh
      real(c_double), intent(in) :: d_in(*)
      real(c_double), intent(inout) :: d_out(*)
      integer(c_int), value :: ncols                  ! 1 (_2D) or mesh%nz (_3D)
      integer(c_int) :: ierr
    end function
    function c_ufe_b200_set_operator( h, op, m_loc, ptr, ind, val) bind(C, name='ufe_b200_set_operator') result( ierr)
      import :: c_int, c_int32_t, c_double, c_ptr
      type(c_ptr), value :: h
      integer(c_int), value :: op, m_loc
      integer(c_int32_t), intent(in) :: ptr(*), ind(*)
      real(c_double), intent(in) :: val(*)
      integer(c_int) :: ierr
    end function
    function c_ufe_b200_get_operator_size( h, op, m_loc, nnz_loc) bind(C, name='ufe_b200_get_operator_size') result( ierr)
      import :: c_int, c_ptr
      type(c_ptr), value :: h
      integer(c_int), value :: op
      integer(c_int), intent(out) :: m_loc, nnz_loc
      integer(c_int) :: ierr
    end function
    function c_ufe_b200_get_operator( h, op, ptr, ind, val) bind(C, name='ufe_b200_get_operator') result( ierr)
      import :: c_int, c_int32_t, c_double, c_ptr
      type(c_ptr), value :: h
      integer(c_int), value :: op
      integer(c_int32_t), intent(inout) :: ptr(*), ind(*)
      real(c_double), intent(inout) :: val(*)
      integer(c_int) :: ierr
    end function
    function c_ufe_b200_mesh_ranges( h, out6) bind(C, name='ufe_b200_mesh_ranges') result( ierr)
      import :: c_int, c_int32_t, c_ptr
      type(c_ptr), value :: h
      integer(c_int32_t), intent(out) :: out6(6)      ! vi1, vi2, ti1, ti2, ei1, ei2: must equal mesh%vi1 ... mesh%ei2
      integer(c_int) :: ierr
    end function
    function c_ufe_b200_kernel_launches( h) bind(C, name='ufe_b200_kernel_launches') result( n)
      import :: c_long_long, c_ptr
      type(c_ptr), value :: h
      integer(c_long_long) :: n
    end function
    subroutine c_ufe_b200_config_defaults( cfg) bind(C, name='ufe_b200_config_defaults')
      import :: ufe_b200_config
      type(ufe_b200_config), intent(out) :: cfg
    end subroutine
    subroutine c_ufe_b200_alter_params_defaults( prm) bind(C, name='ufe_b200_alter_params_defaults')
      import :: c_ptr
      type(c_ptr), value :: prm                       ! c_loc of a ufe_b200_alter_params
    end subroutine

    ! one predictor-corrector step through host arrays (device phases of run_ice_dynamics_model_pc, predictor_corrector_scheme.f90:118-233)
    function c_ufe_b200_pc_step( h, cfg, approximation, dt, dt_prev, ice, state, io, dHi_dt_nm1, n_visc_its, n_Axb_its_velocity, &
      n_Axb_its_thickness) bind(C, name='ufe_b200_pc_step') result( ierr)
      import :: c_int, c_double, c_ptr, ufe_b200_config
      type(c_ptr), value :: h
      type(ufe_b200_config), intent(in) :: cfg
      integer(c_int), value :: approximation
      real(c_double), value :: dt, dt_prev
      type(c_ptr), value :: ice, state, io            ! c_loc of ufe_b200_ice_inputs / _velocity_state / _thickness_io filled with c_loc of the model arrays
      real(c_double), intent(inout) :: dHi_dt_nm1(*)  ! region%ice%pc%dHi_dt_Hnm1_unm1( vi1:vi2) in, %dHi_dt_Hn_un out
      integer(c_int), intent(out) :: n_visc_its, n_Axb_its_velocity, n_Axb_its_thickness
      integer(c_int) :: ierr
    end function

    ! geometry / mask / sub-grid stage (SURVEY 8f-1); logical masks travel as integer(c_int32_t) 0/1 arrays, the structs as c_loc
    function c_ufe_b200_determine_masks( h, cfg, Hi, Hb, SL, masks) bind(C, name='ufe_b200_determine_masks') result( ierr)
      import :: c_int, c_double, c_ptr, ufe_b200_config
      type(c_ptr), value :: h
      type(ufe_b200_config), intent(in) :: cfg
      real(c_double), intent(in) :: Hi(*), Hb(*), SL(*)
      type(c_ptr), value :: masks                     ! c_loc of 11 c_ptr: mask, icefree_land, icefree_ocean, grounded_ice, floating_ice, margin, gl_gr, gl_fl, cf_gr, cf_fl, coastline
      integer(c_int) :: ierr
    end function

    function c_ufe_b200_calc_effective_thickness( h, cfg, Hi, Hb, SL, Hi_eff, fraction_margin) bind(C, name='ufe_b200_calc_effective_thickness') result( ierr)
      import :: c_int, c_double, c_ptr, ufe_b200_config
      type(c_ptr), value :: h
      type(ufe_b200_config), intent(in) :: cfg
      real(c_double), intent(in) :: Hi(*), Hb(*), SL(*)
      real(c_double), intent(out) :: Hi_eff(*), fraction_margin(*)
      integer(c_int) :: ierr
    end function

    function c_ufe_b200_calc_grounded_fractions( h, cfg, choice, nbins, Hi, Hb, SL, dHb, mask_floating_ice, bedrock_cdf, bedrock_cdf_b, &
      fraction_gr, fraction_gr_b) bind(C, name='ufe_b200_calc_grounded_fractions') result( ierr)
      import :: c_int, c_int32_t, c_double, c_ptr, ufe_b200_config
      type(c_ptr), value :: h
      type(ufe_b200_config), intent(in) :: cfg
      integer(c_int), value :: choice, nbins          ! 0 'bilin_interp_TAF', 1 'bedrock_CDF', 2 'bilin_interp_TAF+bedrock_CDF'; C%subgrid_bedrock_cdf_nbins
      real(c_double), intent(in) :: Hi(*), Hb(*), SL(*), dHb(*), bedrock_cdf(*), bedrock_cdf_b(*)
      integer(c_int32_t), intent(in) :: mask_floating_ice(*)
      real(c_double), intent(out) :: fraction_gr(*), fraction_gr_b(*)
      integer(c_int) :: ierr
    end function

    function c_ufe_b200_geometry_resident( h, choice, nbins, dHb, bedrock_cdf, bedrock_cdf_b) bind(C, name='ufe_b200_geometry_resident') result( ierr)
      import :: c_int, c_double, c_ptr
      type(c_ptr), value :: h
      integer(c_int), value :: choice, nbins
      real(c_double), intent(in) :: dHb(*), bedrock_cdf(*), bedrock_cdf_b(*)
      integer(c_int) :: ierr
    end function

    function c_ufe_b200_calc_critical_timestep_adv( h, cfg, do_grounded_only_adv_dt, Hi, mask_floating_ice, u_vav_b, v_vav_b, dt_crit_adv) &
      bind(C, name='ufe_b200_calc_critical_timestep_adv') result( ierr)
      import :: c_int, c_int32_t, c_double, c_ptr, ufe_b200_config
      type(c_ptr), value :: h
      type(ufe_b200_config), intent(in) :: cfg
      integer(c_int), value :: do_grounded_only_adv_dt
      real(c_double), intent(in) :: Hi(*)
      integer(c_int32_t), intent(in) :: mask_floating_ice(*)
      type(c_ptr), value :: u_vav_b, v_vav_b          ! c_loc( ice%u_vav_b) / c_null_ptr = the velocities of the last device solve
      real(c_double), intent(out) :: dt_crit_adv
      integer(c_int) :: ierr
    end function

    function c_ufe_b200_calc_critical_timestep_SIA( h, cfg, Hi, D_3D_b, dt_crit_SIA) bind(C, name='ufe_b200_calc_critical_timestep_SIA') result( ierr)
      import :: c_int, c_double, c_ptr, ufe_b200_config
      type(c_ptr), value :: h
      type(ufe_b200_config), intent(in) :: cfg
      real(c_double), intent(in) :: Hi(*), D_3D_b(*)
      real(c_double), intent(out) :: dt_crit_SIA
      integer(c_int) :: ierr
    end function

    function c_ufe_b200_alter_ice_thickness( h, cfg, prm, time, masks, Hi_old, Hb, SL, refgeo_Hi, refgeo_Hb, Ti_hom, Hi_new) &
      bind(C, name='ufe_b200_alter_ice_thickness') result( ierr)
      import :: c_int, c_double, c_ptr, ufe_b200_config
      type(c_ptr), value :: h
      type(ufe_b200_config), intent(in) :: cfg
      type(c_ptr), value :: prm, masks                ! c_loc of a ufe_b200_alter_params / ufe_b200_masks mirror
      real(c_double), value :: time
      real(c_double), intent(in) :: Hi_old(*), Hb(*), SL(*), refgeo_Hi(*), refgeo_Hb(*), Ti_hom(*)
      real(c_double), intent(inout) :: Hi_new(*)
      integer(c_int) :: ierr
    end function

    function c_ufe_b200_calc_and_apply_spill_over_flux( h, masks, Hi_eff, u_vav_b, v_vav_b, dt, Hi_new, Qspill) &
      bind(C, name='ufe_b200_calc_and_apply_spill_over_flux') result( ierr)
      import :: c_int, c_double, c_ptr
      type(c_ptr), value :: h, masks, u_vav_b, v_vav_b
      real(c_double), intent(in) :: Hi_eff(*)
      real(c_double), value :: dt
      real(c_double), intent(inout) :: Hi_new(*)
      real(c_double), intent(out) :: Qspill(*)
      integer(c_int) :: ierr
    end function

    function c_ufe_b200_set_state_resident( h, on) bind(C, name='ufe_b200_set_state_resident') result( ierr)
      import :: c_int, c_ptr
      type(c_ptr), value :: h
      integer(c_int), value :: on
      integer(c_int) :: ierr
    end function

    ! region%ice%Hi apart from Hi_prev, region%ice%dHi_dt as pc%dHi_dt_Hi_nm1_u_nm1 (ice_dynamics_main.f90:115-140); wt_next of the last action time
    function c_ufe_b200_pc_reference_loop( h, on, wt_next_of_last_action) bind(C, name='ufe_b200_pc_reference_loop') result( ierr)
      import :: c_int, c_ptr, c_double
      type(c_ptr), value :: h
      integer(c_int), value :: on
      real(c_double), value :: wt_next_of_last_action
      integer(c_int) :: ierr
    end function

    ! region%ice%Hi as the host holds it, for the predictor of the next one-call step (predictor_corrector_scheme.f90:129)
    function c_ufe_b200_pc_set_model_thickness( h, Hi_model) bind(C, name='ufe_b200_pc_set_model_thickness') result( ierr)
      import :: c_int, c_ptr
      type(c_ptr), value :: h, Hi_model
      integer(c_int) :: ierr
    end function

    ! alter_ice_thickness inside the one-call steps (predictor_corrector_scheme.f90:145, 233); prm = c_loc of a ufe_b200_alter_params or c_null_ptr
    function c_ufe_b200_alter_resident( h, prm, time, refgeo_Hi, refgeo_Hb, Ti_hom) bind(C, name='ufe_b200_alter_resident') result( ierr)
      import :: c_int, c_ptr, c_double
      type(c_ptr), value :: h, prm
      real(c_double), value :: time
      type(c_ptr), value :: refgeo_Hi, refgeo_Hb, Ti_hom
      integer(c_int) :: ierr
    end function

    ! pc%Hi_star_np1, pc%dHi_dt_Hi_n_u_n, the next pc%dHi_dt_Hi_nm1_u_nm1 and region%ice%Hi of the last one-call step (c_null_ptr = skip)
    function c_ufe_b200_pc_download( h, Hi_star_np1, dHi_dt_Hi_n_u_n, dHi_dt_nm1_next, Hi_model) bind(C, name='ufe_b200_pc_download') result( ierr)
      import :: c_int, c_ptr
      type(c_ptr), value :: h, Hi_star_np1, dHi_dt_Hi_n_u_n, dHi_dt_nm1_next, Hi_model
      integer(c_int) :: ierr
    end function

    ! apply_grounded_fractions_to_bed_roughness (basal_conditions/sliding_laws.f90:568-636); masks = c_loc of a ufe_b200_masks
    function c_ufe_b200_apply_grounded_fractions_to_bed_roughness( h, Hi, Hs_slope, fraction_gr, masks, bed_roughness, bed_roughness_applied) &
      bind(C, name='ufe_b200_apply_grounded_fractions_to_bed_roughness') result( ierr)
      import :: c_int, c_ptr
      type(c_ptr), value :: h, Hi, Hs_slope, fraction_gr, masks, bed_roughness, bed_roughness_applied
      integer(c_int) :: ierr
    end function

    ! calc_pc_truncation_error + the count of unstable vertices (time_stepping/predictor_corrector_scheme.f90:327-361, 254-279);
    ! logical masks as integer(c_int32_t) 0/1 copies; tau_np1, tau_n_guilty may be c_null_ptr
    function c_ufe_b200_calc_pc_truncation_error( h, zeta_t, dt_n, pc_eta_min, pc_epsilon, Hi_np1, Hi_star_np1, mask_grounded_ice, mask_gl_gr, &
      fraction_gr, tau_np1, tau_n_guilty, eta_np1, n_tot, n_guilty) bind(C, name='ufe_b200_calc_pc_truncation_error') result( ierr)
      import :: c_int, c_int32_t, c_double, c_ptr
      type(c_ptr), value :: h
      real(c_double), value :: zeta_t, dt_n, pc_eta_min, pc_epsilon
      type(c_ptr), value :: Hi_np1, Hi_star_np1, mask_grounded_ice, mask_gl_gr, fraction_gr, tau_np1, tau_n_guilty
      real(c_double), intent(out) :: eta_np1
      integer(c_int32_t), intent(out) :: n_tot, n_guilty
      integer(c_int) :: ierr
    end function

    ! No counterpart in the reference (added at the request of the run's operators): the same computation on the n vertices passed, with all device
    ! memory in a fixed array inside the library (n <= UFE_B200_NOALLOC_MAX_VERTICES = 262144); allocates nothing; results of this rank only
    function c_ufe_b200_calc_pc_truncation_error_noalloc( h, n, zeta_t, dt_n, pc_eta_min, pc_epsilon, Hi_np1, Hi_star_np1, mask_grounded_ice, &
      mask_gl_gr, fraction_gr, tau_np1, tau_n_guilty, eta_np1, n_tot, n_guilty) bind(C, name='ufe_b200_calc_pc_truncation_error_noalloc') &
      result( ierr)
      import :: c_int, c_int32_t, c_double, c_ptr
      type(c_ptr), value :: h
      integer(c_int), value :: n
      real(c_double), value :: zeta_t, dt_n, pc_eta_min, pc_epsilon
      type(c_ptr), value :: Hi_np1, Hi_star_np1, mask_grounded_ice, mask_gl_gr, fraction_gr, tau_np1, tau_n_guilty
      real(c_double), intent(out) :: eta_np1
      integer(c_int32_t), intent(out) :: n_tot, n_guilty
      integer(c_int) :: ierr
    end function

    ! remap_DIVA_solver / remap_SSA_solver (DIVA_main.f90:115-222, SSA_main.f90:247-311) on a device-resident solver state: h_old holds the old mesh
    ! and state, h_new the new mesh; map_ptr / map_ind / map_val = c_loc of the rank's rows of the Atlas map (type_map%M) as CSR
    function c_ufe_b200_remap_velocity_solver( h_old, h_new, approximation, m_loc, map_ptr, map_ind, map_val) &
      bind(C, name='ufe_b200_remap_velocity_solver') result( ierr)
      import :: c_int, c_ptr
      type(c_ptr), value :: h_old, h_new
      integer(c_int), value :: approximation, m_loc
      type(c_ptr), value :: map_ptr, map_ind, map_val
      integer(c_int) :: ierr
    end function

    ! set_ice_velocities_to_{SIA,SSA,SIASSA,DIVA}_results (conservation_of_momentum_main.f90:311-511); out = c_loc of a ufe_b200_ice_velocities
    ! (c_loc's of ice%u_3D_b, v_3D_b, du_dx_3D, du_dy_3D, du_dz_3D, dv_dx_3D, dv_dy_3D, dv_dz_3D, dw_dx_3D, dw_dy_3D; c_null_ptr = skipped);
    ! approximation: 0 'SSA', 1 'DIVA', 2 'SIA', 3 'SIA/SSA'; hybrid_scheme: 0 'add', 1 'add_SIA_reduced'
    function c_ufe_b200_set_ice_velocities( h, approximation, hybrid_scheme, out) bind(C, name='ufe_b200_set_ice_velocities') result( ierr)
      import :: c_int, c_ptr
      type(c_ptr), value :: h
      integer(c_int), value :: approximation, hybrid_scheme
      type(c_ptr), value :: out
      integer(c_int) :: ierr
    end function

    ! calc_secondary_velocities (conservation_of_momentum_main.f90:177-246); out = c_loc of a ufe_b200_secondary_velocities
    function c_ufe_b200_calc_secondary_velocities( h, u_3D_b, v_3D_b, out) bind(C, name='ufe_b200_calc_secondary_velocities') result( ierr)
      import :: c_int, c_ptr
      type(c_ptr), value :: h
      type(c_ptr), value :: u_3D_b, v_3D_b            ! c_loc( ice%u_3D_b) / c_null_ptr to use the fields of the last device solve
      type(c_ptr), value :: out                       ! 21 c_loc's in the order of include/ufe_b200.h
      integer(c_int) :: ierr
    end function

    ! calc_vertical_velocities + calc_dw_dz (conservation_of_mass/vertical_velocities.f90:23-243); io = c_loc of a ufe_b200_vertical_velocity_io
    function c_ufe_b200_calc_vertical_velocities( h, cfg, io) bind(C, name='ufe_b200_calc_vertical_velocities') result( ierr)
      import :: c_int, c_ptr, ufe_b200_config
      type(c_ptr), value :: h
      type(ufe_b200_config), intent(in) :: cfg
      type(c_ptr), value :: io
      integer(c_int) :: ierr
    end function

  end interface

contains

  subroutine check( ierr, where)
    ! Map a non-zero status to the reference's crash() (src/UPSY/basic/crash_mod.f90:17-34)
    integer(c_int),   intent(in) :: ierr
    character(len=*), intent(in) :: where
    character(len=1024) :: msg
    character(kind=c_char), pointer :: cmsg(:)
    integer :: i
    if (ierr == 0) return
    msg = ''
    call c_f_pointer( c_ufe_b200_last_error(), cmsg, [1024])
    do i = 1, 1024
      if (cmsg( i) == c_null_char) exit
      msg( i:i) = cmsg( i)
    end do
    call crash( trim( where) // ': ' // trim( msg))
  end subroutine check

  subroutine ufe_b200_initialise
    ! Called once after initialise_parallelisation (src/UFEMISM/main/UFEMISM_program.f90:55): one rank <-> one GPU.
    character(len=1024), parameter :: routine_name = 'ufe_b200_initialise'
    character(kind=c_char) :: id128(128)
    integer :: ierr
    call init_routine( routine_name)
    id128 = c_null_char
    if (par%n > 1) then
      if (par%primary) call check( c_ufe_b200_get_unique_id( id128), routine_name)
      call MPI_BCAST( id128, 128, MPI_BYTE, 0, MPI_COMM_WORLD, ierr)
    end if
    call check( c_ufe_b200_init( -1_c_int, int( par%i, c_int), int( par%n, c_int), id128, handle), routine_name)
    call finalise_routine( routine_name)
  end subroutine ufe_b200_initialise

  subroutine ufe_b200_finalise
    character(len=1024), parameter :: routine_name = 'ufe_b200_finalise'
    call init_routine( routine_name)
    if (c_associated( handle)) call check( c_ufe_b200_finalize( handle), routine_name)
    handle = c_null_ptr
    call finalise_routine( routine_name)
  end subroutine ufe_b200_finalise

  subroutine solve_matrix_equation_CSR_B200( AA, bb, xx, rtol, abstol, n_Axb_its)
    ! Drop-in for solve_matrix_equation_CSR_PETSc (src/UPSY/basic/petsc_basic.f90:32-33); same arguments, same meaning:
    ! AA finalised, xx ignored on input (zero initial guess), n_Axb_its = iteration count, non-convergence silent.
    type(type_sparse_matrix_CSR_dp),    intent(in   ) :: AA
    real(dp), dimension(:),             intent(in   ) :: bb
    real(dp), dimension(:),             intent(inout) :: xx
    real(dp),                           intent(in   ) :: rtol, abstol
    integer,                            intent(  out) :: n_Axb_its
    character(len=1024), parameter :: routine_name = 'solve_matrix_equation_CSR_B200'
    integer(c_int) :: its
    call init_routine( routine_name)
    if (.not. AA%is_finalised) call crash('A should be finalised before calling solve_matrix_equation_CSR_B200')
    ! AA%ptr( AA%i1: AA%i2+1) holds 1-based offsets into the local ind/val (CSR_matrix_basics.f90:63-78)
    call check( c_ufe_b200_ksp_solve( handle, int( AA%m, c_int), int( AA%m_loc, c_int), int( AA%i1, c_int), &
      AA%ptr( AA%i1: AA%i2+1), AA%ind, AA%val, bb, xx, rtol, abstol, 2000_c_int, C%B200_KSP_type, C%B200_PC_type, its), routine_name)
    n_Axb_its = int( its)
    call finalise_routine( routine_name)
  end subroutine solve_matrix_equation_CSR_B200

  subroutine ufe_b200_upload_mesh( mesh)
    ! Called at the end of calc_all_secondary_mesh_data / instead of calc_all_matrix_operators_mesh
    ! (src/UPSY/mesh/mesh_secondary.f90:31-70, mesh_disc_calc_matrix_operators_2D.f90:26-59).
    type(type_mesh), target, intent(in) :: mesh
    character(len=1024), parameter :: routine_name = 'ufe_b200_upload_mesh'
    type(ufe_b200_mesh_desc) :: d
    call init_routine( routine_name)
    d%nV = mesh%nV; d%nTri = mesh%nTri; d%nE = mesh%nE; d%nC_mem = mesh%nC_mem; d%nz = mesh%nz
    d%xmin = mesh%xmin; d%xmax = mesh%xmax; d%ymin = mesh%ymin; d%ymax = mesh%ymax
    d%V = c_loc( mesh%V); d%nC = c_loc( mesh%nC); d%C = c_loc( mesh%C); d%niTri = c_loc( mesh%niTri); d%iTri = c_loc( mesh%iTri)
    d%VBI = c_loc( mesh%VBI); d%Tri = c_loc( mesh%Tri); d%TriC = c_loc( mesh%TriC); d%TriBI = c_loc( mesh%TriBI)
    d%TriGC = c_loc( mesh%TriGC); d%A = c_loc( mesh%A); d%Cw = c_loc( mesh%Cw); d%D_x = c_loc( mesh%D_x); d%D_y = c_loc( mesh%D_y)
    d%D = c_loc( mesh%D); d%VE = c_loc( mesh%VE); d%ETri = c_loc( mesh%ETri); d%zeta = c_loc( mesh%zeta)
    d%Tricc = c_loc( mesh%Tricc)
    call check( c_ufe_b200_mesh_create( handle, d), routine_name)
    call check( c_ufe_b200_calc_matrix_operators( handle), routine_name)
    call finalise_routine( routine_name)
  end subroutine ufe_b200_upload_mesh

  subroutine ufe_b200_destroy_mesh
    ! Called from the mesh finaliser and before update_mesh -> remap_velocity_solver (UFEMISM_main_model.f90:1211)
    if (c_associated( handle)) call check( c_ufe_b200_mesh_destroy( handle), 'ufe_b200_destroy_mesh')
  end subroutine ufe_b200_destroy_mesh

  subroutine remap_velocity_solver_B200( mesh_new, approximation, M_map)
    ! Drop-in for the remap_DIVA_solver / remap_SSA_solver branch of remap_velocity_solver (called after update_mesh, UFEMISM_main_model.f90:1211;
    ! DIVA_main.f90:115-222, SSA_main.f90:247-311) when the solver state lives on the device: a second handle takes the new mesh, the re-used
    ! fields go old b-grid -> old a-grid -> Atlas map -> border fix -> new b-grid on the device, the old handle is released and the new one
    ! becomes the module's handle.  M_map = the rank's rows (mesh_new%vi1:vi2) of the map from mesh_old's to mesh_new's vertices as a finalised
    ! type_sparse_matrix_CSR_dp (mat_petsc2CSR( map%M, M_map), petsc_basic.f90:366; the map itself is still made by create_map_from_mesh_to_mesh_*).
    ! approximation: 0 'SSA', 1 'DIVA'.
    type(type_mesh),                 target, intent(in) :: mesh_new
    integer,                                 intent(in) :: approximation
    type(type_sparse_matrix_CSR_dp), target, intent(in) :: M_map
    character(len=1024), parameter :: routine_name = 'remap_velocity_solver_B200'
    character(kind=c_char) :: id128(128)
    type(c_ptr) :: handle_old
    integer :: ierr
    call init_routine( routine_name)
    handle_old = handle
    id128 = c_null_char
    if (par%n > 1) then
      if (par%primary) call check( c_ufe_b200_get_unique_id( id128), routine_name)
      call MPI_BCAST( id128, 128, MPI_BYTE, 0, MPI_COMM_WORLD, ierr)
    end if
    call check( c_ufe_b200_init( -1_c_int, int( par%i, c_int), int( par%n, c_int), id128, handle), routine_name)
    call ufe_b200_upload_mesh( mesh_new)      ! on the new handle: ufe_b200_mesh_create + ufe_b200_calc_matrix_operators
    call check( c_ufe_b200_remap_velocity_solver( handle_old, handle, int( approximation, c_int), int( M_map%m_loc, c_int), &
      c_loc( M_map%ptr( M_map%i1)), c_loc( M_map%ind), c_loc( M_map%val)), routine_name)   ! ptr( i1: i2+1): 1-based offsets into the local ind / val
    call check( c_ufe_b200_finalize( handle_old), routine_name)
    call finalise_routine( routine_name)
  end subroutine remap_velocity_solver_B200

  subroutine solve_DIVA_infinite_slab_B200( mesh, ice, bed_roughness, DIVA, n_visc_its, n_Axb_its, &
    BC_prescr_mask_b, BC_prescr_u_b, BC_prescr_v_b)
    ! Drop-in for solve_DIVA_infinite_slab (src/UFEMISM/ice_dynamics/conservation_of_momentum/SSA_DIVA/DIVA_solver_infinite_slab.f90:52-53): same
    ! arguments, same meaning; the fields of the solver object are filled exactly as the reference fills them, so the checksum() calls that follow
    ! (:225-226) and restart output keep working.  ice%A_flow must be up to date (calc_ice_rheology_Glen stays on the host).
    type(type_mesh),                                   intent(in   ) :: mesh
    type(type_ice_model),                      target, intent(inout) :: ice
    type(type_bed_roughness_model),            target, intent(in   ) :: bed_roughness
    type(type_ice_velocity_solver_DIVA),       target, intent(inout) :: DIVA
    integer,                                           intent(  out) :: n_visc_its, n_Axb_its
    integer,  dimension(mesh%ti1:mesh%ti2), optional, target, intent(in) :: BC_prescr_mask_b
    real(dp), dimension(mesh%ti1:mesh%ti2), optional, target, intent(in) :: BC_prescr_u_b, BC_prescr_v_b
    character(len=1024), parameter :: routine_name = 'solve_DIVA_infinite_slab_B200'
    type(ufe_b200_config) :: cfg
    type(c_ptr), target :: ice_in(10), state(33)
    integer(c_int32_t), dimension(mesh%vi1:mesh%vi2), target :: mask_grounded_ice, mask_icefree_land
    type(c_ptr) :: bm, bu, bv
    integer(c_int) :: nv, na
    call init_routine( routine_name)
    if (present( BC_prescr_mask_b) .or. present( BC_prescr_u_b) .or. present( BC_prescr_v_b)) then
      if (.not. (present( BC_prescr_mask_b) .and. present( BC_prescr_u_b) .and. present( BC_prescr_v_b))) then
        call crash('need to provide prescribed u,v fields and mask!')
      end if
      bm = c_loc( BC_prescr_mask_b); bu = c_loc( BC_prescr_u_b); bv = c_loc( BC_prescr_v_b)
    else
      bm = c_null_ptr; bu = c_null_ptr; bv = c_null_ptr
    end if
    call ufe_b200_fill_config( cfg)
    mask_grounded_ice = merge( 1_c_int32_t, 0_c_int32_t, ice%mask_grounded_ice( mesh%vi1:mesh%vi2))
    mask_icefree_land = merge( 1_c_int32_t, 0_c_int32_t, ice%mask_icefree_land( mesh%vi1:mesh%vi2))
    ! member order of ufe_b200_ice_inputs (include/ufe_b200.h)
    ice_in = [c_loc( ice%Hi), c_loc( ice%Hs), c_loc( ice%A_flow), c_loc( ice%fraction_gr_b), c_loc( bed_roughness%till_friction_angle), &
              c_loc( bed_roughness%beta_sq), c_loc( bed_roughness%alpha_sq), c_loc( ice%effective_pressure), c_loc( mask_grounded_ice), &
              c_loc( mask_icefree_land)]
    ! member order of ufe_b200_velocity_state (include/ufe_b200.h)
    state = [c_loc( DIVA%u_vav_b), c_loc( DIVA%v_vav_b), c_loc( DIVA%u_base_b), c_loc( DIVA%v_base_b), c_loc( DIVA%tau_bx_b), c_loc( DIVA%tau_by_b), &
             c_loc( DIVA%eta_3D_b), c_loc( DIVA%u_b_prev), c_loc( DIVA%v_b_prev), c_loc( DIVA%tau_dx_b), c_loc( DIVA%tau_dy_b), &
             c_loc( DIVA%u_3D_b), c_loc( DIVA%v_3D_b), c_loc( DIVA%du_dx_a), c_loc( DIVA%du_dy_a), c_loc( DIVA%dv_dx_a), c_loc( DIVA%dv_dy_a), &
             c_loc( DIVA%du_dz_3D_a), c_loc( DIVA%dv_dz_3D_a), c_loc( DIVA%eta_3D_a), c_loc( DIVA%eta_vav_a), c_loc( DIVA%N_a), &
             c_loc( DIVA%N_b), c_loc( DIVA%dN_dx_b), c_loc( DIVA%dN_dy_b), c_loc( DIVA%F1_3D_a), c_loc( DIVA%F2_3D_a), &
             c_loc( DIVA%F1_3D_b), c_loc( DIVA%F2_3D_b), c_loc( ice%basal_friction_coefficient), c_loc( DIVA%basal_friction_coefficient_b), &
             c_loc( DIVA%beta_eff_a), c_loc( DIVA%beta_eff_b)]
    call check( c_ufe_b200_solve_DIVA( handle, cfg, c_loc( ice_in), c_loc( state), nv, na, bm, bu, bv), routine_name)
    n_visc_its = int( nv); n_Axb_its = int( na)
    call finalise_routine( routine_name)
  end subroutine solve_DIVA_infinite_slab_B200

  subroutine solve_SSA_B200( mesh, ice, bed_roughness, SSA, n_visc_its, n_Axb_its, BC_prescr_mask_b, BC_prescr_u_b, BC_prescr_v_b)
    ! Drop-in for solve_SSA (src/UFEMISM/ice_dynamics/conservation_of_momentum/SSA_DIVA/SSA_main.f90:88-89): same arguments, same meaning.
    ! The SSA solver object has no 3-D fields: those members of ufe_b200_velocity_state stay c_null_ptr and are skipped by the library.
    type(type_mesh),                                   intent(in   ) :: mesh
    type(type_ice_model),                      target, intent(inout) :: ice
    type(type_bed_roughness_model),            target, intent(in   ) :: bed_roughness
    type(type_ice_velocity_solver_SSA),        target, intent(inout) :: SSA
    integer,                                           intent(  out) :: n_visc_its, n_Axb_its
    integer,  dimension(mesh%ti1:mesh%ti2), optional, target, intent(in) :: BC_prescr_mask_b
    real(dp), dimension(mesh%ti1:mesh%ti2), optional, target, intent(in) :: BC_prescr_u_b, BC_prescr_v_b
    character(len=1024), parameter :: routine_name = 'solve_SSA_B200'
    type(ufe_b200_config) :: cfg
    type(c_ptr), target :: ice_in_ssa(10), state_ssa(33)
    integer(c_int32_t), dimension(mesh%vi1:mesh%vi2), target :: mask_grounded_ice, mask_icefree_land
    type(c_ptr) :: bm, bu, bv
    integer(c_int) :: nv, na
    call init_routine( routine_name)
    if (present( BC_prescr_mask_b) .or. present( BC_prescr_u_b) .or. present( BC_prescr_v_b)) then
      if (.not. (present( BC_prescr_mask_b) .and. present( BC_prescr_u_b) .and. present( BC_prescr_v_b))) then
        call crash('need to provide prescribed u,v fields and mask!')
      end if
      bm = c_loc( BC_prescr_mask_b); bu = c_loc( BC_prescr_u_b); bv = c_loc( BC_prescr_v_b)
    else
      bm = c_null_ptr; bu = c_null_ptr; bv = c_null_ptr
    end if
    call ufe_b200_fill_config( cfg)
    cfg%stress_balance_PETSc_rtol = SSA%PETSc_rtol; cfg%stress_balance_PETSc_abstol = SSA%PETSc_abstol   ! SSA_main.f90:224
    mask_grounded_ice = merge( 1_c_int32_t, 0_c_int32_t, ice%mask_grounded_ice( mesh%vi1:mesh%vi2))
    mask_icefree_land = merge( 1_c_int32_t, 0_c_int32_t, ice%mask_icefree_land( mesh%vi1:mesh%vi2))
    ice_in_ssa = [c_loc( ice%Hi), c_loc( ice%Hs), c_loc( ice%A_flow), c_loc( ice%fraction_gr_b), c_loc( bed_roughness%till_friction_angle), &
                  c_loc( bed_roughness%beta_sq), c_loc( bed_roughness%alpha_sq), c_loc( ice%effective_pressure), c_loc( mask_grounded_ice), &
                  c_loc( mask_icefree_land)]
    ! member order of ufe_b200_velocity_state; SSA%u_b, v_b travel as u_vav_b, v_vav_b, SSA%eta_a as eta_vav_a
    state_ssa = [c_loc( SSA%u_b), c_loc( SSA%v_b), c_null_ptr, c_null_ptr, c_null_ptr, c_null_ptr, &
                 c_null_ptr, c_loc( SSA%u_b_prev), c_loc( SSA%v_b_prev), c_loc( SSA%tau_dx_b), c_loc( SSA%tau_dy_b), &
                 c_null_ptr, c_null_ptr, c_loc( SSA%du_dx_a), c_loc( SSA%du_dy_a), c_loc( SSA%dv_dx_a), c_loc( SSA%dv_dy_a), &
                 c_null_ptr, c_null_ptr, c_null_ptr, c_loc( SSA%eta_a), c_loc( SSA%N_a), &
                 c_loc( SSA%N_b), c_loc( SSA%dN_dx_b), c_loc( SSA%dN_dy_b), c_null_ptr, c_null_ptr, &
                 c_null_ptr, c_null_ptr, c_loc( ice%basal_friction_coefficient), c_loc( SSA%basal_friction_coefficient_b), &
                 c_null_ptr, c_null_ptr]
    call check( c_ufe_b200_solve_SSA( handle, cfg, c_loc( ice_in_ssa), c_loc( state_ssa), nv, na, bm, bu, bv), routine_name)
    n_visc_its = int( nv); n_Axb_its = int( na)
    call finalise_routine( routine_name)
  end subroutine solve_SSA_B200

  subroutine calc_dHi_dt_B200( mesh, ice, Hi, Hb, SL, u_vav_b, v_vav_b, SMB, BMB, LMB, AMB, fraction_margin, mask_noice, dt, dHi_dt, Hi_tplusdt, &
    divQ, dHi_dt_target, BC_prescr_mask, BC_prescr_Hi)
    ! Drop-in for calc_dHi_dt (src/UFEMISM/ice_dynamics/conservation_of_mass/conservation_of_mass_main.f90:27-28): same arguments, same meaning
    ! (explicit or semi-implicit by C%choice_ice_integration_method, thickness BCs, mask_noice, flux-limited dt, the warning on negative thickness
    ! stays with the caller's checksum).
    type(type_mesh),                                intent(in   ) :: mesh
    type(type_ice_model),                           intent(inout) :: ice
    real(dp), dimension(mesh%vi1:mesh%vi2), target, intent(in   ) :: Hi, Hb, SL
    real(dp), dimension(mesh%ti1:mesh%ti2), target, intent(in   ) :: u_vav_b, v_vav_b
    real(dp), dimension(mesh%vi1:mesh%vi2), target, intent(in   ) :: SMB, BMB, LMB
    real(dp), dimension(mesh%vi1:mesh%vi2), target, intent(inout) :: AMB
    real(dp), dimension(mesh%vi1:mesh%vi2), target, intent(in   ) :: fraction_margin
    logical,  dimension(mesh%vi1:mesh%vi2),         intent(in   ) :: mask_noice
    real(dp),                                       intent(inout) :: dt
    real(dp), dimension(mesh%vi1:mesh%vi2), target, intent(  out) :: dHi_dt, Hi_tplusdt, divQ
    real(dp), dimension(mesh%vi1:mesh%vi2), target, intent(in   ) :: dHi_dt_target
    integer,  dimension(mesh%vi1:mesh%vi2), target, intent(in   ), optional :: BC_prescr_mask
    real(dp), dimension(mesh%vi1:mesh%vi2), target, intent(in   ), optional :: BC_prescr_Hi
    character(len=1024), parameter :: routine_name = 'calc_dHi_dt_B200'
    type(ufe_b200_config) :: cfg
    type(c_ptr), target :: io(17)
    integer(c_int32_t), dimension(mesh%vi1:mesh%vi2), target :: noice
    type(c_ptr) :: bm, bh
    integer(c_int) :: its
    real(c_double) :: dt_c
    call init_routine( routine_name)
    bm = c_null_ptr; bh = c_null_ptr
    if (present( BC_prescr_mask) .and. present( BC_prescr_Hi)) then
      bm = c_loc( BC_prescr_mask); bh = c_loc( BC_prescr_Hi)
    end if
    call ufe_b200_fill_config( cfg)
    noice = merge( 1_c_int32_t, 0_c_int32_t, mask_noice)
    ! member order of ufe_b200_thickness_io (include/ufe_b200.h)
    io = [c_loc( Hi), c_loc( Hb), c_loc( SL), c_loc( u_vav_b), c_loc( v_vav_b), c_loc( SMB), c_loc( BMB), c_loc( LMB), c_loc( fraction_margin), &
          c_loc( noice), c_loc( dHi_dt_target), bm, bh, c_loc( AMB), c_loc( dHi_dt), c_loc( Hi_tplusdt), c_loc( divQ)]
    dt_c = dt
    call check( c_ufe_b200_calc_dHi_dt( handle, cfg, c_loc( io), dt_c, 1_c_int, its), routine_name)
    dt = dt_c
    call finalise_routine( routine_name)
  end subroutine calc_dHi_dt_B200

  subroutine set_ice_velocities_B200( mesh, ice)
    ! Drop-in for the set_ice_velocities_to_{SIA,SSA,SIASSA,DIVA}_results call that follows the solve in solve_stress_balance
    ! (src/UFEMISM/ice_dynamics/conservation_of_momentum/conservation_of_momentum_main.f90:113-152): the solver results are on the device,
    ! the applied velocities and strain rates land in the ice model's own arrays.
    type(type_mesh),              intent(in   ) :: mesh
    type(type_ice_model), target, intent(inout) :: ice
    character(len=1024), parameter :: routine_name = 'set_ice_velocities_B200'
    type(c_ptr), target :: out(10)
    integer(c_int) :: approximation, scheme
    call init_routine( routine_name)
    select case (C%choice_stress_balance_approximation)
    case ('SSA');     approximation = 0
    case ('DIVA');    approximation = 1
    case ('SIA');     approximation = 2
    case ('SIA/SSA'); approximation = 3
    case default
      call crash('unknown choice_stress_balance_approximation "' // trim( C%choice_stress_balance_approximation) // '"!')
    end select
    scheme = 0
    if (approximation == 3) then
      select case (C%choice_hybrid_SIASSA_scheme)
      case ('add');             scheme = 0
      case ('add_SIA_reduced'); scheme = 1
      case default
        call crash('unknown choice_hybrid_SIASSA_scheme_config "' // trim( C%choice_hybrid_SIASSA_scheme) // '"!')
      end select
    end if
    ! member order of ufe_b200_ice_velocities (include/ufe_b200.h)
    out = [c_loc( ice%u_3D_b), c_loc( ice%v_3D_b), c_loc( ice%du_dx_3D), c_loc( ice%du_dy_3D), c_loc( ice%du_dz_3D), &
           c_loc( ice%dv_dx_3D), c_loc( ice%dv_dy_3D), c_loc( ice%dv_dz_3D), c_loc( ice%dw_dx_3D), c_loc( ice%dw_dy_3D)]
    call check( c_ufe_b200_set_ice_velocities( handle, approximation, scheme, c_loc( out)), routine_name)
    call finalise_routine( routine_name)
  end subroutine set_ice_velocities_B200

  subroutine calc_pc_truncation_error_B200( mesh, ice, pc, n_tot, n_guilty)
    ! Drop-in for calc_pc_truncation_error( mesh, ice, pc) plus the loop that counts the unstable vertices in run_ice_dynamics_model_pc
    ! (src/UFEMISM/ice_dynamics/time_stepping/predictor_corrector_scheme.f90:327-361, 254-279); n_tot, n_guilty are already summed over the ranks.
    type(type_mesh),              intent(in   ) :: mesh
    type(type_ice_model), target, intent(in   ) :: ice
    type(type_ice_pc),    target, intent(inout) :: pc
    integer,                      intent(  out) :: n_tot, n_guilty
    character(len=1024), parameter :: routine_name = 'calc_pc_truncation_error_B200'
    integer(c_int32_t), dimension(mesh%vi1:mesh%vi2), target :: mask_grounded_ice, mask_gl_gr
    integer(c_int32_t) :: nt, ng
    real(c_double) :: eta
    call init_routine( routine_name)
    mask_grounded_ice = merge( 1_c_int32_t, 0_c_int32_t, ice%mask_grounded_ice( mesh%vi1:mesh%vi2))
    mask_gl_gr        = merge( 1_c_int32_t, 0_c_int32_t, ice%mask_gl_gr(        mesh%vi1:mesh%vi2))
    call check( c_ufe_b200_calc_pc_truncation_error( handle, pc%zeta_t, pc%dt_n, C%pc_eta_min, C%pc_epsilon, c_loc( pc%Hi_np1), &
      c_loc( pc%Hi_star_np1), c_loc( mask_grounded_ice), c_loc( mask_gl_gr), c_loc( ice%fraction_gr), c_loc( pc%tau_np1), &
      c_loc( pc%tau_n_guilty), eta, nt, ng), routine_name)
    pc%eta_np1 = eta
    n_tot = int( nt); n_guilty = int( ng)
    call finalise_routine( routine_name)
  end subroutine calc_pc_truncation_error_B200

  subroutine ufe_b200_fill_config( cfg)
    ! Copy the C% scalars the path reads into the plain struct that crosses the ABI
    type(ufe_b200_config), intent(out) :: cfg
    cfg%visc_it_norm_dUV_tol = C%visc_it_norm_dUV_tol; cfg%visc_it_relax = C%visc_it_relax; cfg%visc_eff_min = C%visc_eff_min
    cfg%vel_max = C%vel_max; cfg%stress_balance_PETSc_rtol = C%stress_balance_PETSc_rtol
    cfg%stress_balance_PETSc_abstol = C%stress_balance_PETSc_abstol
    cfg%Glens_flow_law_exponent = C%Glens_flow_law_exponent; cfg%Glens_flow_law_epsilon_sq_0 = C%Glens_flow_law_epsilon_sq_0
    cfg%slid_delta_v = C%slid_delta_v; cfg%slid_Weertman_m = C%slid_Weertman_m; cfg%slid_Budd_q_plastic = C%slid_Budd_q_plastic
    cfg%slid_Budd_u_threshold = C%slid_Budd_u_threshold; cfg%slid_ZI_ut = C%slid_ZI_ut; cfg%slid_ZI_p = C%slid_ZI_p
    cfg%slid_beta_max = C%slid_beta_max; cfg%subgrid_friction_exponent_on_B_grid = C%subgrid_friction_exponent_on_B_grid
    cfg%Hi_min = C%Hi_min; cfg%uniform_Glens_flow_factor = C%uniform_Glens_flow_factor
    cfg%refgeo_idealised_SSA_icestream_Hi = C%refgeo_idealised_SSA_icestream_Hi
    cfg%refgeo_idealised_SSA_icestream_dhdx = C%refgeo_idealised_SSA_icestream_dhdx
    cfg%refgeo_idealised_SSA_icestream_L = C%refgeo_idealised_SSA_icestream_L
    cfg%refgeo_idealised_SSA_icestream_m = C%refgeo_idealised_SSA_icestream_m
    cfg%refgeo_idealised_ISMIP_HOM_L = C%refgeo_idealised_ISMIP_HOM_L
    cfg%dHi_semiimplicit_fs = C%dHi_semiimplicit_fs; cfg%dHi_PETSc_rtol = C%dHi_PETSc_rtol; cfg%dHi_PETSc_abstol = C%dHi_PETSc_abstol
    cfg%dt_ice_min = C%dt_ice_min; cfg%dt_ice_max = C%dt_ice_max
    cfg%ice_density = 917.0_dp; cfg%seawater_density = 1027.0_dp; cfg%grav = 9.81_dp   ! parameters.f90:19-24
    cfg%visc_it_nit = C%visc_it_nit
    cfg%do_include_SSADIVA_crossterms = merge( 1, 0, C%do_include_SSADIVA_crossterms)
    cfg%BC_u = [bc_code( C%BC_u_north), bc_code( C%BC_u_east), bc_code( C%BC_u_south), bc_code( C%BC_u_west)]
    cfg%BC_v = [bc_code( C%BC_v_north), bc_code( C%BC_v_east), bc_code( C%BC_v_south), bc_code( C%BC_v_west)]
    cfg%BC_H = [bc_code( C%BC_H_north), bc_code( C%BC_H_east), bc_code( C%BC_H_south), bc_code( C%BC_H_west)]
    cfg%choice_sliding_law = sliding_code( C%choice_sliding_law)
    cfg%do_GL_subgrid_friction = merge( 1, 0, C%do_GL_subgrid_friction)
    cfg%choice_ice_integration_method = merge( 0, 1, C%choice_ice_integration_method == 'explicit')
    cfg%ksp_type = C%B200_KSP_type; cfg%pc_type = C%B200_PC_type; cfg%ksp_maxit = 2000
    cfg%write_diagnostics = merge( 1, 0, C%do_write_checksum_log)
    cfg%ksp_initial_guess_nonzero = merge( 1, 0, C%B200_use_initial_guess)
    cfg%stiffness_format = C%B200_stiffness_format
    cfg%reserved = 0
  end subroutine ufe_b200_fill_config

  pure function bc_code( name) result( code)
    character(len=*), intent(in) :: name
    integer(c_int32_t) :: code
    select case (trim( name))
    case ('infinite'); code = 0
    case ('zero');     code = 1
    case ('periodic_ISMIP-HOM');     code = 2   ! BC_u / BC_v only (copy rows, solve_linearised_SSA_DIVA_infinite_slab.f90:547-563)
    case ('infinite_SSA_icestream'); code = 3   ! BC_u only in the reference (:565-581)
    case default;      code = -1   ! rejected by the library with the reference's own 'unknown choice_BC' message
    end select
  end function bc_code

  pure function sliding_code( name) result( code)
    character(len=*), intent(in) :: name
    integer(c_int32_t) :: code
    select case (trim( name))
    case ('no_sliding');   code = 0
    case ('Weertman');     code = 1
    case ('Coulomb');      code = 2
    case ('Budd');         code = 3
    case ('Tsai2015');     code = 4
    case ('Schoof2005');   code = 5
    case ('Zoet-Iverson'); code = 6
    case ('idealised');    code = 7   ! SSA_icestream only
    case default;          code = -1
    end select
  end function sliding_code

end module ufe_b200_mod
