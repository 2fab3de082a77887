/* ufe_b200.h -- C ABI of the B200-native UFEMISM hot path (libufe_b200.so).
 *
 * Drop-in boundary behind the reference's Fortran module interfaces.  Every entry
 * point names the reference interface it replaces (paths relative to the reference repository).
 * The Fortran side binds these with ISO_C_BINDING (ufemism2.0_b200/fortran/ufe_b200_mod.f90) and
 * is selected by the namelist key  choice_linear_algebra_backend_config = 'B200'.
 *
 * Conventions (exactly what a Fortran caller owns):
 *   - arrays are column-major; indices stored in arrays are 1-based; 0 = "none"
 *   - a field on the a/b/c-grid is passed as the rank's owned slice (vi1:vi2), (ti1:ti2), (ei1:ei2);
 *     3-D fields are (n_loc, nz) column-major, i.e. layer-major in memory
 *   - mesh arrays are the full (global) arrays, which every MPI rank of the reference holds
 *   - all calls are collective over the ranks of the handle and synchronous on return
 *   - every function returns 0 on success; ufe_b200_last_error() describes the last failure.
 *     Krylov non-convergence is NOT an error (iteration count == maxit), exactly like the
 *     reference, which never checks PETSc's return codes (src/UPSY/basic/petsc_basic.f90:199-205).
 *   - there is no CPU fallback: without a CUDA device every compute entry point fails.
 */
#ifndef UFE_B200_H
#define UFE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ufe_b200_ctx* ufe_b200_handle;

/* ---- enumerations ---------------------------------------------------------------------------- */
enum { UFE_KSP_BICGSTAB = 0 };   /* the only Krylov method: neither matrix of the path is symmetric, so north_star (b)'s CG has no use here */
enum { UFE_PC_JACOBI = 0, UFE_PC_NONE = 1,   /* values 2 and 3 are unassigned */
       UFE_PC_MG = 4 };   /* aggregation multigrid, stiffness-matrix solve only (new key B200_PC_type_config = 'multigrid') */
enum { UFE_BC_INFINITE = 0, UFE_BC_ZERO = 1,                          /* C%BC_{u,v,H}_{north,east,south,west} */
       UFE_BC_PERIODIC_ISMIP_HOM = 2, UFE_BC_INFINITE_SSA_ICESTREAM = 3 }; /* BC_u / BC_v only: the copy rows of
       calc_SSA_DIVA_stiffness_matrix_row_BC (solve_linearised_SSA_DIVA_infinite_slab.f90:547-579, 618-636) */
enum { UFE_SLID_NO_SLIDING = 0, UFE_SLID_WEERTMAN = 1, UFE_SLID_COULOMB = 2, UFE_SLID_BUDD = 3, UFE_SLID_TSAI2015 = 4,
       UFE_SLID_SCHOOF2005 = 5, UFE_SLID_ZOET_IVERSON = 6, UFE_SLID_IDEALISED_SSA_ICESTREAM = 7 };
enum { UFE_INTEGRATION_EXPLICIT = 0, UFE_INTEGRATION_SEMI_IMPLICIT = 1 };

/* mesh operators, in the order calc_all_matrix_operators_mesh builds them
 * (src/UPSY/mesh/discretisation/mesh_disc_calc_matrix_operators_2D.f90:26-59; mesh_types.f90:213-232) */
enum { UFE_OP_DDX_A_A = 0, UFE_OP_DDY_A_A, UFE_OP_MAP_A_B, UFE_OP_DDX_A_B, UFE_OP_DDY_A_B, UFE_OP_MAP_B_A, UFE_OP_DDX_B_A,
       UFE_OP_DDY_B_A, UFE_OP_DDX_B_B, UFE_OP_DDY_B_B, UFE_OP2_DDX_B_B, UFE_OP2_DDY_B_B, UFE_OP2_D2DX2_B_B,
       UFE_OP2_D2DXDY_B_B, UFE_OP2_D2DY2_B_B, UFE_OP_COUNT };

/* ---- the C% scalars the path reads (src/UPSY/basic/model_configuration/model_configuration_type_and_namelist.f90) */
typedef struct ufe_b200_config {
  double visc_it_norm_dUV_tol;              /* :306 */
  double visc_it_relax;                     /* :308 */
  double visc_eff_min;                      /* :309 */
  double vel_max;                           /* :310 */
  double stress_balance_PETSc_rtol;         /* :311 */
  double stress_balance_PETSc_abstol;       /* :312 */
  double Glens_flow_law_exponent;           /* :598 */
  double Glens_flow_law_epsilon_sq_0;       /* :599 */
  double slid_delta_v;                      /* :349 */
  double slid_Weertman_m;                   /* :333 */
  double slid_Budd_q_plastic;               /* :334 */
  double slid_Budd_u_threshold;             /* :335 */
  double slid_ZI_ut;                        /* :337 */
  double slid_ZI_p;                         /* :336 */
  double slid_beta_max;                     /* :348 */
  double subgrid_friction_exponent_on_B_grid; /* :345 */
  double Hi_min;                            /* :437 */
  double uniform_Glens_flow_factor;         /* :603 */
  double refgeo_idealised_SSA_icestream_Hi, refgeo_idealised_SSA_icestream_dhdx,
         refgeo_idealised_SSA_icestream_L, refgeo_idealised_SSA_icestream_m;
  double refgeo_idealised_ISMIP_HOM_L;      /* period of BC 'periodic_ISMIP-HOM' (mesh_utilities.f90:2620-2675) */
  double dHi_semiimplicit_fs;               /* :356 */
  double dHi_PETSc_rtol, dHi_PETSc_abstol;  /* :357-358 */
  double dt_ice_min, dt_ice_max;            /* :391-392 */
  double ice_density, seawater_density, grav; /* src/UPSY/basic/parameters.f90:19-24 */
  int32_t visc_it_nit;                      /* :307 */
  int32_t do_include_SSADIVA_crossterms;    /* :275 */
  int32_t BC_u[4], BC_v[4];                 /* north, east, south, west  (:319-326) */
  int32_t choice_sliding_law;               /* UFE_SLID_* */
  int32_t do_GL_subgrid_friction;           /* :340 */
  int32_t choice_ice_integration_method;    /* UFE_INTEGRATION_* (:355) */
  int32_t BC_H[4];                          /* :364-367 */
  int32_t ksp_type, pc_type, ksp_maxit;     /* new keys B200_KSP_type_config / B200_PC_type_config; maxit = 2000 (petsc_basic.f90:166) */
  int32_t write_diagnostics;                /* 1: also fill the intermediate DIVA fields that only checksum() reads (eta_vav_a, F1/F2_3D_b of all layers every iteration) */
  int32_t ksp_initial_guess_nonzero;        /* 0 (default): zero initial guess, the reference's behaviour (petsc_basic.f90:163 is commented out);
                                               1: start the Krylov solve from the current solution (= enabling KSPSetInitialGuessNonzero) */
  int32_t stiffness_format;                 /* storage the Krylov solver streams: 0 = 2x2-blocked sliced ELL (default), 1 = the CSR matrix itself */
  int32_t reserved[4];
} ufe_b200_config;

/* Fill cfg with the reference's namelist defaults. */
void ufe_b200_config_defaults(ufe_b200_config* cfg);

/* ---- life cycle -------------------------------------------------------------------------------- */
/* Replaces initialise_parallelisation + PetscInitialize for this path (src/UFEMISM/main/UFEMISM_program.f90:55-56).
 * One rank <-> one GPU. For nranks > 1 rank 0 calls ufe_b200_get_unique_id() and broadcasts the 128 bytes with
 * the reference's own MPI_BCAST before every rank calls ufe_b200_init().  device < 0 picks rank % device_count. */
int ufe_b200_get_unique_id(void* id128);
int ufe_b200_init(int device, int rank, int nranks, const void* id128, ufe_b200_handle* out);
int ufe_b200_finalize(ufe_b200_handle h);
const char* ufe_b200_last_error(void);
/* number of CUDA kernels this library has launched since init (bench.py's gpu_launches claim) */
long long ufe_b200_kernel_launches(ufe_b200_handle h);

/* ---- narrow seam: linear algebra --------------------------------------------------------------- */
/* Replaces solve_matrix_equation_CSR_PETSc( AA, bb, xx, rtol, abstol, n_Axb_its, ...)  (src/UPSY/basic/petsc_basic.f90:32-215).
 * AA is the rank's finalised type_sparse_matrix_CSR_dp: rows i1..i1+m_loc-1 of an m_glob x m_glob matrix;
 * ptr has m_loc+1 entries holding 1-based offsets into ind/val relative to ptr[0]; ind = global 1-based columns.
 * xx on input is ignored (zero initial guess, petsc_basic.f90:163). */
int ufe_b200_ksp_solve(ufe_b200_handle h, int m_glob, int m_loc, int i1, const int32_t* ptr, const int32_t* ind,
                       const double* val, const double* bb, double* xx, double rtol, double abstol, int maxit,
                       int ksp_type, int pc_type, int* n_Axb_its);

/* Replaces multiply_CSR_matrix_with_vector_{1D,2D}_wrapper (src/UPSY/basic/CSR_matrix_algebra/CSR_matrix_vector_multiplication.f90:24,112):
 * register a matrix once (pattern + values stay on the device), then y = A x for ncols right-hand sides
 * laid out (n_loc, ncols) / (m_loc, ncols) column-major.  j1/n_loc give the rank's owned slice of x. */
int ufe_b200_csr_register(ufe_b200_handle h, int id, int m_glob, int n_glob, int m_loc, int i1, int n_loc, int j1,
                          const int32_t* ptr, const int32_t* ind, const double* val);
int ufe_b200_csr_update_values(ufe_b200_handle h, int id, const double* val);
int ufe_b200_spmv(ufe_b200_handle h, int id, const double* xx, double* yy, int ncols);
int ufe_b200_csr_release(ufe_b200_handle h, int id);

/* ---- wide seam: mesh --------------------------------------------------------------------------- */
/* The members of type_mesh the path reads (src/UPSY/types/mesh_types.f90:40-131). */
typedef struct ufe_b200_mesh_desc {
  int32_t nV, nTri, nE, nC_mem, nz;
  double xmin, xmax, ymin, ymax;
  const double* V;       /* (nV,2)       */
  const int32_t* nC;     /* (nV)         */
  const int32_t* C;      /* (nV,nC_mem)  */
  const int32_t* niTri;  /* (nV)         */
  const int32_t* iTri;   /* (nV,nC_mem)  */
  const int32_t* VBI;    /* (nV)         */
  const int32_t* Tri;    /* (nTri,3)     */
  const int32_t* TriC;   /* (nTri,3)     */
  const int32_t* TriBI;  /* (nTri)       */
  const double* TriGC;   /* (nTri,2)     */
  const double* A;       /* (nV)         */
  const double* Cw;      /* (nV,nC_mem)  */
  const double* D_x;     /* (nV,nC_mem)  */
  const double* D_y;     /* (nV,nC_mem)  */
  const double* D;       /* (nV,nC_mem)  */
  const int32_t* VE;     /* (nV,nC_mem)  */
  const int32_t* ETri;   /* (nE,2)       */
  const double* zeta;    /* (nz)         */
  const double* Tricc;   /* (nTri,2) circumcentres = Voronoi vertices (mesh%Tricc); may be NULL unless ufe_b200_calc_grounded_fractions is used */
} ufe_b200_mesh_desc;

/* Upload the mesh (lifetime = lifetime of type_mesh; destroy on remesh, src/UFEMISM/main/UFEMISM_main_model.f90:1211).
 * Ownership ranges follow partition_list (src/UPSY/basic/mpi_parallelisation/mpi_distributed_memory.f90:47-73). */
int ufe_b200_mesh_create(ufe_b200_handle h, const ufe_b200_mesh_desc* mesh);
int ufe_b200_mesh_destroy(ufe_b200_handle h);
/* out = {vi1,vi2,ti1,ti2,ei1,ei2} (1-based, inclusive) */
int ufe_b200_mesh_ranges(ufe_b200_handle h, int32_t* out6);

/* Replaces calc_all_matrix_operators_mesh (mesh_disc_calc_matrix_operators_2D.f90:26-59): builds the 15 operator
 * matrices on the device (patterns bit-identical to the reference's construction order). */
int ufe_b200_calc_matrix_operators(ufe_b200_handle h);
/* Alternatively hand over an operator the Fortran side already built (rank-local rows, as in ufe_b200_ksp_solve). */
int ufe_b200_set_operator(ufe_b200_handle h, int op, int m_loc, const int32_t* ptr, const int32_t* ind, const double* val);
/* Inspect an operator (parity checks): sizes first (nnz of the local rows), then the arrays. */
int ufe_b200_get_operator_size(ufe_b200_handle h, int op, int* m_loc, int* nnz_loc);
int ufe_b200_get_operator(ufe_b200_handle h, int op, int32_t* ptr, int32_t* ind, double* val);

/* Replaces the 16 wrappers ddx/ddy/map_{a_a,a_b,b_a}_{2D,3D} (+ b_b) (src/UPSY/mesh/discretisation/mesh_disc_apply_operators.f90:18-431).
 * d_in / d_out are the rank's owned slices, (n_loc, ncols) column-major; ncols = 1 (2D) or nz (3D). */
int ufe_b200_apply_operator(ufe_b200_handle h, int op, const double* d_in, double* d_out, int ncols);

/* ---- wide seam: stress balance ------------------------------------------------------------------ */
/* Replaces solve_SSA_DIVA_linearised( mesh, u_b, v_b, N_b, dN_dx_b, dN_dy_b, basal_friction_coefficient_b, tau_dx_b, tau_dy_b,
 *   u_b_prev, v_b_prev, PETSc_rtol, PETSc_abstol, n_Axb_its, BC_prescr_mask_b, BC_prescr_u_b, BC_prescr_v_b)
 * (src/UFEMISM/ice_dynamics/conservation_of_momentum/SSA_DIVA/solve_linearised_SSA_DIVA_infinite_slab.f90:23-178).
 * u_b_prev / v_b_prev are the rank's owned slices here (the reference gathers them globally only for the
 * periodic / ice-stream copy BCs; the library completes the previous solution across ranks itself when cfg selects
 * UFE_BC_PERIODIC_ISMIP_HOM / UFE_BC_INFINITE_SSA_ICESTREAM). BC_* may be NULL (no prescribed velocities). */
int ufe_b200_solve_SSA_DIVA_linearised(ufe_b200_handle h, const ufe_b200_config* cfg, double* u_b, double* v_b,
                                       const double* N_b, const double* dN_dx_b, const double* dN_dy_b,
                                       const double* basal_friction_coefficient_b, const double* tau_dx_b,
                                       const double* tau_dy_b, double* u_b_prev, double* v_b_prev, double rtol,
                                       double abstol, int* n_Axb_its, const int32_t* BC_prescr_mask_b,
                                       const double* BC_prescr_u_b, const double* BC_prescr_v_b);
/* The system assembled by the last linearised solve (rank-local rows): sizes, then arrays (bit-exact ptr/ind parity). */
int ufe_b200_get_stiffness_size(ufe_b200_handle h, int* m_loc, int* nnz_loc);
int ufe_b200_get_stiffness(ufe_b200_handle h, int32_t* ptr, int32_t* ind, double* val, double* bb);
/* bytes of matrix data one SpMV streams: the CSR arrays (12 nnz + 4 (rows+1)) and the 2x2-blocked sliced-ELL copy (36 per block slot) */
int ufe_b200_get_stiffness_storage_bytes(ufe_b200_handle h, long long* csr_bytes, long long* blocked_bytes);
/* Knobs of the multigrid preconditioner (pc_type = UFE_PC_MG) that replaces PETSc's one-level PC behind
 * solve_matrix_equation_CSR_PETSc (src/UPSY/basic/petsc_basic.f90:125-129, 193) for the SSA/DIVA system: block-Jacobi sweeps before / after the
 * coarse correction, damping, correction scaling, number of coarse levels visited twice per cycle, FP32 storage of the preconditioner's
 * matrix values (0/1), largest coarsest level solved by a dense inverse.  Negative / NaN arguments keep the current value. */
int ufe_b200_set_mg_params(ufe_b200_handle h, int nu_pre, int nu_post, double omega, double overcorr, int w_levels, int fp32, int dense_max);
/* Further knobs of the same preconditioner by name (new key B200_MG_<name>_config): "adaptive" (1: the level-0 correction is scaled by the
 * residual-minimising factor instead of `overcorr`), "adaptive_max", "nu_pre" (0 = no pre-smoothing), "nu_post", "coarse_sweeps", "tail_max" /
 * "tail_cluster" (the smallest levels as one kernel per visit, on a thread-block cluster), "fuse", "use_graph".  Unknown names are an error. */
int ufe_b200_set_mg_option(ufe_b200_handle h, const char* name, double value);
/* Hierarchy of the last multigrid build: number of levels, nodes per level (up to 16 entries), matrix bytes one fine-level smoothing sweep
 * streams, total matrix bytes of all coarse levels, host seconds of the one-off pattern phase. */
int ufe_b200_get_mg_info(ufe_b200_handle h, int* n_levels, int* nodes_per_level, long long* fine_bytes, long long* coarse_bytes, double* setup_s);
/* Host logic of the multigrid pattern phase, exposed for the CPU test-suite (no GPU involved, no handle): one coarsening step of a 2x2-block
 * pattern given in the solver's sliced-ELL layout (slices of 32 nodes; slice_ptr[(n+31)/32 + 1] in k-steps; col[k*32 + lane] = extended
 * index of the neighbour node, own node = off + t).  The first call computes and reports sizes, the second copies the arrays out:
 * agg[n], coarse_slice_ptr[(nc+31)/32 + 1], coarse_col[coarse_slots], gal_ptr[coarse_slots + 1], gal_src[gather_entries]. */
int ufe_b200_mg_host_transfer(int n, int off, const int32_t* slice_ptr, const int32_t* col, int* nc, long long* coarse_slots, long long* gather_entries);
int ufe_b200_mg_host_transfer_get(int32_t* agg, int32_t* coarse_slice_ptr, int32_t* coarse_col, int32_t* gal_ptr, int32_t* gal_src);
/* The same for rows split into nparts row blocks (= ranks), emulated in one process with the exchanges replaced by copies: part p owns nodes
 * off[p] .. off[p]+n[p]-1 of one global numbering of n_total nodes (col holds global ids), its pattern starts at slice_ptr[sp_start[p]] /
 * col[col_start[p]].  Outputs: agg_glob[n_total] (global coarse ids), the assembled coarse pattern, and per part a gather list over the
 * assembled coarse slots: gal_ptr[nparts * (coarse_slots + 1)], gal_src (part p at gal_src_start[p] .. gal_src_start[p+1]). */
int ufe_b200_mg_host_transfer_parts(int nparts, int n_total, const int32_t* n, const int32_t* off, const int32_t* sp_start, const int32_t* col_start,
                                    const int32_t* slice_ptr, const int32_t* col, int* nc, long long* coarse_slots, long long* gather_entries);
int ufe_b200_mg_host_transfer_parts_get(int32_t* agg_glob, int32_t* coarse_slice_ptr, int32_t* coarse_col, int32_t* gal_ptr, int32_t* gal_src,
                                        int32_t* gal_src_start);

/* Inputs of the velocity solve taken from type_ice_model / type_bed_roughness_model (owned slices). */
typedef struct ufe_b200_ice_inputs {
  const double* Hi;                  /* (vi1:vi2) ice%Hi */
  const double* Hs;                  /* (vi1:vi2) ice%Hs */
  const double* A_flow;              /* (vi1:vi2,nz) ice%A_flow after calc_ice_rheology_Glen (rheology/constitutive_equation.f90:84-163) */
  const double* fraction_gr_b;       /* (ti1:ti2) ice%fraction_gr_b */
  const double* till_friction_angle; /* (vi1:vi2) bed_roughness%till_friction_angle (may be NULL if the law does not use it) */
  const double* beta_sq;             /* (vi1:vi2) bed_roughness%beta_sq   (may be NULL) */
  const double* alpha_sq;            /* (vi1:vi2) bed_roughness%alpha_sq  (may be NULL) */
  const double* effective_pressure;  /* (vi1:vi2) ice%effective_pressure  (may be NULL) */
  const int32_t* mask_grounded_ice;  /* (vi1:vi2) 0/1 */
  const int32_t* mask_icefree_land;  /* (vi1:vi2) 0/1 */
} ufe_b200_ice_inputs;

/* type_ice_velocity_solver_DIVA (src/UFEMISM/types/ice_model_types.f90:149-197): inout state + outputs.
 * NULL members are skipped on download.  For solve_SSA, u_vav_b/v_vav_b carry SSA%u_b/SSA%v_b. */
typedef struct ufe_b200_velocity_state {
  double *u_vav_b, *v_vav_b;                 /* inout (ti1:ti2) */
  double *u_base_b, *v_base_b;               /* inout */
  double *tau_bx_b, *tau_by_b;               /* inout */
  double *eta_3D_b;                          /* inout (ti1:ti2,nz) */
  double *u_b_prev, *v_b_prev;               /* out   */
  double *tau_dx_b, *tau_dy_b;               /* out   */
  double *u_3D_b, *v_3D_b;                   /* out (ti1:ti2,nz) */
  double *du_dx_a, *du_dy_a, *dv_dx_a, *dv_dy_a;      /* out (vi1:vi2) */
  double *du_dz_3D_a, *dv_dz_3D_a, *eta_3D_a;         /* out (vi1:vi2,nz) */
  double *eta_vav_a, *N_a;                   /* out */
  double *N_b, *dN_dx_b, *dN_dy_b;           /* out */
  double *F1_3D_a, *F2_3D_a;                 /* out (vi1:vi2,nz) */
  double *F1_3D_b, *F2_3D_b;                 /* out (ti1:ti2,nz) */
  double *basal_friction_coefficient_a;      /* out ice%basal_friction_coefficient */
  double *basal_friction_coefficient_b, *beta_eff_a, *beta_eff_b; /* out */
} ufe_b200_velocity_state;

/* Replaces solve_DIVA_infinite_slab( mesh, ice, bed_roughness, DIVA, n_visc_its, n_Axb_its, BC_prescr_*)
 * (SSA_DIVA/DIVA_solver_infinite_slab.f90:52-231) including calc_driving_stress and calc_3D_velocities. */
int ufe_b200_solve_DIVA(ufe_b200_handle h, const ufe_b200_config* cfg, const ufe_b200_ice_inputs* ice,
                        ufe_b200_velocity_state* DIVA, int* n_visc_its, int* n_Axb_its,
                        const int32_t* BC_prescr_mask_b, const double* BC_prescr_u_b, const double* BC_prescr_v_b);
/* Replaces solve_SSA (SSA_DIVA/SSA_main.f90:88-245). */
int ufe_b200_solve_SSA(ufe_b200_handle h, const ufe_b200_config* cfg, const ufe_b200_ice_inputs* ice,
                       ufe_b200_velocity_state* SSA, int* n_visc_its, int* n_Axb_its,
                       const int32_t* BC_prescr_mask_b, const double* BC_prescr_u_b, const double* BC_prescr_v_b);

/* Device-resident variants (inputs stay in HBM between calls; used for the time loop and by bench.py's
 * device-timed leg): upload once, solve any number of times, download when the host needs the fields. */
int ufe_b200_velocity_upload(ufe_b200_handle h, const ufe_b200_ice_inputs* ice, const ufe_b200_velocity_state* st);
int ufe_b200_velocity_solve_resident(ufe_b200_handle h, const ufe_b200_config* cfg, int approximation /*0 SSA, 1 DIVA*/,
                                     int* n_visc_its, int* n_Axb_its);
int ufe_b200_velocity_download(ufe_b200_handle h, ufe_b200_velocity_state* st);

/* Replaces remap_DIVA_solver( mesh_old, mesh_new, DIVA) / remap_SSA_solver( mesh_old, mesh_new, SSA)
 * (src/UFEMISM/ice_dynamics/conservation_of_momentum/SSA_DIVA/DIVA_main.f90:115-222, SSA_main.f90:247-311; called from remap_velocity_solver after
 * update_mesh, UFEMISM_main_model.f90:1211) for a solver state that lives on the device: the fields the viscosity iteration re-uses (DIVA: u_vav_b,
 * v_vav_b, tau_bx_b, tau_by_b, eta_3D_b, u_3D_b, v_3D_b; SSA: u_b, v_b) go  map_b_a( mesh_old) -> Atlas map -> set_border_vertices_to_interior_mean
 * (apply_maps.f90:531-583, mesh_utilities.f90:1648-1757) -> map_a_b( mesh_new)  from the state of h_old into the state of h_new; everything else of
 * the solver object starts from zero (reallocate_bounds / reallocate_clean).  h_old: the handle of the old mesh with its velocity state; h_new: a second
 * handle of the same rank / GPU on which ufe_b200_mesh_create + ufe_b200_calc_matrix_operators have run for the new mesh.  The map is the rank's rows
 * (new vertices vi1:vi2) of type_map%M as CSR (1-based offsets relative to map_ptr[0], 1-based global columns = vertices of the old mesh); building it
 * (create_map_from_mesh_to_mesh_*, line tracing through both meshes) is mesh generation and stays on the host.  Afterwards: ufe_b200_velocity_upload(
 * h_new, ice, NULL) for the inputs on the new mesh, ufe_b200_finalize( h_old) -- the state is NOT copied back to the host in between. */
int ufe_b200_remap_velocity_solver(ufe_b200_handle h_old, ufe_b200_handle h_new, int approximation /*0 SSA, 1 DIVA*/, int m_loc,
                                   const int32_t* map_ptr, const int32_t* map_ind, const double* map_val);

/* ---- wide seam: ice thickness -------------------------------------------------------------------- */
/* Replaces calc_dHi_dt( mesh, ice, Hi, Hb, SL, u_vav_b, v_vav_b, SMB, BMB, LMB, AMB, fraction_margin, mask_noice, dt,
 *   dHi_dt, Hi_tplusdt, divQ, dHi_dt_target, BC_prescr_mask, BC_prescr_Hi)
 * (src/UFEMISM/ice_dynamics/conservation_of_mass/conservation_of_mass_main.f90:27-119), explicit and semi-implicit.
 * calc_and_apply_spill_over_flux (utilities/ice_thickness_safeties.f90:290) is NOT included: callers with
 * calving-front cells set finish = 0, run the spill-over on the host and then call ufe_b200_dHi_dt_finish(). */
typedef struct ufe_b200_thickness_io {
  const double *Hi, *Hb, *SL;               /* (vi1:vi2) */
  const double *u_vav_b, *v_vav_b;          /* (ti1:ti2); NULL = use the device-resident velocity state */
  const double *SMB, *BMB, *LMB;            /* (vi1:vi2) */
  const double *fraction_margin;            /* (vi1:vi2) */
  const int32_t* mask_noice;                /* (vi1:vi2) 0/1 */
  const double* dHi_dt_target;              /* (vi1:vi2) */
  const int32_t* BC_prescr_mask;            /* (vi1:vi2) or NULL */
  const double* BC_prescr_Hi;               /* (vi1:vi2) or NULL */
  double *AMB, *dHi_dt, *Hi_tplusdt, *divQ; /* (vi1:vi2) out */
} ufe_b200_thickness_io;
int ufe_b200_calc_dHi_dt(ufe_b200_handle h, const ufe_b200_config* cfg, ufe_b200_thickness_io* io, double* dt,
                         int finish, int* n_Axb_its);
int ufe_b200_dHi_dt_finish(ufe_b200_handle h, const double* Hi, double* Hi_tplusdt, double dt, double* AMB,
                           double* dHi_dt);

/* Device-resident time stepping: upload the thickness inputs once, then every call advances one predictor-corrector step, i.e. one pass of
 * iterate_pc_timestep of run_ice_dynamics_model_pc (src/UFEMISM/ice_dynamics/time_stepping/predictor_corrector_scheme.f90:86-233; Robinson et al.
 * 2020, Eqs. 30-31) without any field leaving HBM, in the reference's order:
 *   Hi_prev, dHi_dt_nm1 := 0 under mask_noice (:88-93); effective thickness of region%ice%Hi (:107, with ufe_b200_geometry_resident);
 *   calc_dHi_dt( region%ice%Hi, u_n) (:129); Eq. 30 on Hi_prev (:140); alter_ice_thickness (:145); dHi_dt_Hi_n_u_n re-derived from the altered
 *   Hi_star (:152); Hs, masks and grounded fractions of Hi_star (:156-176); solve_stress_balance (:186); masks / effective thickness of Hi_prev
 *   (:198-211); calc_dHi_dt( Hi_prev, u_np1) (:214); Eq. 31 (:226); alter_ice_thickness (:233); Hi := Hi_np1.
 * Without ufe_b200_geometry_resident the masks and sub-grid fractions stay as uploaded and only Hs follows the thickness; without
 * ufe_b200_alter_resident alter_ice_thickness only removes ice whose effective thickness is below cfg->Hi_min (every optional modification off).
 * The truncation error of the finished step is ufe_b200_calc_pc_truncation_error (Hi_star_np1 from ufe_b200_pc_download); the scalar time-step
 * rule (Eq. 33, accept / retry) stays with the caller.  Needs ufe_b200_velocity_upload first. */
int ufe_b200_thickness_upload(ufe_b200_handle h, const ufe_b200_thickness_io* io);
int ufe_b200_pc_step_resident(ufe_b200_handle h, const ufe_b200_config* cfg, int approximation /*0 SSA, 1 DIVA*/, double dt,
                              int* n_visc_its, int* n_Axb_its_velocity, int* n_Axb_its_thickness);
int ufe_b200_thickness_download(ufe_b200_handle h, double* Hi, double* dHi_dt, double* divQ);
/* The same step through HOST buffers -- the widest seam: the device phases of one iterate_pc_timestep pass of run_ice_dynamics_model_pc
 * (src/UFEMISM/ice_dynamics/time_stepping/predictor_corrector_scheme.f90:118-233) in one call.
 *   in : io->Hi (= Hi_n), Hb, SL, SMB, BMB, LMB, fraction_margin, mask_noice, dHi_dt_target [, BC_prescr_*]; ice inputs except Hi/Hs (the
 *        velocity solve sees the predicted geometry); state (u_vav_b, v_vav_b, u_base_b, v_base_b, tau_bx_b, tau_by_b, eta_3D_b);
 *        dHi_dt_nm1 = pc%dHi_dt_Hi_nm1_u_nm1 = region%ice%dHi_dt (vi1:vi2); dt_prev = pc%dt_n (<= 0: first step).  io->Hi is Hi_prev; region%ice%Hi, where
 *        it differs, is tracked by the library or given by ufe_b200_pc_set_model_thickness (see ufe_b200_pc_reference_loop).
 *   out: io->Hi_tplusdt (= pc%Hi_np1 after alter_ice_thickness), io->dHi_dt / divQ / AMB of the corrector's calc_dHi_dt, dHi_dt_nm1 = the rate the next
 *        step's Eq. 30 takes (the applied rate (Hi_np1 - Hi_prev) / dt), every non-NULL state field.
 * Transfers the next phase does not depend on run on a second stream: the ice inputs and the 3-D state go up while the predictor runs, the
 * velocity state comes back while the corrector runs. */
int ufe_b200_pc_step(ufe_b200_handle h, const ufe_b200_config* cfg, int approximation /*0 SSA, 1 DIVA*/, double dt, double dt_prev,
                     const ufe_b200_ice_inputs* ice, ufe_b200_velocity_state* state, const ufe_b200_thickness_io* io, double* dHi_dt_nm1,
                     int* n_visc_its, int* n_Axb_its_velocity, int* n_Axb_its_thickness);
/* on = 1: the host does not modify the velocity solver's state between ufe_b200_pc_step calls (the reference keeps it inside the solver
 * object, type_ice_velocity_solver_DIVA, src/UFEMISM/types/ice_model_types.f90:149-197): after the first call the state fields are no
 * longer copied to the device, and `state` only selects which results are copied back (NULL members are skipped).  Call with 0 after the
 * host changed the state itself (e.g. remapping to a new mesh). */
int ufe_b200_set_state_resident(ufe_b200_handle h, int on);
/* on = 1 (the DEFAULT): the one-call time steps follow run_ice_dynamics_model_pc inside the reference's time loop in the two places where that loop
 * departs from Robinson et al. (2020), Eq. 30 as printed (both visible in the checksum log of integrated_test_Halfar_dome_small;
 * tests/test_reference_halfar_log.py):
 *   - the predictor's calc_dHi_dt (and the effective thickness before it) sees region%ice%Hi, which run_ice_dynamics_model left at the thickness
 *     interpolated at the last action time (src/UFEMISM/ice_dynamics/ice_dynamics_main.f90:115-121), while Eq. 30 / 31 add to Hi_prev
 *     (time_stepping/predictor_corrector_scheme.f90:129 against :140).  The library keeps that field on the device: after a step n -> n+1 it is
 *     (1 - w) Hi_n + w Hi_np1 with w = wt_next_of_last_action, the interpolation weight of the last action time before t_np1 (0 when no other sub-model
 *     acted in between: region%ice%Hi is then one whole step behind); before the first step it equals the uploaded Hi.  A caller that holds
 *     region%ice%Hi itself hands it over with ufe_b200_pc_set_model_thickness before the step instead;
 *   - pc%dHi_dt_Hi_nm1_u_nm1 is region%ice%dHi_dt, the applied rate (Hi_np1 - Hi_n) / dt of the previous step (ice_dynamics_main.f90:140 ->
 *     predictor_corrector_scheme.f90:109), not the previous predictor rate; ufe_b200_pc_step returns that rate in dHi_dt_nm1.
 * on = 0: Eq. 30 as printed -- one thickness, previous predictor rate.  Callable before every step to set w.  The state belongs to the handle and
 * is released with the mesh; ufe_b200_thickness_upload re-initialises the lagged thickness from the uploaded Hi. */
int ufe_b200_pc_reference_loop(ufe_b200_handle h, int on, double wt_next_of_last_action);
/* region%ice%Hi (vi1:vi2) as the host model holds it, for the predictor of the next step (reference-loop mode; see above) */
int ufe_b200_pc_set_model_thickness(ufe_b200_handle h, const double* Hi_model);
/* alter_ice_thickness inside the one-call steps (after Eq. 30 and after Eq. 31; predictor_corrector_scheme.f90:145, 233) with the caller's C% settings:
 * prm as for ufe_b200_alter_ice_thickness (below), time = region%time, refgeo_Hi / refgeo_Hb = refgeo_PD%Hi / %Hb and ice%Ti_hom (vi1:vi2; kept on the
 * device, may be NULL on later calls that only advance `time` or change prm).  The masks it reads are those of the model geometry at that point of the
 * loop (region%ice%Hi for the predictor, Hi_prev for the corrector), formed on the device.  prm = NULL switches back to the minimal form. */
struct ufe_b200_alter_params;
int ufe_b200_alter_resident(ufe_b200_handle h, const struct ufe_b200_alter_params* prm, double time, const double* refgeo_Hi, const double* refgeo_Hb,
                            const double* Ti_hom);
/* fields of the last one-call step that the host model keeps in type_ice_pc / type_ice_model (all (vi1:vi2), NULL = skip): pc%Hi_star_np1,
 * pc%dHi_dt_Hi_n_u_n (re-derived, :152), the rate the next Eq. 30 will use, and region%ice%Hi as the library tracks it */
int ufe_b200_pc_download(ufe_b200_handle h, double* Hi_star_np1, double* dHi_dt_Hi_n_u_n, double* dHi_dt_nm1_next, double* Hi_model);

/* ---- SIA (config 4) ------------------------------------------------------------------------------ */
/* Replaces solve_SIA( mesh, ice, SIA) (src/UFEMISM/ice_dynamics/conservation_of_momentum/SIA/SIA_main.f90:65-158), the analytical SIA
 * solution with Glen's flow law: in ice%Hi, ice%Hs (vi1:vi2), ice%A_flow (vi1:vi2,nz) after calc_ice_rheology_Glen, C%SIA_maximum_diffusivity
 * (model_configuration_type_and_namelist.f90:305, default 1e5); out SIA%u_3D_b, SIA%v_3D_b, SIA%D_3D_b (ti1:ti2,nz) and SIA%du_dz_3D, SIA%dv_dz_3D
 * (vi1:vi2,nz).  NULL outputs are skipped (u_3D_b and v_3D_b are required). */
int ufe_b200_solve_SIA(ufe_b200_handle h, const ufe_b200_config* cfg, double SIA_maximum_diffusivity, const double* Hi, const double* Hs,
                       const double* A_flow, double* u_3D_b, double* v_3D_b, double* D_3D_b, double* du_dz_3D_a, double* dv_dz_3D_a);

/* ---- applied velocities --------------------------------------------------------- */
/* Replaces set_ice_velocities_to_{SIA,SSA,SIASSA,DIVA}_results( mesh, ice, ...) (src/UFEMISM/ice_dynamics/conservation_of_momentum/
 * conservation_of_momentum_main.f90:311-511), the fan-out of the solver results into ice%u_3D_b, ice%v_3D_b and the 3-D strain rates that
 * solve_stress_balance does right after the solve (:113-152).  approximation = C%choice_stress_balance_approximation, hybrid_scheme =
 * C%choice_hybrid_SIASSA_scheme ('add' | 'add_SIA_reduced', only read for SIA/SSA).  The solver results are the ones on the device: for SIA/SSA
 * call ufe_b200_solve_SIA first (with du_dz_3D_a, dv_dz_3D_a if the strain rates are wanted), then ufe_b200_solve_SSA.  Afterwards the device's
 * u_3D_b, v_3D_b ARE ice%u_3D_b, ice%v_3D_b, so ufe_b200_calc_secondary_velocities( h, NULL, NULL, ...) and ufe_b200_calc_vertical_velocities
 * continue from them as calc_secondary_velocities / calc_vertical_velocities do in the reference (:167-170).  NULL outputs are skipped.  The strain
 * rates are those of the last viscosity iteration of the solve, whatever cfg->write_diagnostics was. */
enum { UFE_APPROX_SSA = 0, UFE_APPROX_DIVA = 1, UFE_APPROX_SIA = 2, UFE_APPROX_SIASSA = 3 };
enum { UFE_HYBRID_ADD = 0, UFE_HYBRID_ADD_SIA_REDUCED = 1 };
typedef struct ufe_b200_ice_velocities {
  double *u_3D_b, *v_3D_b;                                   /* (ti1:ti2, nz) */
  double *du_dx_3D, *du_dy_3D, *du_dz_3D, *dv_dx_3D, *dv_dy_3D, *dv_dz_3D, *dw_dx_3D, *dw_dy_3D;   /* (vi1:vi2, nz) */
} ufe_b200_ice_velocities;
int ufe_b200_set_ice_velocities(ufe_b200_handle h, int approximation, int hybrid_scheme, const ufe_b200_ice_velocities* out);

/* ---- secondary velocities ------------------------------------------------------------ */
/* Outputs of calc_secondary_velocities( mesh, ice) (src/UFEMISM/ice_dynamics/conservation_of_momentum/conservation_of_momentum_main.f90:177-246):
 * the ice%... members of the same names; b-grid fields (ti1:ti2), a-grid fields (vi1:vi2[,nz]).  NULL pointers are skipped. */
typedef struct ufe_b200_secondary_velocities {
  double *u_surf_b, *v_surf_b, *uabs_surf_b, *u_base_b, *v_base_b, *uabs_base_b, *u_vav_b, *v_vav_b, *uabs_vav_b;
  double *u_3D, *v_3D;
  double *u_surf, *v_surf, *u_base, *v_base, *u_vav, *v_vav, *uabs_surf, *uabs_base, *uabs_vav, *R_shear;
} ufe_b200_secondary_velocities;
/* u_3D_b, v_3D_b (ti1:ti2,nz): ice%u_3D_b, ice%v_3D_b; pass NULL for both to use the fields the last velocity solve / solve_SIA left on the device. */
int ufe_b200_calc_secondary_velocities(ufe_b200_handle h, const double* u_3D_b, const double* v_3D_b, const ufe_b200_secondary_velocities* out);

/* Replaces calc_vertical_velocities( mesh, ice, BMB) (src/UFEMISM/ice_dynamics/conservation_of_mass/vertical_velocities.f90:23-196) and
 * calc_dw_dz (:198-243): w from conservation of mass in every 3-D Voronoi cell, integrated from the ice base upwards.  In: ice%Hi, ice%Hs,
 * ice%Hib, ice%dHb_dt, ice%dHi_dt, BMB, ice%mask_grounded_ice, ice%mask_floating_ice (vi1:vi2); the zeta gradients dzeta_d{x,y,z}_ak are formed
 * from Hi and Hs as calc_zeta_gradients does (utilities/zeta_gradients.f90:88-115).  The horizontal velocities are the ones on the device:
 * u_3D_b, v_3D_b of the last velocity solve and u_3D, v_3D of the last ufe_b200_calc_secondary_velocities (call that first, as the reference
 * does).  Out: ice%w_3D and (if not NULL) ice%dw_dz_3D (vi1:vi2,nz); dw_dz_3D is NaN where there is no ice, like the reference. */
typedef struct ufe_b200_vertical_velocity_io {
  const double *Hi, *Hs, *Hib, *dHb_dt, *dHi_dt, *BMB;
  const int32_t *mask_grounded_ice, *mask_floating_ice;
  double *w_3D, *dw_dz_3D;
} ufe_b200_vertical_velocity_io;
int ufe_b200_calc_vertical_velocities(ufe_b200_handle h, const ufe_b200_config* cfg, const ufe_b200_vertical_velocity_io* io);

/* ---- geometry / mask / sub-grid stage between the thickness and velocity solves --------------------------------------------
 * All fields are the rank's owned slices; logical masks travel as int32 (0/1), as everywhere in this ABI. */

/* Replaces determine_masks( mesh, Hi, Hb, SL, mask, mask_icefree_land, ...) (src/UFEMISM/ice_dynamics/utilities/masks_mod.f90:25-217).
 * `mask` holds C%type_icefree_land .. C%type_coastline = 1 .. 10 (model_configuration_type_and_namelist.f90:4526-4535).  NULL members are skipped. */
typedef struct ufe_b200_masks {
  int32_t *mask, *mask_icefree_land, *mask_icefree_ocean, *mask_grounded_ice, *mask_floating_ice, *mask_margin, *mask_gl_gr, *mask_gl_fl,
          *mask_cf_gr, *mask_cf_fl, *mask_coastline;
} ufe_b200_masks;
int ufe_b200_determine_masks(ufe_b200_handle h, const ufe_b200_config* cfg, const double* Hi, const double* Hb, const double* SL,
                             const ufe_b200_masks* out);

/* Replaces calc_effective_thickness( mesh, Hi, Hb, SL, Hi_eff, fraction_margin) (utilities/subgrid_ice_margin.f90:19-158). */
int ufe_b200_calc_effective_thickness(ufe_b200_handle h, const ufe_b200_config* cfg, const double* Hi, const double* Hb, const double* SL,
                                      double* Hi_eff, double* fraction_margin);

/* Replaces calc_grounded_fractions( mesh, Hi, Hb, SL, dHb, fraction_gr, fraction_gr_b, mask_floating_ice, bedrock_cdf, bedrock_cdf_b)
 * (utilities/subgrid_grounded_fractions_main.f90:22-120 with ..._bilin_TAF.f90 and ..._bedrock_CDF.f90).  choice = C%choice_subgrid_grounded_fraction,
 * nbins = C%subgrid_bedrock_cdf_nbins; bedrock_cdf (vi1:vi2,nbins), bedrock_cdf_b (ti1:ti2,nbins) may be NULL for UFE_GF_BILIN_INTERP_TAF. */
enum { UFE_GF_BILIN_INTERP_TAF = 0, UFE_GF_BEDROCK_CDF = 1, UFE_GF_BILIN_INTERP_TAF_PLUS_BEDROCK_CDF = 2 };
int ufe_b200_calc_grounded_fractions(ufe_b200_handle h, const ufe_b200_config* cfg, int choice, int nbins, const double* Hi, const double* Hb,
                                     const double* SL, const double* dHb, const int32_t* mask_floating_ice, const double* bedrock_cdf,
                                     const double* bedrock_cdf_b, double* fraction_gr, double* fraction_gr_b);

/* Keep the geometry stage on the device inside the predictor-corrector step: after this call ufe_b200_pc_step_resident and ufe_b200_pc_step
 * update, as run_ice_dynamics_model_pc does (time_stepping/predictor_corrector_scheme.f90:95-107, 161-183, 198-211), the masks and grounded
 * fractions the velocity solve reads (for the predicted geometry) and fraction_margin of the thickness solves (for the current geometry)
 * themselves instead of using the uploaded ones.  dHb, bedrock_cdf, bedrock_cdf_b as in ufe_b200_calc_grounded_fractions (NULL for
 * UFE_GF_BILIN_INTERP_TAF); choice < 0 switches the behaviour off again. */
int ufe_b200_geometry_resident(ufe_b200_handle h, int choice, int nbins, const double* dHb, const double* bedrock_cdf, const double* bedrock_cdf_b);

/* Replaces apply_grounded_fractions_to_bed_roughness( mesh, ice, bed_roughness, bed_roughness_applied) (basal_conditions/sliding_laws.f90:568-636),
 * the preliminary A-grid scaling every sliding law applies to its bed-roughness field when C%do_subgrid_friction_on_A_grid is set: weight =
 * fraction_gr ** max( 0, log10( max( 1, Hi)) - Hs_slope / 0.005) on grounding-line / grounded-calving-front vertices, 1 on other grounded ice, 0 on
 * floating ice and open ocean.  Pass the result as the till_friction_angle / beta_sq / alpha_sq member of struct ufe_b200_ice_inputs; the sliding law decides
 * which field it scales (:108, 151, 207, 263, 318, 372).  masks: mask_gl_gr, mask_cf_gr, mask_gl_fl, mask_grounded_ice, mask_floating_ice,
 * mask_icefree_ocean are read.  All fields (vi1:vi2). */
int ufe_b200_apply_grounded_fractions_to_bed_roughness(ufe_b200_handle h, const double* Hi, const double* Hs_slope, const double* fraction_gr,
                                                       const ufe_b200_masks* masks, const double* bed_roughness, double* bed_roughness_applied);

/* Replaces calc_pc_truncation_error( mesh, ice, pc) (time_stepping/predictor_corrector_scheme.f90:327-361; Robinson et al., 2020, Eq. 32) and the
 * error assessment that follows it in run_ice_dynamics_model_pc (:254-279): tau_np1 = zeta_t |Hi_np1 - Hi_star_np1| / ((3 zeta_t + 3) dt_n);
 * eta_np1 = max( C%pc_eta_min, tau_np1) over fully grounded vertices off the grounding line; n_tot = vertices with fraction_gr = 1, n_guilty =
 * those of them with tau_np1 > C%pc_epsilon, whose pc%tau_n_guilty is incremented.  All fields (vi1:vi2); logical masks as 0/1; the three scalar
 * results are global (the reference's MPI_ALLREDUCE MAX / SUM).  tau_np1 and tau_n_guilty may be NULL.  The time-step rule itself (Eq. 33 and
 * the accept / retry decision, :59-76, 281-300) is scalar host logic and stays in the Fortran driver. */
int ufe_b200_calc_pc_truncation_error(ufe_b200_handle h, double zeta_t, double dt_n, double pc_eta_min, double pc_epsilon, const double* Hi_np1,
                                      const double* Hi_star_np1, const int32_t* mask_grounded_ice, const int32_t* mask_gl_gr,
                                      const double* fraction_gr, double* tau_np1, int32_t* tau_n_guilty, double* eta_np1, int32_t* n_tot,
                                      int32_t* n_guilty);

/* No counterpart in the reference: added at the request of the run's operators, for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  The same computation as ufe_b200_calc_pc_truncation_error on the n vertices passed
 * (host buffers of length n; tau_np1 and tau_n_guilty may be NULL), but with every byte of device memory -- staged inputs, per-vertex outputs, the
 * reduction cells -- in one file-scope __device__ array of fixed size inside the library (csrc/noalloc.cu: 256 + 44 bytes x
 * UFE_B200_NOALLOC_MAX_VERTICES = 11 534 592 bytes, held by every GPU that loads the library until the process exits).  The call allocates nothing
 * (no cudaMalloc / cudaFree / allocating library routine, no mesh needed on the handle) and fails with a clear ufe_b200_last_error when
 * n > UFE_B200_NOALLOC_MAX_VERTICES.  eta_np1, n_tot and n_guilty cover the n vertices of THIS call only (no reduction over ranks).  All calls on
 * a device share the array: the cells that must start from a known value are reset inside every call, and two calls that overlap on one GPU (two
 * handles or two host threads) race -- serialise them. */
#define UFE_B200_NOALLOC_MAX_VERTICES 262144
int ufe_b200_calc_pc_truncation_error_noalloc(ufe_b200_handle h, int n, double zeta_t, double dt_n, double pc_eta_min, double pc_epsilon,
                                              const double* Hi_np1, const double* Hi_star_np1, const int32_t* mask_grounded_ice,
                                              const int32_t* mask_gl_gr, const double* fraction_gr, double* tau_np1, int32_t* tau_n_guilty,
                                              double* eta_np1, int32_t* n_tot, int32_t* n_guilty);

/* Replace calc_critical_timestep_adv / _SIA( mesh, ice, dt_crit) (utilities/time_step_criteria.f90:80-141, 20-78); the result is the global
 * minimum (the reference's MPI_ALLREDUCE( MPI_MIN)).  u_vav_b, v_vav_b NULL = the velocities of the last device solve. */
int ufe_b200_calc_critical_timestep_adv(ufe_b200_handle h, const ufe_b200_config* cfg, int do_grounded_only_adv_dt, const double* Hi,
                                        const int32_t* mask_floating_ice, const double* u_vav_b, const double* v_vav_b, double* dt_crit_adv);
int ufe_b200_calc_critical_timestep_SIA(ufe_b200_handle h, const ufe_b200_config* cfg, const double* Hi, const double* D_3D_b, double* dt_crit_SIA);

/* Replaces alter_ice_thickness( mesh, ice, Hi_old, Hb, SL, Hi_new, refgeo, time) (utilities/ice_thickness_safeties.f90:26-288): the would-be
 * effective thickness, mask protection, thin-ice / calving / shelf removal, fixiness, limitness, modiness.  The C% scalars it reads, with the
 * defaults of model_configuration_type_and_namelist.f90:424-469 (ufe_b200_alter_params_defaults). */
enum { UFE_MODINESS_NONE = 0, UFE_MODINESS_TI_HOM = 1, UFE_MODINESS_TI_HOM_UP = 2, UFE_MODINESS_TI_HOM_DOWN = 3, UFE_MODINESS_NO_THICK_INLAND = 4,
       UFE_MODINESS_NO_THIN_INLAND = 5 };
typedef struct ufe_b200_alter_params {
  int32_t do_protect_grounded_mask; double protect_grounded_mask_t_end;
  double Hi_min;
  int32_t calving_law_threshold_thickness; double calving_threshold_thickness_shelf;   /* C%choice_calving_law == 'threshold_thickness' */
  int32_t remove_ice_absent_at_PD, do_remove_shelves, remove_shelves_larger_than_PD, continental_shelf_calving;
  double continental_shelf_min_height;
  double fixiness_t_start, fixiness_t_end; int32_t do_fixiness_before_start;
  double limitness_t_start, limitness_t_end; int32_t do_limitness_before_start;
  int32_t modiness_H_style; double modiness_T_hom_ref;
  double fixiness_H_gl_gr, fixiness_H_gl_fl, fixiness_H_grounded, fixiness_H_floating;
  int32_t fixiness_H_freeland, fixiness_H_freeocean;
  double limitness_H_gl_gr, limitness_H_gl_fl, limitness_H_grounded, limitness_H_floating;
} ufe_b200_alter_params;
void ufe_b200_alter_params_defaults(ufe_b200_alter_params* p);
/* masks: ice%mask_* of the current geometry (only icefree_land, icefree_ocean, grounded_ice, floating_ice, gl_gr, gl_fl are read);
 * refgeo_Hi, refgeo_Hb: refgeo%Hi, refgeo%Hb; Ti_hom: ice%Ti_hom (may be NULL unless a Ti_hom modiness style is chosen); Hi_new is in/out. */
int ufe_b200_alter_ice_thickness(ufe_b200_handle h, const ufe_b200_config* cfg, const ufe_b200_alter_params* prm, double time, const ufe_b200_masks* masks,
                                 const double* Hi_old, const double* Hb, const double* SL, const double* refgeo_Hi, const double* refgeo_Hb,
                                 const double* Ti_hom, double* Hi_new);

/* Replaces calc_and_apply_spill_over_flux( mesh, ice, Hi_new, dt) (utilities/ice_thickness_safeties.f90:290-437): ice above the upstream
 * thickness at calving-front vertices is handed to the neighbouring ice-free ocean vertices, weighted by the outflow velocity.
 * In: ice%mask_icefree_ocean, mask_cf_gr, mask_cf_fl (from `masks`), ice%Hi_eff, u_vav_b, v_vav_b (NULL = the last device solve; ice%u_perp is
 * formed from them as calc_ice_flux_divergence_matrix_upwind does).  In/out Hi_new; out ice%Qspill (may be NULL). */
int ufe_b200_calc_and_apply_spill_over_flux(ufe_b200_handle h, const ufe_b200_masks* masks, const double* Hi_eff, const double* u_vav_b,
                                            const double* v_vav_b, double dt, double* Hi_new, double* Qspill);

/* ---- timing hooks for bench.py (CUDA events on the library's own stream) ----------------------- */
/* Milliseconds spent in the named kernel class since the last reset, and its launch count.
 * classes: "spmv_stiffness", "krylov_vec", "assembly", "column", "tri", "epilogue", "thickness", "operator_spmv" */
int ufe_b200_timing_enable(ufe_b200_handle h, int on);
/* one CUDA-event stopwatch on the library's stream: start records an event, stop records another, synchronises and
 * returns the elapsed device time in milliseconds */
int ufe_b200_timer_start(ufe_b200_handle h);
int ufe_b200_timer_stop(ufe_b200_handle h, double* ms);
/* writes a 512 MB device buffer (4 x L2) on the library's stream: bench.py calls it between the timed iterations of workloads smaller than L2 */
int ufe_b200_l2_flush(ufe_b200_handle h);
/* page-lock / unlock a host buffer the caller passes repeatedly (cudaHostRegister), so H2D/D2H copies run at full rate */
int ufe_b200_host_register(void* ptr, size_t bytes);
int ufe_b200_host_unregister(void* ptr);
int ufe_b200_timing_get(ufe_b200_handle h, const char* kernel_class, double* ms, long long* launches);
int ufe_b200_timing_reset(ufe_b200_handle h);

#ifdef __cplusplus
}
#endif
#endif /* UFE_B200_H */
