// TEST INFRASTRUCTURE ONLY -- CPU restatement ("oracle") of UFEMISM's stress-balance / ice-thickness
// hot path.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg
// may link or call this; the product (ufemism2.0_b200/csrc) never does.
//
// Parity status: the reference is Fortran 2018 + PETSc + MPI + NetCDF and cannot be built in this
// image (no Fortran compiler).  The oracle is pinned against the reference's own
// known-answer vectors that are usable as text (tests/golden/README.md): the 6x6 and two 7x7 CSR
// systems, the partition_list table, the dummy-mesh connectivity, polynomial exactness of the
// operators (what ct_discretisation_mapping_derivatives checks), the Laplace component test's
// analytic solution, and the Halfar / Schoof analytic solutions; on output of the reference itself:
// element-wise / to 1e-11 on the three meshes it ships and on what its component tests computed on
// them (secondary mesh data, 15 operators, Laplace solve through the KSP seam; tests/golden/h5lite.py
// reads its NetCDF-4 files, tests/test_reference_meshes.py), and statistically on the per-iteration
// checksum logs of four of its integrated tests.  Field-level parity of a stress-balance solve with a
// *running* reference (PETSc KSP output on a mesh it ships) remains a statistical pin.
//
// Conventions: all arrays are the reference's own (column-major, 1-based indices, 0 = none).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace orc {

// type_sparse_matrix_CSR_dp restricted to one process (CSR_sparse_matrix_type.f90:15-38)
struct Csr {
  int m = 0, n = 0;
  std::vector<int> ptr;     // size m+1, 1-based offsets into ind/val (ptr[0] == 1)
  std::vector<int> ind;     // 1-based column indices, construction order (never sorted, never merged)
  std::vector<double> val;
  void init(int m_, int n_, size_t nnz_est = 0) {
    m = m_; n = n_; ptr.assign((size_t)m + 1, 1); ind.clear(); val.clear(); ind.reserve(nnz_est); val.reserve(nnz_est);
  }
  // add_entry_CSR_dist (CSR_matrix_basics.f90:251-277): rows must be filled in ascending order
  void add(int i, int j, double v) { ind.push_back(j); val.push_back(v); ptr[i] = (int)ind.size() + 1; }
  // rows that received no entry inherit the running offset (what crop/finalise leave behind)
  void finalise() { for (int i = 1; i <= m; ++i) if (ptr[i] < ptr[i - 1]) ptr[i] = ptr[i - 1]; }
  size_t nnz() const { return ind.size(); }
};

// y = A x  (CSR_matrix_vector_multiplication.f90:318-330)
void spmv(const Csr& A, const double* x, double* y);

struct MeshView {  // borrowed pointers to the Fortran arrays
  int nV, nTri, nE, nC_mem, nz;
  double xmin, xmax, ymin, ymax;
  const double* V; const int* nC; const int* C; const int* niTri; const int* iTri; const int* VBI;
  const int* Tri; const int* TriC; const int* TriBI; const double* TriGC;
  const double* A; const double* Cw; const double* D_x; const double* D_y; const double* D;
  const int* VE; const int* EV; const int* ETri; const double* zeta;
  const double* Tricc = nullptr;   // (nTri,2) circumcentres (Voronoi vertices); only the grounded fractions need them
};

enum OpId {
  OP_ddx_a_a = 0, OP_ddy_a_a, OP_map_a_b, OP_ddx_a_b, OP_ddy_a_b, OP_map_b_a, OP_ddx_b_a, OP_ddy_b_a,
  OP_ddx_b_b, OP_ddy_b_b, OP2_ddx_b_b, OP2_ddy_b_b, OP2_d2dx2_b_b, OP2_d2dxdy_b_b, OP2_d2dy2_b_b, OP_COUNT
};

struct Operators { Csr M[OP_COUNT]; };

// calc_all_matrix_operators_mesh (mesh_disc_calc_matrix_operators_2D.f90:26-59)
void calc_all_matrix_operators(const MeshView& mesh, Operators& ops);

// --- configuration scalars the path reads (model_configuration_type_and_namelist.f90) -----------------
enum SlidingLaw { SLID_no_sliding = 0, SLID_Weertman = 1, SLID_Coulomb = 2, SLID_Budd = 3, SLID_Tsai2015 = 4,
                  SLID_Schoof2005 = 5, SLID_ZoetIverson = 6, SLID_idealised_SSA_icestream = 7 };
enum BC { BC_infinite = 0, BC_zero = 1, BC_periodic_ISMIP_HOM = 2, BC_infinite_SSA_icestream = 3 };
enum KspType { KSP_bicgstab_jacobi = 0, KSP_gmres_ilu0 = 1, KSP_gmres_bjacobi_ilu0 = 2 };

struct Config {
  // viscosity iteration (…type_and_namelist.f90:300-312)
  double visc_it_norm_dUV_tol = 5e-5; int visc_it_nit = 50; double visc_it_relax = 0.2;
  double visc_eff_min = 1e4; double vel_max = 5000.0;
  double stress_balance_PETSc_rtol = 1e-7, stress_balance_PETSc_abstol = 1e-5;
  double Glens_flow_law_exponent = 3.0, Glens_flow_law_epsilon_sq_0 = 1e-8;
  int do_include_SSADIVA_crossterms = 1;
  int BC_u[4] = {0, 0, 0, 0}, BC_v[4] = {0, 0, 0, 0};  // north, east, south, west
  // sliding
  int choice_sliding_law = SLID_ZoetIverson;
  double slid_delta_v = 1e-3, slid_Weertman_m = 3.0, slid_Budd_q_plastic = 0.3, slid_Budd_u_threshold = 100.0;
  double slid_ZI_ut = 200.0, slid_ZI_p = 5.0, slid_beta_max = 1e20;
  int do_GL_subgrid_friction = 1; double subgrid_friction_exponent_on_B_grid = 2.0;
  double Hi_min = 0.0;
  // SSA_icestream idealised law (Schoof 2006)
  double uniform_Glens_flow_factor = 1e-16;
  double icestream_Hi = 2000.0, icestream_dhdx = -0.0003, icestream_L = 150e3, icestream_m = 1.0;
  double ISMIP_HOM_L = 0.0;  // refgeo_idealised_ISMIP_HOM_L: period of the 'periodic_ISMIP-HOM' copy rows
  // thickness
  int choice_ice_integration_method = 1;  // 0 explicit, 1 semi-implicit
  double dHi_semiimplicit_fs = 1.5, dHi_PETSc_rtol = 1e-8, dHi_PETSc_abstol = 1e-6;
  double dt_ice_min = 0.1, dt_ice_max = 10.0;
  int BC_H[4] = {1, 1, 1, 1};             // 0 infinite, 1 zero
  // physical constants (parameters.f90)
  double ice_density = 917.0, seawater_density = 1027.0, grav = 9.81;
  // which Krylov method the oracle uses
  int ksp_type = KSP_bicgstab_jacobi; int ksp_maxit = 2000;
  int ksp_initial_guess_nonzero = 0;  // petsc_basic.f90:163 (commented out in the reference => 0)
};

// --- linear solvers ---------------------------------------------------------------------------------
// Both follow PETSc semantics at the reference's call site (petsc_basic.f90:125-237): zero initial
// guess, left preconditioning, convergence on ||B r||_2 <= max(rtol ||B b||_2, abstol), <= maxit its.
int ksp_bicgstab_jacobi(const Csr& A, const double* b, double* x, double rtol, double abstol, int maxit, int* its, bool use_guess = false);
int ksp_gmres_ilu0(const Csr& A, const double* b, double* x, double rtol, double abstol, int maxit, int restart, int* its, bool use_guess = false);
// the same with PETSc's default PC on `nblocks` MPI ranks: block Jacobi, one ILU(0) per rank's row block (partition_list)
void set_bjacobi_blocks(int n);   // 0 = one block per OpenMP thread
int bjacobi_blocks();
int ksp_solve(int ksp_type, const Csr& A, const double* b, double* x, double rtol, double abstol, int maxit, int* its, bool use_guess = false);
int ksp_gmres_bjacobi_ilu0(const Csr& A, int nblocks, const double* b, double* x, double rtol, double abstol, int maxit, int restart, int* its,
                           bool use_guess = false);

// --- linearised SSA/DIVA (solve_linearised_SSA_DIVA_infinite_slab.f90) --------------------------------
void assemble_SSA_DIVA(const MeshView& mesh, const Operators& ops, const Config& C,
                       const double* N_b, const double* dN_dx_b, const double* dN_dy_b, const double* beta_b,
                       const double* tau_dx_b, const double* tau_dy_b,
                       const int* BC_prescr_mask_b, const double* BC_prescr_u_b, const double* BC_prescr_v_b,
                       Csr& A, std::vector<double>& bb, const double* u_b_prev = nullptr, const double* v_b_prev = nullptr);
// find_containing_vertex (mesh_utilities.f90:1368), find_ti_copy_{ISMIP_HOM_periodic,SSA_icestream_infinite} (:2620, :2681)
int find_containing_vertex(const MeshView& mesh, const double p[2], int vi);
void find_ti_copy(const MeshView& mesh, int choice, double L, int ti, int* ti_copy, double* wti_copy);

struct DivaState {  // type_ice_velocity_solver_DIVA (ice_model_types.f90:149-197), fields on the path
  std::vector<double> u_vav_b, v_vav_b, u_base_b, v_base_b, u_b_prev, v_b_prev;
  std::vector<double> tau_dx_b, tau_dy_b, tau_bx_b, tau_by_b;
  std::vector<double> du_dx_a, du_dy_a, dv_dx_a, dv_dy_a, du_dz_3D_a, dv_dz_3D_a;
  std::vector<double> eta_3D_a, eta_3D_b, eta_vav_a, N_a, N_b, dN_dx_b, dN_dy_b;
  std::vector<double> F1_3D_a, F2_3D_a, F1_3D_b, F2_3D_b;
  std::vector<double> basal_friction_coefficient_a, basal_friction_coefficient_b, beta_eff_a, beta_eff_b;
  std::vector<double> u_3D_b, v_3D_b;
  void init(int nV, int nTri, int nz);
};

struct IceInputs {  // the type_ice_model / bed_roughness fields the velocity solve reads
  const double* Hi; const double* Hs; const double* A_flow_3D;   // (nV), (nV), (nV,nz)
  const double* fraction_gr_b;                                    // (nTri)
  const double* till_friction_angle; const double* beta_sq; const double* alpha_sq;  // (nV) bed roughness
  const double* effective_pressure;                               // (nV)
  const int* mask_grounded_ice; const int* mask_icefree_land;      // (nV) 0/1
};

// solve_DIVA_infinite_slab (DIVA_solver_infinite_slab.f90:52-231); returns 0 ok
int solve_DIVA(const MeshView& mesh, const Operators& ops, const Config& C, const IceInputs& ice, DivaState& D,
               const int* BC_prescr_mask_b, const double* BC_prescr_u_b, const double* BC_prescr_v_b,
               int* n_visc_its, int* n_Axb_its, int max_visc_its_override = -1);
// solve_SSA (SSA_main.f90:88-245); uses u_vav_b/v_vav_b of D as u_b/v_b
int solve_SSA(const MeshView& mesh, const Operators& ops, const Config& C, const IceInputs& ice, DivaState& D,
              const int* BC_prescr_mask_b, const double* BC_prescr_u_b, const double* BC_prescr_v_b,
              int* n_visc_its, int* n_Axb_its, int max_visc_its_override = -1);

// --- ice thickness (conservation_of_mass_*.f90) -------------------------------------------------------
struct ThicknessIO {
  const double* Hi; const double* Hb; const double* SL; const double* u_vav_b; const double* v_vav_b;
  const double* SMB; const double* BMB; const double* LMB; const double* fraction_margin; const int* mask_noice;
  const double* dHi_dt_target; const int* BC_prescr_mask; const double* BC_prescr_Hi;  // may be null
  double* AMB; double* dHi_dt; double* Hi_tplusdt; double* divQ;
};
int calc_dHi_dt(const MeshView& mesh, const Config& C, ThicknessIO& io, double* dt, int* n_Axb_its);
void calc_flux_divergence_matrix_upwind(const MeshView& mesh, const double* u_vav_b, const double* v_vav_b,
                                        const double* fraction_margin, Csr& M_divQ, std::vector<double>* u_perp = nullptr);

// --- apply_grounded_fractions_to_bed_roughness (basal_conditions/sliding_laws.f90:568-636), do_subgrid_friction_on_A_grid = .true. ----------
void apply_grounded_fractions_to_bed_roughness(int nV, const double* Hi, const double* Hs_slope, const double* fraction_gr, const int* mask_gl_gr,
                                               const int* mask_cf_gr, const int* mask_gl_fl, const int* mask_grounded_ice, const int* mask_floating_ice,
                                               const int* mask_icefree_ocean, const double* bed_roughness, double* bed_roughness_applied);

// --- calc_pc_truncation_error + error assessment (time_stepping/predictor_corrector_scheme.f90:327-361, 254-279) ---------
void calc_pc_truncation_error(int nV, double zeta_t, double dt_n, double pc_eta_min, double pc_epsilon, const double* Hi_np1, const double* Hi_star_np1,
                              const int* mask_grounded_ice, const int* mask_gl_gr, const double* fraction_gr, double* tau_np1, int* tau_n_guilty,
                              double* eta_np1, int* n_tot, int* n_guilty);

// --- set_ice_velocities_to_{SIA,SSA,SIASSA,DIVA}_results (conservation_of_momentum_main.f90:311-511) -------
// approximation: 0 SSA, 1 DIVA, 2 SIA, 3 SIA/SSA; scheme: 0 'add', 1 'add_SIA_reduced'.  Inputs: the solver objects' fields (NULL where the
// approximation has none); outputs: ice%u_3D_b, v_3D_b (nTri,nz) and the eight 3-D strain rates (nV,nz), NULL = skipped.
struct SolverResults {
  const double *u_3D_b = nullptr, *v_3D_b = nullptr;               // SIA% or DIVA% (nTri,nz)
  const double *u_b = nullptr, *v_b = nullptr;                     // SSA% (nTri)
  const double *du_dx_a = nullptr, *du_dy_a = nullptr, *dv_dx_a = nullptr, *dv_dy_a = nullptr;   // SSA% / DIVA% (nV)
  const double *du_dz_3D = nullptr, *dv_dz_3D = nullptr;           // SIA% / DIVA% (nV,nz)
};
struct IceVelocities { double *u_3D_b, *v_3D_b, *du_dx_3D, *du_dy_3D, *du_dz_3D, *dv_dx_3D, *dv_dy_3D, *dv_dz_3D, *dw_dx_3D, *dw_dy_3D; };
int set_ice_velocities(const MeshView& mesh, int approximation, int scheme, const SolverResults& R, IceVelocities& I);

// --- secondary velocities (conservation_of_momentum_main.f90:177-246): every pointer is an output the caller owns
struct SecondaryVelocities {
  double *u_surf_b, *v_surf_b, *uabs_surf_b, *u_base_b, *v_base_b, *uabs_base_b, *u_vav_b, *v_vav_b, *uabs_vav_b;   // (nTri)
  double *u_3D, *v_3D;                                                                                             // (nV,nz)
  double *u_surf, *v_surf, *u_base, *v_base, *u_vav, *v_vav, *uabs_surf, *uabs_base, *uabs_vav, *R_shear;           // (nV)
};
void calc_secondary_velocities(const MeshView& mesh, const Operators& ops, const double* u_3D_b, const double* v_3D_b, SecondaryVelocities& S);

// --- geometry / mask / sub-grid stage between the device phases (geometry.cpp)
struct MaskSet { int *mask, *icefree_land, *icefree_ocean, *grounded_ice, *floating_ice, *margin, *gl_gr, *gl_fl, *cf_gr, *cf_fl, *coastline; };
enum GroundedFractionChoice { GF_bilin_interp_TAF = 0, GF_bedrock_CDF = 1, GF_bilin_interp_TAF_plus_bedrock_CDF = 2 };
void determine_masks(const MeshView& mesh, const Config& C, const double* Hi, const double* Hb, const double* SL, const MaskSet& out);
void calc_effective_thickness(const MeshView& mesh, const Config& C, const double* Hi, const double* Hb, const double* SL, double* Hi_eff,
                              double* fraction_margin);
void calc_grounded_fractions(const MeshView& mesh, const Operators& ops, const Config& C, int choice, int nbins, const double* Hi, const double* Hb,
                             const double* SL, const double* dHb, const int* mask_floating_ice, const double* bedrock_cdf,
                             const double* bedrock_cdf_b, double* fraction_gr, double* fraction_gr_b);
// alter_ice_thickness / calc_and_apply_spill_over_flux (ice_thickness_safeties.f90:26-437); AlterParams = the C% scalars (same members as
// ufe_b200_alter_params, include/ufe_b200.h)
struct AlterParams {
  int do_protect_grounded_mask; double protect_grounded_mask_t_end;
  double Hi_min;
  int calving_law_threshold_thickness; double calving_threshold_thickness_shelf;
  int remove_ice_absent_at_PD, do_remove_shelves, remove_shelves_larger_than_PD, continental_shelf_calving;
  double continental_shelf_min_height;
  double fixiness_t_start, fixiness_t_end; int do_fixiness_before_start;
  double limitness_t_start, limitness_t_end; int do_limitness_before_start;
  int modiness_H_style; double modiness_T_hom_ref;
  double fixiness_H_gl_gr, fixiness_H_gl_fl, fixiness_H_grounded, fixiness_H_floating;
  int fixiness_H_freeland, fixiness_H_freeocean;
  double limitness_H_gl_gr, limitness_H_gl_fl, limitness_H_grounded, limitness_H_floating;
};
void alter_ice_thickness(const MeshView& mesh, const Config& C, const AlterParams& P, double time, const MaskSet& masks, const double* Hi_old,
                         const double* Hb, const double* SL, const double* refgeo_Hi, const double* refgeo_Hb, const double* Ti_hom, double* Hi_new);
void calc_and_apply_spill_over_flux(const MeshView& mesh, const MaskSet& masks, const double* Hi_eff, const double* u_vav_b, const double* v_vav_b,
                                    double dt, double* Hi_new, double* Qspill);
double calc_critical_timestep_SIA(const MeshView& mesh, const Config& C, const double* Hi, const double* D_3D_b);
double calc_critical_timestep_adv(const MeshView& mesh, const Config& C, int do_grounded_only_adv_dt, const double* Hi, const int* mask_floating_ice,
                                  const double* u_vav_b, const double* v_vav_b);

// --- vertical velocities (conservation_of_mass/vertical_velocities.f90:23-243); the zeta gradients follow zeta_gradients.f90:88-115
struct VerticalVelocityIO {
  const double *Hi, *Hs, *Hib, *dHb_dt, *dHi_dt, *BMB;     // (nV)
  const int *mask_grounded_ice, *mask_floating_ice;         // (nV)
  double *w_3D, *dw_dz_3D;                                  // (nV,nz) out (dw_dz_3D may be null)
};
void calc_vertical_velocities(const MeshView& mesh, const Operators& ops, const Config& C, const VerticalVelocityIO& io, const double* u_3D_b,
                              const double* v_3D_b, const double* u_3D, const double* v_3D);
void ddzeta_operator(const double* zeta, int nz, std::vector<int>& nb, std::vector<double>& coef);

// --- SIA (SIA_main.f90:65-158) and analytic solutions ------------------------------------------------
void solve_SIA(const MeshView& mesh, const Operators& ops, const Config& C, double SIA_maximum_diffusivity, const double* Hi, const double* Hs,
               const double* A_flow_3D, double* u_3D_b, double* v_3D_b, double* D_3D_b, double* du_dz_3D_a, double* dv_dz_3D_a);
double Halfar_H(double A, double n, double H0, double R0, double x, double y, double t_yr);
void Schoof2006_icestream(double A, double n, double H, double dhdx, double L, double m, double y, double* u, double* tauc);

}  // namespace orc
