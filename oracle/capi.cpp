// TEST INFRASTRUCTURE ONLY (see oracle.hpp).  C entry points of liboracle.so for ctypes.
#include <cstring>
#include <memory>

#include <omp.h>

#include "../include/ufe_b200.h"
#include "oracle.hpp"

using namespace orc;

namespace orc_capi {

struct OracleMesh {
  MeshView mv;
  Operators ops;
  bool have_ops = false;
  Csr A; std::vector<double> bb;   // last assembled stiffness system
  Csr MdivQ;
  DivaState D;
};

Config to_config(const ufe_b200_config* c) {
  Config C;
  C.visc_it_norm_dUV_tol = c->visc_it_norm_dUV_tol; C.visc_it_nit = c->visc_it_nit; C.visc_it_relax = c->visc_it_relax;
  C.visc_eff_min = c->visc_eff_min; C.vel_max = c->vel_max;
  C.stress_balance_PETSc_rtol = c->stress_balance_PETSc_rtol; C.stress_balance_PETSc_abstol = c->stress_balance_PETSc_abstol;
  C.Glens_flow_law_exponent = c->Glens_flow_law_exponent; C.Glens_flow_law_epsilon_sq_0 = c->Glens_flow_law_epsilon_sq_0;
  C.do_include_SSADIVA_crossterms = c->do_include_SSADIVA_crossterms;
  for (int k = 0; k < 4; ++k) { C.BC_u[k] = c->BC_u[k]; C.BC_v[k] = c->BC_v[k]; C.BC_H[k] = c->BC_H[k]; }
  C.choice_sliding_law = c->choice_sliding_law;
  C.slid_delta_v = c->slid_delta_v; C.slid_Weertman_m = c->slid_Weertman_m; C.slid_Budd_q_plastic = c->slid_Budd_q_plastic;
  C.slid_Budd_u_threshold = c->slid_Budd_u_threshold; C.slid_ZI_ut = c->slid_ZI_ut; C.slid_ZI_p = c->slid_ZI_p;
  C.slid_beta_max = c->slid_beta_max; C.do_GL_subgrid_friction = c->do_GL_subgrid_friction;
  C.subgrid_friction_exponent_on_B_grid = c->subgrid_friction_exponent_on_B_grid; C.Hi_min = c->Hi_min;
  C.uniform_Glens_flow_factor = c->uniform_Glens_flow_factor;
  C.icestream_Hi = c->refgeo_idealised_SSA_icestream_Hi; C.icestream_dhdx = c->refgeo_idealised_SSA_icestream_dhdx;
  C.icestream_L = c->refgeo_idealised_SSA_icestream_L; C.icestream_m = c->refgeo_idealised_SSA_icestream_m;
  C.ISMIP_HOM_L = c->refgeo_idealised_ISMIP_HOM_L;
  C.choice_ice_integration_method = c->choice_ice_integration_method;
  C.dHi_semiimplicit_fs = c->dHi_semiimplicit_fs; C.dHi_PETSc_rtol = c->dHi_PETSc_rtol; C.dHi_PETSc_abstol = c->dHi_PETSc_abstol;
  C.dt_ice_min = c->dt_ice_min; C.dt_ice_max = c->dt_ice_max;
  C.ice_density = c->ice_density; C.seawater_density = c->seawater_density; C.grav = c->grav;
  C.ksp_maxit = c->ksp_maxit;
  C.ksp_initial_guess_nonzero = c->ksp_initial_guess_nonzero;
  return C;
}

Csr borrow_csr(int m, int n, const int* ptr, const int* ind, const double* val) {
  Csr A; A.m = m; A.n = n;
  A.ptr.assign(ptr, ptr + m + 1);
  const int nnz = ptr[m] - 1;
  A.ind.assign(ind, ind + nnz); A.val.assign(val, val + nnz);
  return A;
}

}  // namespace orc_capi
using namespace orc_capi;

extern "C" {

void orc_spmv(int m, int n, const int* ptr, const int* ind, const double* val, const double* x, double* y) {
  Csr A = borrow_csr(m, n, ptr, ind, val);
  spmv(A, x, y);
}

// ksp: 0 = BiCGStab + Jacobi (the CUDA algorithm), 1 = GMRES(30) + ILU(0) (the reference's PETSc default on one rank), 2 = GMRES(30) + block Jacobi of
// ILU(0) blocks (the default on several ranks; orc_set_bjacobi_blocks, default one block per OpenMP thread)
void orc_set_num_threads(int n) { if (n > 0) omp_set_num_threads(n); }   // (torch.distributed.run exports OMP_NUM_THREADS=1 to its children)
int orc_max_threads() { return omp_get_max_threads(); }
void orc_set_bjacobi_blocks(int n) { set_bjacobi_blocks(n); }
int orc_bjacobi_blocks() { return bjacobi_blocks(); }
int orc_ksp_solve(int m, const int* ptr, const int* ind, const double* val, const double* b, double* x, double rtol,
                  double abstol, int maxit, int ksp, int* its) {
  Csr A = borrow_csr(m, m, ptr, ind, val);
  return ksp_solve(ksp, A, b, x, rtol, abstol, maxit, its);
}

void* orc_mesh_create(const ufe_b200_mesh_desc* d) {
  auto* o = new OracleMesh();
  MeshView& m = o->mv;
  m.nV = d->nV; m.nTri = d->nTri; m.nE = d->nE; m.nC_mem = d->nC_mem; m.nz = d->nz;
  m.xmin = d->xmin; m.xmax = d->xmax; m.ymin = d->ymin; m.ymax = d->ymax;
  m.V = d->V; m.nC = d->nC; m.C = d->C; m.niTri = d->niTri; m.iTri = d->iTri; m.VBI = d->VBI; m.Tri = d->Tri; m.TriC = d->TriC;
  m.TriBI = d->TriBI; m.TriGC = d->TriGC; m.A = d->A; m.Cw = d->Cw; m.D_x = d->D_x; m.D_y = d->D_y; m.D = d->D; m.VE = d->VE;
  m.EV = nullptr; m.ETri = d->ETri; m.zeta = d->zeta; m.Tricc = d->Tricc;
  return o;
}
void orc_mesh_destroy(void* h) { delete static_cast<OracleMesh*>(h); }

void orc_calc_matrix_operators(void* h) {
  auto* o = static_cast<OracleMesh*>(h);
  calc_all_matrix_operators(o->mv, o->ops);
  o->have_ops = true;
}
void orc_get_operator_size(void* h, int op, int* m, int* n, int* nnz) {
  const Csr& M = static_cast<OracleMesh*>(h)->ops.M[op];
  *m = M.m; *n = M.n; *nnz = (int)M.nnz();
}
void orc_get_operator(void* h, int op, int* ptr, int* ind, double* val) {
  const Csr& M = static_cast<OracleMesh*>(h)->ops.M[op];
  std::memcpy(ptr, M.ptr.data(), sizeof(int) * M.ptr.size());
  std::memcpy(ind, M.ind.data(), sizeof(int) * M.ind.size());
  std::memcpy(val, M.val.data(), sizeof(double) * M.val.size());
}
void orc_apply_operator(void* h, int op, const double* x, double* y, int ncols) {
  const Csr& M = static_cast<OracleMesh*>(h)->ops.M[op];
  for (int k = 0; k < ncols; ++k) spmv(M, x + (size_t)k * M.n, y + (size_t)k * M.m);
}

void orc_assemble_SSA_DIVA(void* h, const ufe_b200_config* cfg, const double* N_b, const double* dN_dx_b, const double* dN_dy_b,
                           const double* beta_b, const double* tau_dx_b, const double* tau_dy_b, const int* BC_mask,
                           const double* BC_u, const double* BC_v, int* m, int* nnz) {
  auto* o = static_cast<OracleMesh*>(h);
  assemble_SSA_DIVA(o->mv, o->ops, to_config(cfg), N_b, dN_dx_b, dN_dy_b, beta_b, tau_dx_b, tau_dy_b, BC_mask, BC_u, BC_v, o->A, o->bb);
  *m = o->A.m; *nnz = (int)o->A.nnz();
}
// same with the previous solution the copy rows ('periodic_ISMIP-HOM', 'infinite_SSA_icestream') read
void orc_assemble_SSA_DIVA_prev(void* h, const ufe_b200_config* cfg, const double* N_b, const double* dN_dx_b, const double* dN_dy_b,
                                const double* beta_b, const double* tau_dx_b, const double* tau_dy_b, const int* BC_mask,
                                const double* BC_u, const double* BC_v, const double* u_b_prev, const double* v_b_prev, int* m, int* nnz) {
  auto* o = static_cast<OracleMesh*>(h);
  assemble_SSA_DIVA(o->mv, o->ops, to_config(cfg), N_b, dN_dx_b, dN_dy_b, beta_b, tau_dx_b, tau_dy_b, BC_mask, BC_u, BC_v, o->A, o->bb,
                    u_b_prev, v_b_prev);
  *m = o->A.m; *nnz = (int)o->A.nnz();
}
void orc_find_ti_copy(void* h, int choice, double L, int ti, int* ti_copy, double* wti_copy) {
  find_ti_copy(static_cast<OracleMesh*>(h)->mv, choice, L, ti, ti_copy, wti_copy);
}
void orc_get_stiffness(void* h, int* ptr, int* ind, double* val, double* bb) {
  auto* o = static_cast<OracleMesh*>(h);
  std::memcpy(ptr, o->A.ptr.data(), sizeof(int) * o->A.ptr.size());
  std::memcpy(ind, o->A.ind.data(), sizeof(int) * o->A.ind.size());
  std::memcpy(val, o->A.val.data(), sizeof(double) * o->A.val.size());
  std::memcpy(bb, o->bb.data(), sizeof(double) * o->bb.size());
}

static IceInputs to_ice(const ufe_b200_ice_inputs* i) {
  IceInputs ice;
  ice.Hi = i->Hi; ice.Hs = i->Hs; ice.A_flow_3D = i->A_flow; ice.fraction_gr_b = i->fraction_gr_b;
  ice.till_friction_angle = i->till_friction_angle; ice.beta_sq = i->beta_sq; ice.alpha_sq = i->alpha_sq;
  ice.effective_pressure = i->effective_pressure; ice.mask_grounded_ice = i->mask_grounded_ice; ice.mask_icefree_land = i->mask_icefree_land;
  return ice;
}

static void state_in(DivaState& D, const ufe_b200_velocity_state* s, int nV, int nTri, int nz) {
  D.init(nV, nTri, nz);
  auto cp = [](std::vector<double>& dst, const double* src) { if (src) std::memcpy(dst.data(), src, sizeof(double) * dst.size()); };
  cp(D.u_vav_b, s->u_vav_b); cp(D.v_vav_b, s->v_vav_b); cp(D.u_base_b, s->u_base_b); cp(D.v_base_b, s->v_base_b);
  cp(D.tau_bx_b, s->tau_bx_b); cp(D.tau_by_b, s->tau_by_b); cp(D.eta_3D_b, s->eta_3D_b);
}
static void state_out(const DivaState& D, ufe_b200_velocity_state* s) {
  auto cp = [](double* dst, const std::vector<double>& src) { if (dst) std::memcpy(dst, src.data(), sizeof(double) * src.size()); };
  cp(s->u_vav_b, D.u_vav_b); cp(s->v_vav_b, D.v_vav_b); cp(s->u_base_b, D.u_base_b); cp(s->v_base_b, D.v_base_b);
  cp(s->tau_bx_b, D.tau_bx_b); cp(s->tau_by_b, D.tau_by_b); cp(s->eta_3D_b, D.eta_3D_b);
  cp(s->u_b_prev, D.u_b_prev); cp(s->v_b_prev, D.v_b_prev); cp(s->tau_dx_b, D.tau_dx_b); cp(s->tau_dy_b, D.tau_dy_b);
  cp(s->u_3D_b, D.u_3D_b); cp(s->v_3D_b, D.v_3D_b);
  cp(s->du_dx_a, D.du_dx_a); cp(s->du_dy_a, D.du_dy_a); cp(s->dv_dx_a, D.dv_dx_a); cp(s->dv_dy_a, D.dv_dy_a);
  cp(s->du_dz_3D_a, D.du_dz_3D_a); cp(s->dv_dz_3D_a, D.dv_dz_3D_a); cp(s->eta_3D_a, D.eta_3D_a);
  cp(s->eta_vav_a, D.eta_vav_a); cp(s->N_a, D.N_a); cp(s->N_b, D.N_b); cp(s->dN_dx_b, D.dN_dx_b); cp(s->dN_dy_b, D.dN_dy_b);
  cp(s->F1_3D_a, D.F1_3D_a); cp(s->F2_3D_a, D.F2_3D_a); cp(s->F1_3D_b, D.F1_3D_b); cp(s->F2_3D_b, D.F2_3D_b);
  cp(s->basal_friction_coefficient_a, D.basal_friction_coefficient_a);
  cp(s->basal_friction_coefficient_b, D.basal_friction_coefficient_b); cp(s->beta_eff_a, D.beta_eff_a); cp(s->beta_eff_b, D.beta_eff_b);
}

// approximation: 0 SSA, 1 DIVA.  ksp: see orc_ksp_solve.  max_visc_its < 0: use cfg->visc_it_nit semantics of the reference
int orc_solve_velocity(void* h, const ufe_b200_config* cfg, int approximation, int ksp, const ufe_b200_ice_inputs* ice,
                       ufe_b200_velocity_state* st, int* n_visc_its, int* n_Axb_its, const int* BC_mask, const double* BC_u,
                       const double* BC_v, int max_visc_its) {
  auto* o = static_cast<OracleMesh*>(h);
  Config C = to_config(cfg);
  C.ksp_type = ksp;
  IceInputs in = to_ice(ice);
  state_in(o->D, st, o->mv.nV, o->mv.nTri, o->mv.nz);
  int rc = approximation == 1
               ? solve_DIVA(o->mv, o->ops, C, in, o->D, BC_mask, BC_u, BC_v, n_visc_its, n_Axb_its, max_visc_its)
               : solve_SSA(o->mv, o->ops, C, in, o->D, BC_mask, BC_u, BC_v, n_visc_its, n_Axb_its, max_visc_its);
  state_out(o->D, st);
  return rc;
}

int orc_calc_dHi_dt(void* h, const ufe_b200_config* cfg, int ksp, ufe_b200_thickness_io* io, double* dt, int* n_Axb_its) {
  auto* o = static_cast<OracleMesh*>(h);
  Config C = to_config(cfg);
  C.ksp_type = ksp;
  ThicknessIO t;
  t.Hi = io->Hi; t.Hb = io->Hb; t.SL = io->SL; t.u_vav_b = io->u_vav_b; t.v_vav_b = io->v_vav_b; t.SMB = io->SMB; t.BMB = io->BMB;
  t.LMB = io->LMB; t.fraction_margin = io->fraction_margin; t.mask_noice = io->mask_noice; t.dHi_dt_target = io->dHi_dt_target;
  t.BC_prescr_mask = io->BC_prescr_mask; t.BC_prescr_Hi = io->BC_prescr_Hi;
  t.AMB = io->AMB; t.dHi_dt = io->dHi_dt; t.Hi_tplusdt = io->Hi_tplusdt; t.divQ = io->divQ;
  return calc_dHi_dt(o->mv, C, t, dt, n_Axb_its);
}

void orc_flux_divergence_matrix(void* h, const double* u_b, const double* v_b, const double* fraction_margin, int* m, int* nnz) {
  auto* o = static_cast<OracleMesh*>(h);
  calc_flux_divergence_matrix_upwind(o->mv, u_b, v_b, fraction_margin, o->MdivQ);
  *m = o->MdivQ.m; *nnz = (int)o->MdivQ.nnz();
}
void orc_get_flux_divergence_matrix(void* h, int* ptr, int* ind, double* val) {
  const Csr& M = static_cast<OracleMesh*>(h)->MdivQ;
  std::memcpy(ptr, M.ptr.data(), sizeof(int) * M.ptr.size());
  std::memcpy(ind, M.ind.data(), sizeof(int) * M.ind.size());
  std::memcpy(val, M.val.data(), sizeof(double) * M.val.size());
}

double orc_Halfar_H(double A, double n, double H0, double R0, double x, double y, double t) { return Halfar_H(A, n, H0, R0, x, y, t); }
void orc_Schoof2006_icestream(double A, double n, double H, double dhdx, double L, double m, double y, double* u, double* tauc) {
  Schoof2006_icestream(A, n, H, dhdx, L, m, y, u, tauc);
}

}  // extern "C"

extern "C" void orc_solve_SIA(void* h, const ufe_b200_config* cfg, double SIA_maximum_diffusivity, const double* Hi, const double* Hs,
                              const double* A_flow, double* u_3D_b, double* v_3D_b, double* D_3D_b, double* du_dz_3D_a, double* dv_dz_3D_a) {
  auto* o = static_cast<OracleMesh*>(h);
  solve_SIA(o->mv, o->ops, to_config(cfg), SIA_maximum_diffusivity, Hi, Hs, A_flow, u_3D_b, v_3D_b, D_3D_b, du_dz_3D_a, dv_dz_3D_a);
}

extern "C" void orc_calc_secondary_velocities(void* h, const double* u_3D_b, const double* v_3D_b, const ufe_b200_secondary_velocities* out) {
  auto* o = static_cast<OracleMesh*>(h);
  static_assert(sizeof(SecondaryVelocities) == sizeof(ufe_b200_secondary_velocities), "the two structs list the same 21 pointers in the same order");
  SecondaryVelocities S;
  std::memcpy(&S, out, sizeof(S));
  calc_secondary_velocities(o->mv, o->ops, u_3D_b, v_3D_b, S);
}

extern "C" void orc_calc_vertical_velocities(void* h, const ufe_b200_config* cfg, const ufe_b200_vertical_velocity_io* io, const double* u_3D_b,
                                             const double* v_3D_b, const double* u_3D, const double* v_3D) {
  auto* o = static_cast<OracleMesh*>(h);
  static_assert(sizeof(VerticalVelocityIO) == sizeof(ufe_b200_vertical_velocity_io), "same 10 pointers in the same order");
  VerticalVelocityIO V;
  std::memcpy(&V, io, sizeof(V));
  calc_vertical_velocities(o->mv, o->ops, to_config(cfg), V, u_3D_b, v_3D_b, u_3D, v_3D);
}

extern "C" void orc_determine_masks(void* h, const ufe_b200_config* cfg, const double* Hi, const double* Hb, const double* SL, const ufe_b200_masks* out) {
  auto* o = static_cast<OracleMesh*>(h);
  static_assert(sizeof(MaskSet) == sizeof(ufe_b200_masks), "same 11 pointers in the same order");
  MaskSet M;
  std::memcpy(&M, out, sizeof(M));
  determine_masks(o->mv, to_config(cfg), Hi, Hb, SL, M);
}
extern "C" void orc_calc_effective_thickness(void* h, const ufe_b200_config* cfg, const double* Hi, const double* Hb, const double* SL, double* Hi_eff,
                                             double* fraction_margin) {
  auto* o = static_cast<OracleMesh*>(h);
  calc_effective_thickness(o->mv, to_config(cfg), Hi, Hb, SL, Hi_eff, fraction_margin);
}
extern "C" int orc_calc_grounded_fractions(void* h, const ufe_b200_config* cfg, int choice, int nbins, const double* Hi, const double* Hb, const double* SL,
                                           const double* dHb, const int* mask_floating_ice, const double* cdf, const double* cdf_b, double* fr,
                                           double* fr_b) {
  auto* o = static_cast<OracleMesh*>(h);
  try {
    calc_grounded_fractions(o->mv, o->ops, to_config(cfg), choice, nbins, Hi, Hb, SL, dHb, mask_floating_ice, cdf, cdf_b, fr, fr_b);
  } catch (const std::exception&) { return 1; }
  return 0;
}
extern "C" double orc_calc_critical_timestep_adv(void* h, const ufe_b200_config* cfg, int grounded_only, const double* Hi, const int* mask_floating_ice,
                                                 const double* u_vav_b, const double* v_vav_b) {
  auto* o = static_cast<OracleMesh*>(h);
  return calc_critical_timestep_adv(o->mv, to_config(cfg), grounded_only, Hi, mask_floating_ice, u_vav_b, v_vav_b);
}
extern "C" double orc_calc_critical_timestep_SIA(void* h, const ufe_b200_config* cfg, const double* Hi, const double* D_3D_b) {
  auto* o = static_cast<OracleMesh*>(h);
  return calc_critical_timestep_SIA(o->mv, to_config(cfg), Hi, D_3D_b);
}

extern "C" int orc_alter_ice_thickness(void* h, const ufe_b200_config* cfg, const ufe_b200_alter_params* prm, double time, const ufe_b200_masks* masks,
                                       const double* Hi_old, const double* Hb, const double* SL, const double* refgeo_Hi, const double* refgeo_Hb,
                                       const double* Ti_hom, double* Hi_new) {
  auto* o = static_cast<OracleMesh*>(h);
  MaskSet M; std::memcpy(&M, masks, sizeof(M));
  AlterParams P;
  P.do_protect_grounded_mask = prm->do_protect_grounded_mask; P.protect_grounded_mask_t_end = prm->protect_grounded_mask_t_end; P.Hi_min = prm->Hi_min;
  P.calving_law_threshold_thickness = prm->calving_law_threshold_thickness; P.calving_threshold_thickness_shelf = prm->calving_threshold_thickness_shelf;
  P.remove_ice_absent_at_PD = prm->remove_ice_absent_at_PD; P.do_remove_shelves = prm->do_remove_shelves;
  P.remove_shelves_larger_than_PD = prm->remove_shelves_larger_than_PD; P.continental_shelf_calving = prm->continental_shelf_calving;
  P.continental_shelf_min_height = prm->continental_shelf_min_height;
  P.fixiness_t_start = prm->fixiness_t_start; P.fixiness_t_end = prm->fixiness_t_end; P.do_fixiness_before_start = prm->do_fixiness_before_start;
  P.limitness_t_start = prm->limitness_t_start; P.limitness_t_end = prm->limitness_t_end; P.do_limitness_before_start = prm->do_limitness_before_start;
  P.modiness_H_style = prm->modiness_H_style; P.modiness_T_hom_ref = prm->modiness_T_hom_ref;
  P.fixiness_H_gl_gr = prm->fixiness_H_gl_gr; P.fixiness_H_gl_fl = prm->fixiness_H_gl_fl; P.fixiness_H_grounded = prm->fixiness_H_grounded;
  P.fixiness_H_floating = prm->fixiness_H_floating; P.fixiness_H_freeland = prm->fixiness_H_freeland; P.fixiness_H_freeocean = prm->fixiness_H_freeocean;
  P.limitness_H_gl_gr = prm->limitness_H_gl_gr; P.limitness_H_gl_fl = prm->limitness_H_gl_fl; P.limitness_H_grounded = prm->limitness_H_grounded;
  P.limitness_H_floating = prm->limitness_H_floating;
  try {
    alter_ice_thickness(o->mv, to_config(cfg), P, time, M, Hi_old, Hb, SL, refgeo_Hi, refgeo_Hb, Ti_hom, Hi_new);
  } catch (const std::exception&) { return 1; }
  return 0;
}
extern "C" void orc_calc_and_apply_spill_over_flux(void* h, const ufe_b200_masks* masks, const double* Hi_eff, const double* u_vav_b, const double* v_vav_b,
                                                   double dt, double* Hi_new, double* Qspill) {
  auto* o = static_cast<OracleMesh*>(h);
  MaskSet M; std::memcpy(&M, masks, sizeof(M));
  calc_and_apply_spill_over_flux(o->mv, M, Hi_eff, u_vav_b, v_vav_b, dt, Hi_new, Qspill);
}

extern "C" int orc_set_ice_velocities(void* h, int approximation, int scheme, const double* u_3D_b, const double* v_3D_b, const double* u_b,
                                      const double* v_b, const double* du_dx_a, const double* du_dy_a, const double* dv_dx_a,
                                      const double* dv_dy_a, const double* du_dz_3D, const double* dv_dz_3D, const ufe_b200_ice_velocities* out) {
  auto* o = static_cast<OracleMesh*>(h);
  SolverResults R;
  R.u_3D_b = u_3D_b; R.v_3D_b = v_3D_b; R.u_b = u_b; R.v_b = v_b; R.du_dx_a = du_dx_a; R.du_dy_a = du_dy_a; R.dv_dx_a = dv_dx_a; R.dv_dy_a = dv_dy_a;
  R.du_dz_3D = du_dz_3D; R.dv_dz_3D = dv_dz_3D;
  IceVelocities I{out->u_3D_b, out->v_3D_b, out->du_dx_3D, out->du_dy_3D, out->du_dz_3D, out->dv_dx_3D, out->dv_dy_3D, out->dv_dz_3D, out->dw_dx_3D, out->dw_dy_3D};
  return set_ice_velocities(o->mv, approximation, scheme, R, I);
}

extern "C" void orc_calc_pc_truncation_error(int nV, double zeta_t, double dt_n, double pc_eta_min, double pc_epsilon, const double* Hi_np1,
                                             const double* Hi_star_np1, const int* mask_grounded_ice, const int* mask_gl_gr,
                                             const double* fraction_gr, double* tau_np1, int* tau_n_guilty, double* eta_np1, int* n_tot, int* n_guilty) {
  calc_pc_truncation_error(nV, zeta_t, dt_n, pc_eta_min, pc_epsilon, Hi_np1, Hi_star_np1, mask_grounded_ice, mask_gl_gr, fraction_gr, tau_np1,
                           tau_n_guilty, eta_np1, n_tot, n_guilty);
}

extern "C" void orc_apply_grounded_fractions_to_bed_roughness(int nV, const double* Hi, const double* Hs_slope, const double* fraction_gr,
                                                              const ufe_b200_masks* mk, const double* bed_roughness, double* applied) {
  apply_grounded_fractions_to_bed_roughness(nV, Hi, Hs_slope, fraction_gr, mk->mask_gl_gr, mk->mask_cf_gr, mk->mask_gl_fl, mk->mask_grounded_ice,
                                            mk->mask_floating_ice, mk->mask_icefree_ocean, bed_roughness, applied);
}
