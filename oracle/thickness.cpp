// TEST INFRASTRUCTURE ONLY (see oracle.hpp).
// CPU restatement of the ice-thickness step (paths relative to src/UFEMISM/ice_dynamics):
//   calc_dHi_dt                              conservation_of_mass/conservation_of_mass_main.f90:27-119
//   calc_dHi_dt_explicit + BC                conservation_of_mass/conservation_of_mass_explicit.f90:24-297
//   calc_dHi_dt_semiimplicit + BC            conservation_of_mass/conservation_of_mass_semiimplicit.f90:25-322
//   calc_ice_flux_divergence_matrix_upwind, calc_flux_limited_timestep, calc_n_interior_neighbours
//                                            conservation_of_mass/conservation_of_mass_utilities.f90:23-239
//   map_velocities_from_b_to_c_2D            utilities/map_velocities_to_c_grid.f90:20-77
//   ice_surface_elevation, Hi_from_Hb_Hs_and_SL   src/UPSY/basic/math_utilities/ice_geometry_basics.f90:33-94
//   SIA                                      conservation_of_momentum/SIA/SIA_main.f90:65-158
// calc_and_apply_spill_over_flux (ice_thickness_safeties.f90:290) is a "next" row: it is a
// no-op when there are no calving-front cells, which is how every config here is set up.
#include <algorithm>
#include <cmath>

#include "oracle.hpp"

namespace orc {

namespace {
double ice_surface_elevation(const Config& C, double Hi, double Hb, double SL) {
  return Hi + std::max(SL - C.ice_density / C.seawater_density * Hi, Hb);
}
double Hi_from_Hb_Hs_and_SL(const Config& C, double Hb, double Hs, double SL) {
  const double Hi_float = std::max(0.0, (SL - Hb) * (C.seawater_density / C.ice_density));
  const double Hs_float = Hb + Hi_float;
  if (Hs > Hs_float) return Hs - Hb;
  return std::min(Hi_float, (Hs - SL) / (1.0 - (C.ice_density / C.seawater_density)));
}
int side_of(int bi) { return (bi == 1 || bi == 2) ? 0 : (bi == 3 || bi == 4) ? 1 : (bi == 5 || bi == 6) ? 2 : 3; }
}  // namespace

void calc_flux_divergence_matrix_upwind(const MeshView& mesh, const double* u_b, const double* v_b,
                                        const double* fraction_margin, Csr& M, std::vector<double>* u_perp_out) {
  const int nV = mesh.nV, nE = mesh.nE, ncm = mesh.nC_mem;
  std::vector<double> u_c(nE), v_c(nE);
  for (int ei = 1; ei <= nE; ++ei) {
    const int til = mesh.ETri[ei - 1], tir = mesh.ETri[(size_t)nE + ei - 1];
    if (til == 0 && tir > 0) { u_c[ei - 1] = u_b[tir - 1]; v_c[ei - 1] = v_b[tir - 1]; }
    else if (tir == 0 && til > 0) { u_c[ei - 1] = u_b[til - 1]; v_c[ei - 1] = v_b[til - 1]; }
    else { u_c[ei - 1] = (u_b[til - 1] + u_b[tir - 1]) / 2.0; v_c[ei - 1] = (v_b[til - 1] + v_b[tir - 1]) / 2.0; }
  }
  long nnz_est = nV;
  for (int v = 0; v < nV; ++v) nnz_est += mesh.nC[v];
  M.init(nV, nV, nnz_est);
  if (u_perp_out) u_perp_out->assign((size_t)nV * ncm, 0.0);
  std::vector<double> cM(ncm + 1);
  for (int vi = 1; vi <= nV; ++vi) {
    std::fill(cM.begin(), cM.end(), 0.0);
    const double A_i = mesh.A[vi - 1];
    for (int ci = 1; ci <= mesh.nC[vi - 1]; ++ci) {
      const size_t k = (size_t)(ci - 1) * nV + vi - 1;
      const int ei = mesh.VE[k], vj = mesh.C[k];
      const double L_c = mesh.Cw[k];
      const double u_perp = u_c[ei - 1] * mesh.D_x[k] / mesh.D[k] + v_c[ei - 1] * mesh.D_y[k] / mesh.D[k];
      if (u_perp_out) (*u_perp_out)[k] = u_perp;
      if (fraction_margin[vi - 1] >= 1.0) cM[0] = cM[0] + L_c * std::max(0.0, u_perp) / A_i;
      if (fraction_margin[vj - 1] >= 1.0) cM[ci] = L_c * std::min(0.0, u_perp) / A_i;
    }
    M.add(vi, vi, cM[0]);
    for (int ci = 1; ci <= mesh.nC[vi - 1]; ++ci) M.add(vi, mesh.C[(size_t)(ci - 1) * nV + vi - 1], cM[ci]);
  }
  M.finalise();
}

namespace {

// apply_ice_thickness_BC_explicit (conservation_of_mass_explicit.f90:159-297)
void apply_BC_explicit(const MeshView& mesh, const Config& C, const int* mask_noice, const double* Hb, const double* SL, double* Hi_new) {
  const int nV = mesh.nV;
  std::vector<double> Hs(nV), Hs_tot(nV);
  std::vector<int> n_int(nV, 0);
  for (int v = 0; v < nV; ++v) Hs[v] = ice_surface_elevation(C, Hi_new[v], Hb[v], SL[v]);
  Hs_tot = Hs;
  for (int vi = 1; vi <= nV; ++vi)
    for (int ci = 1; ci <= mesh.nC[vi - 1]; ++ci) {
      int vj = mesh.C[(size_t)(ci - 1) * nV + vi - 1];
      if (mesh.VBI[vj - 1] == 0 && !mask_noice[vj - 1]) n_int[vi - 1]++;
    }
  for (int pass = 1; pass <= 2; ++pass) {
    for (int vi = 1; vi <= nV; ++vi) {
      const int bi = mesh.VBI[vi - 1];
      if (bi == 0) continue;
      if (C.BC_H[side_of(bi)] == BC_zero) { Hi_new[vi - 1] = 0.0; continue; }
      const bool has_int = n_int[vi - 1] > 0;
      if ((pass == 1) != has_int) continue;
      double Hs_sum = 0.0;
      for (int ci = 1; ci <= mesh.nC[vi - 1]; ++ci) {
        int vj = mesh.C[(size_t)(ci - 1) * nV + vi - 1];
        if (pass == 2 || (mesh.VBI[vj - 1] == 0 && !mask_noice[vj - 1])) Hs_sum += Hs_tot[vj - 1];
      }
      const double Hs_av = Hs_sum / (double)(pass == 1 ? n_int[vi - 1] : mesh.nC[vi - 1]);
      Hs[vi - 1] = std::max(Hb[vi - 1], Hs_av);
      Hi_new[vi - 1] = Hi_from_Hb_Hs_and_SL(C, Hb[vi - 1], Hs[vi - 1], SL[vi - 1]);
    }
    Hs_tot = Hs;  // second gather_to_all (:232)
  }
}

struct ExplicitOut { std::vector<double> AMB, dHi_dt, Hi_new, divQ; double dt; };

void dHi_dt_explicit(const MeshView& mesh, const Config& C, const ThicknessIO& io, double dt_in, double dt_ice_max, ExplicitOut& o) {
  const int nV = mesh.nV;
  Csr M;
  calc_flux_divergence_matrix_upwind(mesh, io.u_vav_b, io.v_vav_b, io.fraction_margin, M);
  o.divQ.resize(nV); o.dHi_dt.resize(nV); o.AMB.resize(nV); o.Hi_new.resize(nV);
  spmv(M, io.Hi, o.divQ.data());
  for (int v = 0; v < nV; ++v) {
    o.dHi_dt[v] = -o.divQ[v] + io.fraction_margin[v] * (io.SMB[v] + io.BMB[v] - io.dHi_dt_target[v]) + io.LMB[v];
    o.AMB[v] = o.dHi_dt[v];
  }
  // calc_flux_limited_timestep (conservation_of_mass_utilities.f90:160-205), restated verbatim
  double dt_max = dt_ice_max;
  for (int v = 0; v < nV; ++v)
    if (io.Hi[v] > C.Hi_min && o.dHi_dt[v] < 0.0) dt_max = std::min(dt_max, io.Hi[v] / std::max(o.dHi_dt[v], 1e-9));
  dt_max = std::max(C.dt_ice_min, dt_max);
  o.dt = std::min(dt_in, dt_max);
  for (int v = 0; v < nV; ++v) o.Hi_new[v] = std::max(0.0, io.Hi[v] + o.dHi_dt[v] * o.dt);
  apply_BC_explicit(mesh, C, io.mask_noice, io.Hb, io.SL, o.Hi_new.data());
  if (io.BC_prescr_mask)
    for (int v = 0; v < nV; ++v) if (io.BC_prescr_mask[v] == 1) o.Hi_new[v] = std::max(0.0, io.BC_prescr_Hi[v]);
  for (int v = 0; v < nV; ++v) if (io.mask_noice[v]) o.Hi_new[v] = 0.0;
  for (int v = 0; v < nV; ++v) {
    o.dHi_dt[v] = (o.Hi_new[v] - io.Hi[v]) / o.dt;
    o.AMB[v] = o.dHi_dt[v] - o.AMB[v];
  }
}

void row_to_identity(Csr& A, int i) {  // set_diagonal_to_one_and_rest_of_row_to_zero (CSR_matrix_basics.f90:574-)
  for (int k = A.ptr[i - 1]; k <= A.ptr[i] - 1; ++k) A.val[k - 1] = (A.ind[k - 1] == i) ? 1.0 : 0.0;
}

}  // namespace

int calc_dHi_dt(const MeshView& mesh, const Config& C, ThicknessIO& io, double* dt, int* n_Axb_its) {
  const int nV = mesh.nV;
  const double dt_ice_max = C.dt_ice_max;  // model_configuration_type_and_namelist.f90:391
  *n_Axb_its = 0;
  for (int v = 0; v < nV; ++v) io.AMB[v] = 0.0;
  ExplicitOut ex;
  dHi_dt_explicit(mesh, C, io, *dt, dt_ice_max, ex);
  if (C.choice_ice_integration_method == 0) {
    *dt = ex.dt;
    for (int v = 0; v < nV; ++v) { io.AMB[v] = ex.AMB[v]; io.dHi_dt[v] = ex.dHi_dt[v]; io.Hi_tplusdt[v] = ex.Hi_new[v]; io.divQ[v] = ex.divQ[v]; }
  } else {
    // calc_dHi_dt_semiimplicit (conservation_of_mass_semiimplicit.f90:104-181): note dt itself is NOT reduced to dt_ex
    const double dts = *dt, fs = C.dHi_semiimplicit_fs;
    Csr M, AA;
    calc_flux_divergence_matrix_upwind(mesh, io.u_vav_b, io.v_vav_b, io.fraction_margin, M);
    spmv(M, io.Hi, io.divQ);
    AA = M;
    for (auto& v : AA.val) v = v * dts * fs;
    for (int vi = 1; vi <= nV; ++vi)
      for (int k = AA.ptr[vi - 1]; k <= AA.ptr[vi] - 1; ++k) if (M.ind[k - 1] == vi) AA.val[k - 1] += 1.0;
    std::vector<double> bb(nV);
    for (int v = 0; v < nV; ++v)
      bb[v] = io.Hi[v] - (dts * (1.0 - fs) * io.divQ[v]) +
              std::max(-1.0 * io.Hi[v], dts * (io.fraction_margin[v] * (io.SMB[v] + io.BMB[v] - io.dHi_dt_target[v]) + io.LMB[v]));
    // BCs: domain border (uses the explicit solution, BC re-applied), prescribed, no-ice (:184-322)
    apply_BC_explicit(mesh, C, io.mask_noice, io.Hb, io.SL, ex.Hi_new.data());
    for (int vi = 1; vi <= nV; ++vi) if (mesh.VBI[vi - 1] > 0) { row_to_identity(AA, vi); bb[vi - 1] = ex.Hi_new[vi - 1]; }
    if (io.BC_prescr_mask)
      for (int vi = 1; vi <= nV; ++vi) if (io.BC_prescr_mask[vi - 1] == 1) { row_to_identity(AA, vi); bb[vi - 1] = io.BC_prescr_Hi[vi - 1]; }
    for (int vi = 1; vi <= nV; ++vi) if (io.mask_noice[vi - 1]) { row_to_identity(AA, vi); bb[vi - 1] = 0.0; }
    const bool g = C.ksp_initial_guess_nonzero != 0;
    if (g) for (int v = 0; v < nV; ++v) io.Hi_tplusdt[v] = ex.Hi_new[v];
    int rc = ksp_solve(C.ksp_type, AA, bb.data(), io.Hi_tplusdt, C.dHi_PETSc_rtol, C.dHi_PETSc_abstol, C.ksp_maxit, n_Axb_its, g);
    if (rc != 0) return rc;
    for (int v = 0; v < nV; ++v) { io.dHi_dt[v] = (io.Hi_tplusdt[v] - io.Hi[v]) / dts; io.AMB[v] = io.dHi_dt[v] - (io.Hi_tplusdt[v] - io.Hi[v]) / dts; }
  }
  // tail of calc_dHi_dt (conservation_of_mass_main.f90:83-109); spill-over flux: see file header
  for (int v = 0; v < nV; ++v) {
    if (io.Hi_tplusdt[v] < 0.0) io.Hi_tplusdt[v] = 0.0;
    io.AMB[v] = io.AMB[v] + (io.Hi_tplusdt[v] - io.Hi[v]) / *dt - io.dHi_dt[v];
    io.dHi_dt[v] = (io.Hi_tplusdt[v] - io.Hi[v]) / *dt;
  }
  return 0;
}

}  // namespace orc

// ---------------------------------------------------------------------------------------------------
// calc_pc_truncation_error (predictor_corrector_scheme.f90:327-361; Robinson et al., 2020, Eq. 32) and the count of unstable
// vertices that run_ice_dynamics_model_pc takes right after it (:254-279)
// ---------------------------------------------------------------------------------------------------
void orc::calc_pc_truncation_error(int nV, double zeta_t, double dt_n, double pc_eta_min, double pc_epsilon, const double* Hi_np1,
                                   const double* Hi_star_np1, const int* mask_grounded_ice, const int* mask_gl_gr, const double* fraction_gr,
                                   double* tau_np1, int* tau_n_guilty, double* eta_np1, int* n_tot, int* n_guilty) {
  for (int vi = 0; vi < nV; ++vi)
    tau_np1[vi] = zeta_t * std::fabs(Hi_np1[vi] - Hi_star_np1[vi]) / ((3.0 * zeta_t + 3.0) * dt_n);
  double eta = pc_eta_min;
  for (int vi = 0; vi < nV; ++vi)
    if (mask_grounded_ice[vi] && !mask_gl_gr[vi] && fraction_gr[vi] == 1.0) eta = std::max(eta, tau_np1[vi]);
  *eta_np1 = eta;
  int tot = 0, guilty = 0;
  for (int vi = 0; vi < nV; ++vi) {
    if (fraction_gr[vi] < 1.0) continue;      // only fully grounded vertices
    ++tot;
    if (tau_np1[vi] > pc_epsilon) { ++guilty; if (tau_n_guilty) tau_n_guilty[vi] = tau_n_guilty[vi] + 1; }
  }
  *n_tot = tot; *n_guilty = guilty;
}
