// TEST INFRASTRUCTURE ONLY (see oracle.hpp).
// calc_secondary_velocities (src/UFEMISM/ice_dynamics/conservation_of_momentum/conservation_of_momentum_main.f90:177-246)
#include <algorithm>
#include <cmath>
#include <vector>

#include "oracle.hpp"

namespace orc {

// vertical_average (src/UPSY/mesh/mesh_zeta.f90:269-295)
static double vertical_average(const double* zeta, const double* f, int nz, size_t stride) {
  double s = 0.0;
  for (int k = 0; k < nz - 1; ++k) s = s + 0.5 * (f[(k + 1) * stride] + f[k * stride]) * (zeta[k + 1] - zeta[k]);
  return s;
}

void calc_secondary_velocities(const MeshView& mesh, const Operators& ops, const double* u_3D_b, const double* v_3D_b, SecondaryVelocities& S) {
  const int nV = mesh.nV, nTri = mesh.nTri, nz = mesh.nz;
  for (int ti = 0; ti < nTri; ++ti) {   // :192-213
    S.u_surf_b[ti] = u_3D_b[ti]; S.v_surf_b[ti] = v_3D_b[ti];
    S.uabs_surf_b[ti] = std::sqrt(S.u_surf_b[ti] * S.u_surf_b[ti] + S.v_surf_b[ti] * S.v_surf_b[ti]);
    S.u_base_b[ti] = u_3D_b[ti + (size_t)(nz - 1) * nTri]; S.v_base_b[ti] = v_3D_b[ti + (size_t)(nz - 1) * nTri];
    S.uabs_base_b[ti] = std::sqrt(S.u_base_b[ti] * S.u_base_b[ti] + S.v_base_b[ti] * S.v_base_b[ti]);
    S.u_vav_b[ti] = vertical_average(mesh.zeta, u_3D_b + ti, nz, nTri);
    S.v_vav_b[ti] = vertical_average(mesh.zeta, v_3D_b + ti, nz, nTri);
    S.uabs_vav_b[ti] = std::sqrt(S.u_vav_b[ti] * S.u_vav_b[ti] + S.v_vav_b[ti] * S.v_vav_b[ti]);
  }
  const Csr& M = ops.M[OP_map_b_a];   // :217-232
  for (int k = 0; k < nz; ++k) {
    spmv(M, u_3D_b + (size_t)k * nTri, S.u_3D + (size_t)k * nV);
    spmv(M, v_3D_b + (size_t)k * nTri, S.v_3D + (size_t)k * nV);
  }
  spmv(M, S.u_surf_b, S.u_surf); spmv(M, S.v_surf_b, S.v_surf);
  spmv(M, S.u_base_b, S.u_base); spmv(M, S.v_base_b, S.v_base);
  spmv(M, S.u_vav_b, S.u_vav); spmv(M, S.v_vav_b, S.v_vav);
  for (int vi = 0; vi < nV; ++vi) {   // :235-244
    S.uabs_surf[vi] = std::sqrt(S.u_surf[vi] * S.u_surf[vi] + S.v_surf[vi] * S.v_surf[vi]);
    S.uabs_base[vi] = std::sqrt(S.u_base[vi] * S.u_base[vi] + S.v_base[vi] * S.v_base[vi]);
    S.uabs_vav[vi] = std::sqrt(S.u_vav[vi] * S.u_vav[vi] + S.v_vav[vi] * S.v_vav[vi]);
    S.R_shear[vi] = (S.uabs_base[vi] + 0.1) / (S.uabs_surf[vi] + 0.1);
  }
}

}  // namespace orc

// ---- calc_vertical_velocities (src/UFEMISM/ice_dynamics/conservation_of_mass/vertical_velocities.f90:23-196) ------------------------
namespace orc {

// calc_shape_functions_1D_reg_2nd_order (src/UPSY/basic/math_utilities/shape_functions.f90:17-79), two neighbours, q = 1.5
static void shape_1D_reg_2nd(double x, const double xc[2], double* Nfx_i, double Nfx_c[2]) {
  const double q = 1.5;
  double dx[2], w[2], a11 = 0, a12 = 0, a21 = 0, a22 = 0;
  for (int ci = 0; ci < 2; ++ci) { dx[ci] = xc[ci] - x; w[ci] = 1.0 / std::pow(std::fabs(dx[ci]), q); }
  for (int ci = 0; ci < 2; ++ci) {
    const double w2 = w[ci] * w[ci];
    a11 += w2 * dx[ci] * dx[ci];
    a12 += w2 * dx[ci] * 1 / 2 * (dx[ci] * dx[ci]);
    a21 += w2 * 1 / 2 * (dx[ci] * dx[ci]) * dx[ci];
    a22 += w2 * 1 / 2 * (dx[ci] * dx[ci]) * 1 / 2 * (dx[ci] * dx[ci]);
  }
  const double det = a11 * a22 - a12 * a21;   // calc_matrix_inverse_2_by_2 (matrix_algebra.f90:110-131)
  const double m11 = a22 / det, m12 = -a12 / det;
  for (int ci = 0; ci < 2; ++ci) Nfx_c[ci] = w[ci] * w[ci] * ((m11 * dx[ci]) + (m12 * 1 / 2 * (dx[ci] * dx[ci])));
  *Nfx_i = -(Nfx_c[0] + Nfx_c[1]);
}

// M_ddzeta_k_k_1D (calc_vertical_operators_reg_1D, src/UPSY/mesh/mesh_zeta.f90:300-376): per layer the diagonal and two neighbours
void ddzeta_operator(const double* zeta, int nz, std::vector<int>& nb, std::vector<double>& coef) {
  nb.assign((size_t)nz * 2, 0); coef.assign((size_t)nz * 3, 0.0);
  for (int k = 0; k < nz; ++k) {
    int ic[2];
    if (k == 0) { ic[0] = 1; ic[1] = 2; }
    else if (k == nz - 1) { ic[0] = nz - 3; ic[1] = nz - 2; }
    else { ic[0] = k - 1; ic[1] = k + 1; }
    const double zc[2] = {zeta[ic[0]], zeta[ic[1]]};
    double ni, ncf[2];
    shape_1D_reg_2nd(zeta[k], zc, &ni, ncf);
    nb[2 * k] = ic[0]; nb[2 * k + 1] = ic[1];
    coef[3 * k] = ni; coef[3 * k + 1] = ncf[0]; coef[3 * k + 2] = ncf[1];
  }
}

void calc_vertical_velocities(const MeshView& mesh, const Operators& ops, const Config& C, const VerticalVelocityIO& io, const double* u_3D_b,
                              const double* v_3D_b, const double* u_3D, const double* v_3D) {
  const int nV = mesh.nV, nTri = mesh.nTri, nE = mesh.nE, nz = mesh.nz, ncm = mesh.nC_mem;
  std::vector<double> dHib_dt(nV), dHib_dx(nV), dHib_dy(nV), dHs_dx(nV), dHs_dy(nV), dHi_dx(nV), dHi_dy(nV);
  for (int vi = 0; vi < nV; ++vi) {   // :93-107
    if (io.mask_grounded_ice[vi]) dHib_dt[vi] = io.dHb_dt[vi];
    else if (io.mask_floating_ice[vi]) dHib_dt[vi] = -io.dHi_dt[vi] * C.ice_density / C.seawater_density;
    else dHib_dt[vi] = 0.0;
  }
  spmv(ops.M[OP_ddx_a_a], io.Hib, dHib_dx.data()); spmv(ops.M[OP_ddy_a_a], io.Hib, dHib_dy.data());   // :110-111
  // zeta gradients on the a-grid (calc_zeta_gradients, src/UFEMISM/ice_dynamics/utilities/zeta_gradients.f90:88-115)
  spmv(ops.M[OP_ddx_a_a], io.Hs, dHs_dx.data()); spmv(ops.M[OP_ddy_a_a], io.Hs, dHs_dy.data());
  spmv(ops.M[OP_ddx_a_a], io.Hi, dHi_dx.data()); spmv(ops.M[OP_ddy_a_a], io.Hi, dHi_dy.data());
  // map_velocities_from_b_to_c_3D (map_velocities_to_c_grid.f90:72-123)
  std::vector<double> u_c((size_t)nE * nz), v_c((size_t)nE * nz);
  for (int ei = 0; ei < nE; ++ei) {
    const int til = mesh.ETri[ei], tir = mesh.ETri[ei + nE];
    for (int k = 0; k < nz; ++k) {
      const size_t o = (size_t)k * nTri;
      if (til == 0 && tir > 0) { u_c[ei + (size_t)k * nE] = u_3D_b[tir - 1 + o]; v_c[ei + (size_t)k * nE] = v_3D_b[tir - 1 + o]; }
      else if (tir == 0 && til > 0) { u_c[ei + (size_t)k * nE] = u_3D_b[til - 1 + o]; v_c[ei + (size_t)k * nE] = v_3D_b[til - 1 + o]; }
      else {
        u_c[ei + (size_t)k * nE] = (u_3D_b[til - 1 + o] + u_3D_b[tir - 1 + o]) / 2.0;
        v_c[ei + (size_t)k * nE] = (v_3D_b[til - 1 + o] + v_3D_b[tir - 1 + o]) / 2.0;
      }
    }
  }
  auto W = [&](int vi, int k) -> double& { return io.w_3D[vi + (size_t)k * nV]; };
  for (int vi = 0; vi < nV; ++vi) {   // :118-188
    if (!(io.mask_grounded_ice[vi] || io.mask_floating_ice[vi])) { for (int k = 0; k < nz; ++k) W(vi, k) = 0.0; continue; }
    W(vi, nz - 1) = (u_3D[vi + (size_t)(nz - 1) * nV] * dHib_dx[vi]) + (v_3D[vi + (size_t)(nz - 1) * nV] * dHib_dy[vi]) + dHib_dt[vi] + std::min(0.0, io.BMB[vi]);
    if (io.Hi[vi] < 10.0) { for (int k = 0; k < nz; ++k) W(vi, k) = W(vi, nz - 1); continue; }
    const double Hi = std::max(10.0, io.Hi[vi]);
    for (int ks = nz - 2; ks >= 0; --ks) {
      const double dzeta = mesh.zeta[ks + 1] - mesh.zeta[ks];
      double cint_un_dS = 0.0;
      for (int ci = 0; ci < mesh.nC[vi]; ++ci) {
        const int vj = mesh.C[vi + (size_t)ci * nV] - 1, ei = mesh.VE[vi + (size_t)ci * nV] - 1;
        const double u_ks = 0.5 * (u_c[ei + (size_t)ks * nE] + u_c[ei + (size_t)(ks + 1) * nE]);
        const double v_ks = 0.5 * (v_c[ei + (size_t)ks * nE] + v_c[ei + (size_t)(ks + 1) * nE]);
        const double dS = mesh.Cw[vi + (size_t)ci * nV];
        double nx = mesh.V[vj] - mesh.V[vi], ny = mesh.V[vj + nV] - mesh.V[vi + nV];
        const double nn = std::sqrt(nx * nx + ny * ny);
        nx = nx / nn; ny = ny / nn;
        cint_un_dS = cint_un_dS + (u_ks * nx + v_ks * ny) * dS;
      }
      const double grad_uv_ks = cint_un_dS / mesh.A[vi];
      const double du_dzeta_ks = (u_3D[vi + (size_t)(ks + 1) * nV] - u_3D[vi + (size_t)ks * nV]) / dzeta;
      const double dv_dzeta_ks = (v_3D[vi + (size_t)(ks + 1) * nV] - v_3D[vi + (size_t)ks * nV]) / dzeta;
      auto dzdx = [&](int k) { return (1.0 / Hi) * (dHs_dx[vi] - mesh.zeta[k] * dHi_dx[vi]); };
      auto dzdy = [&](int k) { return (1.0 / Hi) * (dHs_dy[vi] - mesh.zeta[k] * dHi_dy[vi]); };
      const double dzeta_dx_ks = 0.5 * (dzdx(ks) + dzdx(ks + 1));
      const double dzeta_dy_ks = 0.5 * (dzdy(ks) + dzdy(ks + 1));
      const double dzeta_dz_ks = 0.5 * ((-1.0 / Hi) + (-1.0 / Hi));
      const double dw_dzeta_ks = -1.0 / dzeta_dz_ks * (grad_uv_ks + dzeta_dx_ks * du_dzeta_ks + dzeta_dy_ks * dv_dzeta_ks);
      W(vi, ks) = W(vi, ks + 1) - dzeta * dw_dzeta_ks;
    }
  }
  if (io.dw_dz_3D) {   // calc_dw_dz (:198-243)
    std::vector<int> nb; std::vector<double> coef;
    ddzeta_operator(mesh.zeta, nz, nb, coef);
    for (int vi = 0; vi < nV; ++vi)
      for (int k = 0; k < nz; ++k) {
        double& o = io.dw_dz_3D[vi + (size_t)k * nV];
        if (!(io.mask_grounded_ice[vi] || io.mask_floating_ice[vi])) { o = std::nan(""); continue; }
        const double d = coef[3 * k] * W(vi, k) + coef[3 * k + 1] * W(vi, nb[2 * k]) + coef[3 * k + 2] * W(vi, nb[2 * k + 1]);
        o = -1.0 / io.Hi[vi] * d;
      }
  }
  (void)ncm;
}

}  // namespace orc

// ---------------------------------------------------------------------------------------------------
// set_ice_velocities_to_SIA_results (conservation_of_momentum_main.f90:311-347), _SSA_ (:349-385), _SIASSA_ (:387-463), _DIVA_ (:465-511)
// ---------------------------------------------------------------------------------------------------
int orc::set_ice_velocities(const MeshView& mesh, int approximation, int scheme, const SolverResults& R, IceVelocities& I) {
  const int nV = mesh.nV, nTri = mesh.nTri, nz = mesh.nz;
  const double pi = 3.141592653589793238462643383279502884197;
  auto zero_a = [&](double* d) { if (d) std::fill(d, d + (size_t)nV * nz, 0.0); };
  auto layers = [&](double* d, const double* s2) { if (d) for (int k = 0; k < nz; ++k) for (int v = 0; v < nV; ++v) d[(size_t)k * nV + v] = s2[v]; };
  auto copy_a = [&](double* d, const double* s3) { if (d) std::copy(s3, s3 + (size_t)nV * nz, d); };
  switch (approximation) {
    case 2:  // SIA
      for (size_t j = 0; j < (size_t)nTri * nz; ++j) { I.u_3D_b[j] = R.u_3D_b[j]; I.v_3D_b[j] = R.v_3D_b[j]; }
      copy_a(I.du_dz_3D, R.du_dz_3D); copy_a(I.dv_dz_3D, R.dv_dz_3D);
      zero_a(I.du_dx_3D); zero_a(I.du_dy_3D); zero_a(I.dv_dx_3D); zero_a(I.dv_dy_3D);
      break;
    case 0:  // SSA
      for (int k = 0; k < nz; ++k) for (int t = 0; t < nTri; ++t) { I.u_3D_b[(size_t)k * nTri + t] = R.u_b[t]; I.v_3D_b[(size_t)k * nTri + t] = R.v_b[t]; }
      layers(I.du_dx_3D, R.du_dx_a); layers(I.du_dy_3D, R.du_dy_a); layers(I.dv_dx_3D, R.dv_dx_a); layers(I.dv_dy_3D, R.dv_dy_a);
      zero_a(I.du_dz_3D); zero_a(I.dv_dz_3D);
      break;
    case 3:  // SIA/SSA
      if (scheme != 0 && scheme != 1) return 1;  // crash('unknown choice_hybrid_SIASSA_scheme_config')
      for (int t = 0; t < nTri; ++t) {
        double w_sia_u = 1.0, w_sia_v = 1.0;
        if (scheme == 1) {
          w_sia_u = 1.0 - (2.0 / pi) * std::atan(std::pow(std::fabs(R.u_b[t]), 2.0) / std::pow(30.0, 2.0));
          w_sia_v = 1.0 - (2.0 / pi) * std::atan(std::pow(std::fabs(R.v_b[t]), 2.0) / std::pow(30.0, 2.0));
        }
        for (int k = 0; k < nz; ++k) {
          const size_t j = (size_t)k * nTri + t;
          I.u_3D_b[j] = w_sia_u * R.u_3D_b[j] + R.u_b[t];
          I.v_3D_b[j] = w_sia_v * R.v_3D_b[j] + R.v_b[t];
        }
      }
      copy_a(I.du_dz_3D, R.du_dz_3D); copy_a(I.dv_dz_3D, R.dv_dz_3D);
      layers(I.du_dx_3D, R.du_dx_a); layers(I.du_dy_3D, R.du_dy_a); layers(I.dv_dx_3D, R.dv_dx_a); layers(I.dv_dy_3D, R.dv_dy_a);
      break;
    case 1:  // DIVA
      for (size_t j = 0; j < (size_t)nTri * nz; ++j) { I.u_3D_b[j] = R.u_3D_b[j]; I.v_3D_b[j] = R.v_3D_b[j]; }
      layers(I.du_dx_3D, R.du_dx_a); layers(I.du_dy_3D, R.du_dy_a); layers(I.dv_dx_3D, R.dv_dx_a); layers(I.dv_dy_3D, R.dv_dy_a);
      copy_a(I.du_dz_3D, R.du_dz_3D); copy_a(I.dv_dz_3D, R.dv_dz_3D);
      break;
    default: return 2;  // crash('unknown choice_stress_balance_approximation')
  }
  zero_a(I.dw_dx_3D); zero_a(I.dw_dy_3D);
  return 0;
}
