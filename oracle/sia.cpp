// TEST INFRASTRUCTURE ONLY (see oracle.hpp).
// solve_SIA (src/UFEMISM/ice_dynamics/conservation_of_momentum/SIA/SIA_main.f90:65-158): the analytical solution of the shallow ice
// approximation with Glen's flow law (Bueler and Brown, 2009, Eqs. 12-13).
#include <algorithm>
#include <cmath>
#include <vector>

#include "oracle.hpp"

namespace orc {

void solve_SIA(const MeshView& mesh, const Operators& ops, const Config& C, double SIA_maximum_diffusivity, const double* Hi, const double* Hs,
               const double* A_flow_3D, double* u_3D_b, double* v_3D_b, double* D_3D_b, double* du_dz_3D_a, double* dv_dz_3D_a) {
  const int nV = mesh.nV, nTri = mesh.nTri, nz = mesh.nz;
  const double n = C.Glens_flow_law_exponent;
  const double rg_n = std::pow(C.ice_density * C.grav, n);
  // b-grid fields (SIA_main.f90:101-108): map_a_b_2D, ddx_a_a_2D, ddy_a_a_2D, ddx_a_b_2D, ddy_a_b_2D, map_a_b_3D
  std::vector<double> Hi_b(nTri), Hs_b(nTri), dHs_dx(nV), dHs_dy(nV), dHs_dx_b(nTri), dHs_dy_b(nTri), A_flow_b((size_t)nTri * nz);
  spmv(ops.M[OP_map_a_b], Hi, Hi_b.data());
  spmv(ops.M[OP_map_a_b], Hs, Hs_b.data());
  spmv(ops.M[OP_ddx_a_a], Hs, dHs_dx.data());
  spmv(ops.M[OP_ddy_a_a], Hs, dHs_dy.data());
  spmv(ops.M[OP_ddx_a_b], Hs, dHs_dx_b.data());
  spmv(ops.M[OP_ddy_a_b], Hs, dHs_dy_b.data());
  for (int k = 0; k < nz; ++k) spmv(ops.M[OP_map_a_b], A_flow_3D + (size_t)k * nV, A_flow_b.data() + (size_t)k * nTri);
  std::vector<double> z(nz), f(nz), integral(nz);
  for (int ti = 0; ti < nTri; ++ti) {   // :119-140
    for (int k = 0; k < nz; ++k) {
      z[k] = Hs_b[ti] - mesh.zeta[k] * Hi_b[ti];
      f[k] = A_flow_b[ti + (size_t)k * nTri] * std::pow(Hs_b[ti] - z[k], n);
    }
    // integrate_from_zeta_is_one_to_zeta_is_zetap( z, f) (mesh_zeta.f90:175-206), called with z in the place of zeta
    integral[nz - 1] = 0.0;
    for (int k = nz - 2; k >= 0; --k) integral[k] = integral[k + 1] - 0.5 * (f[k + 1] + f[k]) * (z[k + 1] - z[k]);
    const double abs_grad_Hs = std::sqrt(dHs_dx_b[ti] * dHs_dx_b[ti] + dHs_dy_b[ti] * dHs_dy_b[ti]);
    for (int k = 0; k < nz; ++k) {
      double D = -2.0 * rg_n * std::pow(abs_grad_Hs, n - 1.0) * integral[k];
      D = std::max(-SIA_maximum_diffusivity, D);   // :132
      if (D_3D_b) D_3D_b[ti + (size_t)k * nTri] = D;
      u_3D_b[ti + (size_t)k * nTri] = D * dHs_dx_b[ti];
      v_3D_b[ti + (size_t)k * nTri] = D * dHs_dy_b[ti];
    }
  }
  if (du_dz_3D_a && dv_dz_3D_a)
    for (int vi = 0; vi < nV; ++vi) {   // :142-155
      const double abs_grad_Hs = std::sqrt(dHs_dx[vi] * dHs_dx[vi] + dHs_dy[vi] * dHs_dy[vi]);
      for (int k = 0; k < nz; ++k) {
        const double zk = Hs[vi] - mesh.zeta[k] * Hi[vi];
        const double c = -2.0 * rg_n * std::pow(abs_grad_Hs, n - 1.0) * A_flow_3D[vi + (size_t)k * nV] * std::pow(Hs[vi] - zk, n);
        du_dz_3D_a[vi + (size_t)k * nV] = c * dHs_dx[vi];
        dv_dz_3D_a[vi + (size_t)k * nV] = c * dHs_dy[vi];
      }
    }
}

}  // namespace orc
