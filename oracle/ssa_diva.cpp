// TEST INFRASTRUCTURE ONLY (see oracle.hpp).
// CPU restatement of the SSA / DIVA velocity solve:
//   solve_SSA_DIVA_linearised + row builders   SSA_DIVA/solve_linearised_SSA_DIVA_infinite_slab.f90:23-641
//   solve_DIVA_infinite_slab + helpers         SSA_DIVA/DIVA_solver_infinite_slab.f90:52-568
//   solve_SSA + helpers                        SSA_DIVA/SSA_main.f90:88-245, 315-460
//   driving stress, strain rates, relax, limits, L2 norm   SSA_DIVA/SSA_DIVA_utilities.f90:24-206
//   Glen viscosity                             rheology/constitutive_equation.f90:25-80
//   sliding laws                               conservation_of_momentum/sliding_laws.f90:25-695
//   zeta integrals                             src/UPSY/mesh/mesh_zeta.f90:175-295
// (paths relative to src/UFEMISM/ice_dynamics/conservation_of_momentum unless noted)
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include <omp.h>

#include "oracle.hpp"

namespace orc {

void DivaState::init(int nV, int nTri, int nz) {
  auto z = [](std::vector<double>& v, size_t n) { v.assign(n, 0.0); };
  for (auto* v : {&u_vav_b, &v_vav_b, &u_base_b, &v_base_b, &u_b_prev, &v_b_prev, &tau_dx_b, &tau_dy_b, &tau_bx_b, &tau_by_b,
                  &N_b, &dN_dx_b, &dN_dy_b, &basal_friction_coefficient_b, &beta_eff_b}) z(*v, nTri);
  for (auto* v : {&du_dx_a, &du_dy_a, &dv_dx_a, &dv_dy_a, &eta_vav_a, &N_a, &basal_friction_coefficient_a, &beta_eff_a}) z(*v, nV);
  for (auto* v : {&du_dz_3D_a, &dv_dz_3D_a, &eta_3D_a, &F1_3D_a, &F2_3D_a}) z(*v, (size_t)nV * nz);
  for (auto* v : {&eta_3D_b, &F1_3D_b, &F2_3D_b, &u_3D_b, &v_3D_b}) z(*v, (size_t)nTri * nz);
  // allocate_DIVA_solver zero-fills everything (DIVA_main.f90:298-350); 'zero' initial velocities (:66-68)
}

// ---------------------------------------------------------------------------------------------------
// Copy rows of the ISMIP-HOM / SSA_icestream experiments
// ---------------------------------------------------------------------------------------------------
// find_containing_vertex (mesh_utilities.f90:1368-1416): greedy walk to the vertex nearest to p
int find_containing_vertex(const MeshView& mesh, const double p[2], int vi) {
  const int nV = mesh.nV;
  auto dist = [&](int v) { const double dx = mesh.V[v - 1] - p[0], dy = mesh.V[(size_t)nV + v - 1] - p[1]; return std::sqrt(dx * dx + dy * dy); };
  int vi_prev = vi;
  for (;;) {
    const double d = dist(vi);
    double dcmin = d + 10.0; int vcmin = 0;
    for (int ci = 1; ci <= mesh.nC[vi - 1]; ++ci) {
      const int vc = mesh.C[(size_t)(ci - 1) * nV + vi - 1];
      if (vc == vi_prev) continue;  // the neighbour we just came from
      const double dc = dist(vc);
      if (dc < dcmin) { dcmin = dc; vcmin = vc; }
    }
    if (dcmin < d) { vi_prev = vi; vi = vcmin; } else return vi;
  }
}

// find_ti_copy_ISMIP_HOM_periodic (mesh_utilities.f90:2620-2677) and find_ti_copy_SSA_icestream_infinite (:2681-2730):
// the point the previous solution is copied from, the vertex whose Voronoi cell holds it, inverse-square-distance weights
// over that vertex's triangles.  ti_copy / wti_copy hold nC_mem entries (0 = unused).
void find_ti_copy(const MeshView& mesh, int choice, double L, int ti, int* ti_copy, double* wti_copy) {
  const int nV = mesh.nV, nTri = mesh.nTri;
  const double gc[2] = {mesh.TriGC[ti - 1], mesh.TriGC[(size_t)nTri + ti - 1]};
  double p[2];
  if (choice == BC_periodic_ISMIP_HOM) {
    p[0] = (gc[0] > 0.0) ? gc[0] - L / 2.0 : gc[0] + L / 2.0;
    p[1] = (gc[1] > 0.0) ? gc[1] - L / 2.0 : gc[1] + L / 2.0;
  } else {
    p[1] = gc[1];
    p[0] = (gc[0] < 0.0) ? mesh.xmin + (mesh.xmax - mesh.xmin) * 1.0 / 3.0 : mesh.xmin + (mesh.xmax - mesh.xmin) * 2.0 / 3.0;
  }
  const int vi = find_containing_vertex(mesh, p, 5);
  for (int n = 0; n < mesh.nC_mem; ++n) { ti_copy[n] = 0; wti_copy[n] = 0.0; }
  const int nt = mesh.niTri[vi - 1];
  double wsum = 0.0;
  for (int iti = 1; iti <= nt; ++iti) {
    const int tj = mesh.iTri[(size_t)(iti - 1) * nV + vi - 1];
    const double dx = p[0] - mesh.TriGC[tj - 1], dy = p[1] - mesh.TriGC[(size_t)nTri + tj - 1];
    const double dist = std::sqrt(dx * dx + dy * dy);
    ti_copy[iti - 1] = tj;
    wti_copy[iti - 1] = 1.0 / (dist * dist);
    wsum += wti_copy[iti - 1];
  }
  for (int iti = 0; iti < nt; ++iti) wti_copy[iti] = wti_copy[iti] / wsum;
}

// ---------------------------------------------------------------------------------------------------
// Assembly of the linearised SSA/DIVA (solve_linearised_SSA_DIVA_infinite_slab.f90:61-153)
// ---------------------------------------------------------------------------------------------------
void assemble_SSA_DIVA(const MeshView& mesh, const Operators& ops, const Config& C,
                       const double* N_b, const double* dN_dx_b, const double* dN_dy_b, const double* beta_b,
                       const double* tau_dx_b, const double* tau_dy_b,
                       const int* BC_mask, const double* BC_u, const double* BC_v,
                       Csr& A, std::vector<double>& bb, const double* u_b_prev, const double* v_b_prev) {
  const int nTri = mesh.nTri, R = 2 * nTri;
  const Csr& Mx = ops.M[OP2_ddx_b_b]; const Csr& My = ops.M[OP2_ddy_b_b];
  const Csr& Mxx = ops.M[OP2_d2dx2_b_b]; const Csr& Mxy = ops.M[OP2_d2dxdy_b_b]; const Csr& Myy = ops.M[OP2_d2dy2_b_b];
  bb.assign(R, 0.0);
  // Like the reference, where every MPI rank appends the rows of its own block (partition_list) to its own CSR object: one block of rows per OpenMP
  // thread, each appended in ascending row order, the blocks concatenated in rank order.  Entry order and values do not depend on the thread count.
  const int nblk = std::max(1, std::min(omp_get_max_threads(), R / 64));
  std::vector<std::vector<int>> b_len(nblk), b_ind(nblk); std::vector<std::vector<double>> b_val(nblk);
  std::vector<int> r_lo(nblk + 1);
  for (int b = 0; b <= nblk; ++b) r_lo[b] = 1 + (int)((long long)R * b / nblk);
#pragma omp parallel for schedule(static, 1)
  for (int blk = 0; blk < nblk; ++blk) {
  std::vector<int> ti_copy(mesh.nC_mem); std::vector<double> wti_copy(mesh.nC_mem);
  struct Local { std::vector<int>& len; std::vector<int>& ind; std::vector<double>& val; int row0;
                 void add(int i, int j, double v) { ind.push_back(j); val.push_back(v); ++len[i - row0]; } } A{b_len[blk], b_ind[blk], b_val[blk], r_lo[blk]};
  A.len.assign(r_lo[blk + 1] - r_lo[blk], 0);
  A.ind.reserve((size_t)(r_lo[blk + 1] - r_lo[blk]) * 22); A.val.reserve((size_t)(r_lo[blk + 1] - r_lo[blk]) * 22);
  for (int row = r_lo[blk]; row < r_lo[blk + 1]; ++row) {
    // n2tiuv: row = 2(ti-1)+uv (mesh_translation_tables.f90:190-198)
    const int ti = (row + 1) / 2, uv = (row % 2 == 1) ? 1 : 2;
    if (BC_mask && BC_mask[ti - 1] == 1) {
      A.add(row, row, 1.0);
      bb[row - 1] = (uv == 1) ? BC_u[ti - 1] : BC_v[ti - 1];
    } else if (mesh.TriBI[ti - 1] > 0) {
      int side;  // north, east, south, west
      switch (mesh.TriBI[ti - 1]) { case 1: case 2: side = 0; break; case 3: case 4: side = 1; break; case 5: case 6: side = 2; break; default: side = 3; }
      const int choice = (uv == 1) ? C.BC_u[side] : C.BC_v[side];
      if (choice == BC_infinite) {
        int n_neighbours = 0;
        for (int n = 1; n <= 3; ++n) {
          int tj = mesh.TriC[(size_t)(n - 1) * nTri + ti - 1];
          if (tj == 0) continue;
          ++n_neighbours;
          A.add(row, 2 * (tj - 1) + uv, 1.0);
        }
        A.add(row, row, -1.0 * (double)n_neighbours);
        bb[row - 1] = 0.0;
      } else if (choice == BC_periodic_ISMIP_HOM || (choice == BC_infinite_SSA_icestream && uv == 1)) {
        // value at ti = relaxed, weighted previous solution around the displaced point (:547-581, :618-636)
        if (!u_b_prev || !v_b_prev) { std::fprintf(stderr, "oracle: copy BC rows need u_b_prev, v_b_prev\n"); std::abort(); }
        const double* prev = (uv == 1) ? u_b_prev : v_b_prev;
        find_ti_copy(mesh, choice, C.ISMIP_HOM_L, ti, ti_copy.data(), wti_copy.data());
        A.add(row, row, 1.0);
        double fixed = 0.0;
        for (int n = 0; n < mesh.nC_mem; ++n) {
          const int tj = ti_copy[n];
          if (tj == 0) continue;
          fixed = fixed + wti_copy[n] * prev[tj - 1];
        }
        fixed = (C.visc_it_relax * fixed) + ((1.0 - C.visc_it_relax) * prev[ti - 1]);
        bb[row - 1] = fixed;
      } else if (choice != BC_zero) {  // crash('unknown choice_BC_u/v') in the reference
        std::fprintf(stderr, "oracle: unknown choice_BC_%c %d\n", uv == 1 ? 'u' : 'v', choice); std::abort();
      } else {  // 'zero'
        A.add(row, row, 1.0);
        bb[row - 1] = 0.0;
      }
    } else {
      const double N = N_b[ti - 1], dNx = dN_dx_b[ti - 1], dNy = dN_dy_b[ti - 1], beta = beta_b[ti - 1];
      const int k1 = Mx.ptr[ti - 1], k2 = Mx.ptr[ti] - 1;
      for (int k = k1; k <= k2; ++k) {
        const int tj = Mx.ind[k - 1];
        const double ddx = Mx.val[k - 1], ddy = My.val[k - 1], dxx = Mxx.val[k - 1], dxy = Mxy.val[k - 1], dyy = Myy.val[k - 1];
        double Au, Av;
        if (C.do_include_SSADIVA_crossterms) {
          if (uv == 1) {
            Au = 4.0 * N * dxx + 4.0 * dNx * ddx + N * dyy + dNy * ddy;
            if (tj == ti) Au = Au - beta;
            Av = 3.0 * N * dxy + 2.0 * dNx * ddy + dNy * ddx;
          } else {
            Av = 4.0 * N * dyy + 4.0 * dNy * ddy + N * dxx + dNx * ddx;
            if (tj == ti) Av = Av - beta;
            Au = 3.0 * N * dxy + 2.0 * dNy * ddx + dNx * ddy;
          }
        } else {
          if (uv == 1) {
            Au = 4.0 * dxx + dyy;
            if (tj == ti) Au = Au - beta / N;
            Av = 3.0 * dxy;
          } else {
            Av = 4.0 * dyy + dxx;
            if (tj == ti) Av = Av - beta / N;
            Au = 3.0 * dxy;
          }
        }
        A.add(row, 2 * (tj - 1) + 1, Au);
        A.add(row, 2 * (tj - 1) + 2, Av);
      }
      const double tau = (uv == 1) ? tau_dx_b[ti - 1] : tau_dy_b[ti - 1];
      bb[row - 1] = C.do_include_SSADIVA_crossterms ? -tau : -tau / N;
    }
  }
  }
  // concatenate (ptr as add_entry_CSR_dist + finalise_matrix_CSR_dist leave it: 1-based running offsets)
  size_t nnz = 0;
  for (int b = 0; b < nblk; ++b) nnz += b_ind[b].size();
  A.init(R, R, 0);
  A.ind.resize(nnz); A.val.resize(nnz);
  size_t off = 0;
  for (int b = 0; b < nblk; ++b) {
    int run = (int)off + 1;
    for (int r = r_lo[b]; r < r_lo[b + 1]; ++r) { run += b_len[b][r - r_lo[b]]; A.ptr[r] = run; }
    std::memcpy(A.ind.data() + off, b_ind[b].data(), sizeof(int) * b_ind[b].size());
    std::memcpy(A.val.data() + off, b_val[b].data(), sizeof(double) * b_val[b].size());
    off += b_ind[b].size();
  }
}

namespace {

int linear_solve(const Config& C, const Csr& A, const double* b, double* x, double rtol, double abstol, int* its) {
  const bool g = C.ksp_initial_guess_nonzero != 0;
  return ksp_solve(C.ksp_type, A, b, x, rtol, abstol, C.ksp_maxit, its, g);
}

// solve_SSA_DIVA_linearised (solve_linearised_SSA_DIVA_infinite_slab.f90:23-178)
void solve_linearised(const MeshView& mesh, const Operators& ops, const Config& C, double* u_b, double* v_b,
                      const double* N_b, const double* dN_dx_b, const double* dN_dy_b, const double* beta_b,
                      const double* tau_dx_b, const double* tau_dy_b, double* u_b_prev, double* v_b_prev,
                      int* n_Axb_its, const int* BC_mask, const double* BC_u, const double* BC_v) {
  const int nTri = mesh.nTri;
  for (int t = 0; t < nTri; ++t) { u_b_prev[t] = u_b[t]; v_b_prev[t] = v_b[t]; }
  Csr A; std::vector<double> bb, uv(2 * (size_t)nTri);
  assemble_SSA_DIVA(mesh, ops, C, N_b, dN_dx_b, dN_dy_b, beta_b, tau_dx_b, tau_dy_b, BC_mask, BC_u, BC_v, A, bb, u_b_prev, v_b_prev);
  for (int t = 0; t < nTri; ++t) { uv[2 * t] = u_b[t]; uv[2 * t + 1] = v_b[t]; }
  linear_solve(C, A, bb.data(), uv.data(), C.stress_balance_PETSc_rtol, C.stress_balance_PETSc_abstol, n_Axb_its);
  for (int t = 0; t < nTri; ++t) { u_b[t] = uv[2 * t]; v_b[t] = uv[2 * t + 1]; }
}

// multiply_CSR_matrix_with_vector_2D: nz independent 1-D products (CSR_matrix_vector_multiplication.f90:337-365)
void spmv_3D(const Csr& M, const double* x, double* y, int nz) {
  for (int k = 0; k < nz; ++k) spmv(M, x + (size_t)k * M.n, y + (size_t)k * M.m);
}

double vertical_average(const double* zeta, const double* f, int nz, size_t stride) {
  double s = 0.0;
  for (int k = 0; k < nz - 1; ++k) s = s + 0.5 * (f[(k + 1) * stride] + f[k * stride]) * (zeta[k + 1] - zeta[k]);
  return s;
}

// calc_basal_friction_coefficient (sliding_laws.f90:25-76); do_subgrid_friction_on_A_grid = .false. (default, :344)
void sliding_law(const MeshView& mesh, const Config& C, const IceInputs& ice, const double* u_a, const double* v_a, double* beta_a) {
  const int nV = mesh.nV;
  const double pi = 4.0 * std::atan(1.0);
  std::vector<double> tauc;
  const int law = C.choice_sliding_law;
  if (law == SLID_Coulomb || law == SLID_Budd || law == SLID_ZoetIverson) {
    tauc.resize(nV);
    for (int v = 0; v < nV; ++v) tauc[v] = ice.effective_pressure[v] * std::tan(pi / 180.0) * ice.till_friction_angle[v];
    // extend_till_yield_stress_to_neighbours (:638-695)
    std::vector<double> tot(tauc);
    for (int vi = 1; vi <= nV; ++vi) {
      if (!ice.mask_icefree_land[vi - 1]) continue;
      bool found = false;
      double mn = 1000.0 * C.ice_density * C.grav;
      for (int ci = 1; ci <= mesh.nC[vi - 1]; ++ci) {
        int vc = mesh.C[(size_t)(ci - 1) * nV + vi - 1];
        if (ice.mask_grounded_ice[vc - 1]) { mn = std::min(mn, tot[vc - 1]); found = true; }
      }
      tauc[vi - 1] = found ? mn : C.Hi_min * C.ice_density * C.grav;
    }
  }
#pragma omp parallel for schedule(static)
  for (int v = 0; v < nV; ++v) {
    const double uabs = std::sqrt(C.slid_delta_v * C.slid_delta_v + u_a[v] * u_a[v] + v_a[v] * v_a[v]);
    double b = 0.0;
    switch (law) {
      case SLID_no_sliding: b = 0.0; break;
      case SLID_Weertman: b = ice.beta_sq[v] * std::pow(uabs, 1.0 / C.slid_Weertman_m - 1.0); break;
      case SLID_Coulomb: b = tauc[v] / uabs; break;
      case SLID_Budd: b = tauc[v] * std::pow(uabs, C.slid_Budd_q_plastic - 1.0) / std::pow(C.slid_Budd_u_threshold, C.slid_Budd_q_plastic); break;
      case SLID_Tsai2015:
        b = std::min(ice.alpha_sq[v] * ice.effective_pressure[v], ice.beta_sq[v] * std::pow(uabs, 1.0 / C.slid_Weertman_m)) * std::pow(uabs, -1.0);
        break;
      case SLID_Schoof2005: {
        const double m = C.slid_Weertman_m, aN = ice.alpha_sq[v] * ice.effective_pressure[v];
        b = ((ice.beta_sq[v] * std::pow(uabs, 1.0 / m) * aN) /
             std::pow(std::pow(ice.beta_sq[v], m) * uabs + std::pow(aN, m), 1.0 / m)) * std::pow(uabs, -1.0);
        break;
      }
      case SLID_ZoetIverson:
        b = tauc[v] * std::pow(uabs, 1.0 / C.slid_ZI_p - 1.0) * std::pow(uabs + C.slid_ZI_ut, -1.0 / C.slid_ZI_p);
        break;
      case SLID_idealised_SSA_icestream: {
        double u, ty;
        Schoof2006_icestream(C.uniform_Glens_flow_factor, C.Glens_flow_law_exponent, C.icestream_Hi, C.icestream_dhdx,
                             C.icestream_L, C.icestream_m, mesh.V[(size_t)nV + v], &u, &ty);
        b = ty / uabs;
        break;
      }
    }
    beta_a[v] = std::min(C.slid_beta_max, b);
  }
}

void driving_stress(const MeshView& mesh, const Operators& ops, const Config& C, const IceInputs& ice, double* tdx, double* tdy) {
  const int nTri = mesh.nTri;
  std::vector<double> Hi_b(nTri), dx(nTri), dy(nTri);
  spmv(ops.M[OP_map_a_b], ice.Hi, Hi_b.data());
  spmv(ops.M[OP_ddx_a_b], ice.Hs, dx.data());
  spmv(ops.M[OP_ddy_a_b], ice.Hs, dy.data());
  for (int t = 0; t < nTri; ++t) {
    tdx[t] = -C.ice_density * C.grav * Hi_b[t] * dx[t];
    tdy[t] = -C.ice_density * C.grav * Hi_b[t] * dy[t];
  }
}

void velocity_limits_relax(const Config& C, int nTri, double* u, double* v, const double* up, const double* vp, double relax) {
  for (int t = 0; t < nTri; ++t) {
    double uabs = std::sqrt(u[t] * u[t] + v[t] * v[t]);
    if (uabs > C.vel_max) { u[t] = u[t] * C.vel_max / uabs; v[t] = v[t] * C.vel_max / uabs; }
  }
  for (int t = 0; t < nTri; ++t) {
    u[t] = (relax * u[t]) + ((1.0 - relax) * up[t]);
    v[t] = (relax * v[t]) + ((1.0 - relax) * vp[t]);
  }
}

double L2_norm_uv(int nTri, const double* u, const double* v, const double* up, const double* vp) {
  double res1 = 0.0, res2 = 0.0;
  for (int t = 0; t < nTri; ++t)
    if (u[t] == u[t] && v[t] == v[t]) {
      res1 += (u[t] - up[t]) * (u[t] - up[t]); res1 += (v[t] - vp[t]) * (v[t] - vp[t]);
      res2 += (u[t] + up[t]) * (u[t] + up[t]); res2 += (v[t] + vp[t]) * (v[t] + vp[t]);
    }
  return 2.0 * res1 / std::max(res2, 1e-8);
}

}  // namespace

int solve_DIVA(const MeshView& mesh, const Operators& ops, const Config& C, const IceInputs& ice, DivaState& D,
               const int* BC_mask, const double* BC_u, const double* BC_v, int* n_visc_its, int* n_Axb_its,
               int max_visc_its_override) {
  const int nV = mesh.nV, nTri = mesh.nTri, nz = mesh.nz;
  const double* zeta = mesh.zeta;
  *n_visc_its = 0; *n_Axb_its = 0;
  bool grounded = false;
  for (int v = 0; v < nV; ++v) grounded = grounded || ice.mask_grounded_ice[v];
  if (!grounded) {
    for (auto* f : {&D.u_vav_b, &D.v_vav_b, &D.u_base_b, &D.v_base_b, &D.u_3D_b, &D.v_3D_b}) std::fill(f->begin(), f->end(), 0.0);
    return 0;
  }
  driving_stress(mesh, ops, C, ice, D.tau_dx_b.data(), D.tau_dy_b.data());

  double L2_uv = 1e9, relax = C.visc_it_relax, eps0 = C.Glens_flow_law_epsilon_sq_0;
  int nit_diverg = 0, it = 0;
  const double n_glen = C.Glens_flow_law_exponent;
  std::vector<double> dz_b((size_t)nTri * nz), dz2_b((size_t)nTri * nz), u_base_a(nV), v_base_a(nV);
  if (nz > 64) return 5;
  bool converged = false;
  while (!converged) {
    ++it;
    // calc_horizontal_strain_rates (SSA_DIVA_utilities.f90:65-94)
    spmv(ops.M[OP_ddx_b_a], D.u_vav_b.data(), D.du_dx_a.data());
    spmv(ops.M[OP_ddy_b_a], D.u_vav_b.data(), D.du_dy_a.data());
    spmv(ops.M[OP_ddx_b_a], D.v_vav_b.data(), D.dv_dx_a.data());
    spmv(ops.M[OP_ddy_b_a], D.v_vav_b.data(), D.dv_dy_a.data());
    // calc_vertical_shear_strain_rates (DIVA_solver_infinite_slab.f90:235-273)
#pragma omp parallel for collapse(2) schedule(static)
    for (int k = 0; k < nz; ++k)
      for (int t = 0; t < nTri; ++t) {
        const double e = std::max(C.visc_eff_min, D.eta_3D_b[(size_t)k * nTri + t]);
        dz_b[(size_t)k * nTri + t] = D.tau_bx_b[t] * zeta[k] / e;
        dz2_b[(size_t)k * nTri + t] = D.tau_by_b[t] * zeta[k] / e;
      }
    spmv_3D(ops.M[OP_map_b_a], dz_b.data(), D.du_dz_3D_a.data(), nz);
    spmv_3D(ops.M[OP_map_b_a], dz2_b.data(), D.dv_dz_3D_a.data(), nz);
    // calc_effective_viscosity (:275-351)
    const double A_min = 1e-18;
    const double eta_max = 0.5 * std::pow(A_min, -1.0 / n_glen) * std::pow(eps0, (1.0 - n_glen) / (2.0 * n_glen));
#pragma omp parallel for collapse(2) schedule(static)
    for (int k = 0; k < nz; ++k)
      for (int v = 0; v < nV; ++v) {
        const size_t i = (size_t)k * nV + v;
        const double eps_sq = D.du_dx_a[v] * D.du_dx_a[v] + D.dv_dy_a[v] * D.dv_dy_a[v] + D.du_dx_a[v] * D.dv_dy_a[v] +
                              0.25 * (D.du_dy_a[v] + D.dv_dx_a[v]) * (D.du_dy_a[v] + D.dv_dx_a[v]) +
                              0.25 * (D.du_dz_3D_a[i] * D.du_dz_3D_a[i] + D.dv_dz_3D_a[i] * D.dv_dz_3D_a[i]) + eps0;
        double eta = 0.5 * std::pow(ice.A_flow_3D[i], -1.0 / n_glen) * std::pow(eps_sq, (1.0 - n_glen) / (2.0 * n_glen));
        D.eta_3D_a[i] = std::min(std::max(eta, C.visc_eff_min), eta_max);
      }
    spmv_3D(ops.M[OP_map_a_b], D.eta_3D_a.data(), D.eta_3D_b.data(), nz);
#pragma omp parallel for schedule(static)
    for (int v = 0; v < nV; ++v) {
      D.eta_vav_a[v] = vertical_average(zeta, &D.eta_3D_a[v], nz, nV);
      D.N_a[v] = D.eta_vav_a[v] * std::max((double)0.1f, ice.Hi[v]);  // max( 0.1, Hi): default-real literal (:341)
    }
    spmv(ops.M[OP_map_a_b], D.N_a.data(), D.N_b.data());
    spmv(ops.M[OP_ddx_a_b], D.N_a.data(), D.dN_dx_b.data());
    spmv(ops.M[OP_ddy_a_b], D.N_a.data(), D.dN_dy_b.data());
    // calc_F_integrals (:353-395)
#pragma omp parallel for schedule(static)
    for (int v = 0; v < nV; ++v) {
      const double Hc = -std::max(0.1, ice.Hi[v]);
      double prof[64], integ[64];   // nz <= 64
      for (int pw = 1; pw <= 2; ++pw) {
        for (int k = 0; k < nz; ++k) prof[k] = ((pw == 1) ? zeta[k] : zeta[k] * zeta[k]) / D.eta_3D_a[(size_t)k * nV + v];
        integ[nz - 1] = 0.0;
        for (int k = nz - 2; k >= 0; --k) integ[k] = integ[k + 1] - 0.5 * (prof[k + 1] + prof[k]) * (zeta[k + 1] - zeta[k]);
        double* F = (pw == 1) ? D.F1_3D_a.data() : D.F2_3D_a.data();
        for (int k = 0; k < nz; ++k) F[(size_t)k * nV + v] = Hc * integ[k];
      }
    }
    spmv_3D(ops.M[OP_map_a_b], D.F1_3D_a.data(), D.F1_3D_b.data(), nz);
    spmv_3D(ops.M[OP_map_a_b], D.F2_3D_a.data(), D.F2_3D_b.data(), nz);
    // calc_effective_basal_friction_coefficient (:397-457)
    spmv(ops.M[OP_map_b_a], D.u_base_b.data(), u_base_a.data());
    spmv(ops.M[OP_map_b_a], D.v_base_b.data(), v_base_a.data());
    sliding_law(mesh, C, ice, u_base_a.data(), v_base_a.data(), D.basal_friction_coefficient_a.data());
#pragma omp parallel for schedule(static)
    for (int v = 0; v < nV; ++v) {
      const double b = D.basal_friction_coefficient_a[v];
      D.beta_eff_a[v] = (C.choice_sliding_law == SLID_no_sliding) ? 1.0 / D.F2_3D_a[v] : b / (1.0 + b * D.F2_3D_a[v]);
    }
    spmv(ops.M[OP_map_a_b], D.basal_friction_coefficient_a.data(), D.basal_friction_coefficient_b.data());
    spmv(ops.M[OP_map_a_b], D.beta_eff_a.data(), D.beta_eff_b.data());
    if (C.do_GL_subgrid_friction)
      for (int t = 0; t < nTri; ++t) D.beta_eff_b[t] = D.beta_eff_b[t] * std::pow(ice.fraction_gr_b[t], C.subgrid_friction_exponent_on_B_grid);
    // linearised solve
    int its = 0;
    solve_linearised(mesh, ops, C, D.u_vav_b.data(), D.v_vav_b.data(), D.N_b.data(), D.dN_dx_b.data(), D.dN_dy_b.data(),
                     D.beta_eff_b.data(), D.tau_dx_b.data(), D.tau_dy_b.data(), D.u_b_prev.data(), D.v_b_prev.data(), &its,
                     BC_mask, BC_u, BC_v);
    *n_Axb_its += its;
    velocity_limits_relax(C, nTri, D.u_vav_b.data(), D.v_vav_b.data(), D.u_b_prev.data(), D.v_b_prev.data(), relax);
    // calc_basal_velocities (:487-523), calc_basal_shear_stress (:459-485)
#pragma omp parallel for schedule(static)
    for (int t = 0; t < nTri; ++t) {
      if (C.choice_sliding_law == SLID_no_sliding) { D.u_base_b[t] = 0.0; D.v_base_b[t] = 0.0; }
      else {
        const double den = 1.0 + D.basal_friction_coefficient_b[t] * D.F2_3D_b[t];
        D.u_base_b[t] = D.u_vav_b[t] / den; D.v_base_b[t] = D.v_vav_b[t] / den;
      }
      D.tau_bx_b[t] = D.u_vav_b[t] * D.beta_eff_b[t];
      D.tau_by_b[t] = D.v_vav_b[t] * D.beta_eff_b[t];
    }
    const double L2_prev = L2_uv;
    L2_uv = L2_norm_uv(nTri, D.u_vav_b.data(), D.v_vav_b.data(), D.u_b_prev.data(), D.v_b_prev.data());
    if (L2_uv > L2_prev) ++nit_diverg; else nit_diverg = 0;
    if (nit_diverg > 2) { nit_diverg = 0; relax *= 0.9; eps0 *= 1.2; }
    if (relax < 0.05 || eps0 > 1e-5) return 4;  // crash() in the reference
    converged = L2_uv < C.visc_it_norm_dUV_tol;
    const int nit = (max_visc_its_override >= 0) ? max_visc_its_override : C.visc_it_nit;
    if (max_visc_its_override >= 0 ? it >= nit : it > nit) break;
  }
  // calc_3D_velocities (:525-568)
#pragma omp parallel for collapse(2) schedule(static)
  for (int k = 0; k < nz; ++k)
    for (int t = 0; t < nTri; ++t) {
      const size_t i = (size_t)k * nTri + t;
      if (C.choice_sliding_law == SLID_no_sliding) {
        D.u_3D_b[i] = D.tau_bx_b[t] * D.F1_3D_b[i]; D.v_3D_b[i] = D.tau_by_b[t] * D.F1_3D_b[i];
      } else {
        D.u_3D_b[i] = D.u_base_b[t] * (1.0 + D.basal_friction_coefficient_b[t] * D.F1_3D_b[i]);
        D.v_3D_b[i] = D.v_base_b[t] * (1.0 + D.basal_friction_coefficient_b[t] * D.F1_3D_b[i]);
      }
    }
  *n_visc_its = it;
  return 0;
}

int solve_SSA(const MeshView& mesh, const Operators& ops, const Config& C, const IceInputs& ice, DivaState& D,
              const int* BC_mask, const double* BC_u, const double* BC_v, int* n_visc_its, int* n_Axb_its,
              int max_visc_its_override) {
  const int nV = mesh.nV, nTri = mesh.nTri, nz = mesh.nz;
  *n_visc_its = 0; *n_Axb_its = 0;
  bool grounded = false;
  for (int v = 0; v < nV; ++v) grounded = grounded || ice.mask_grounded_ice[v];
  if (!grounded || C.choice_sliding_law == SLID_no_sliding) {
    std::fill(D.u_vav_b.begin(), D.u_vav_b.end(), 0.0); std::fill(D.v_vav_b.begin(), D.v_vav_b.end(), 0.0);
    return 0;
  }
  driving_stress(mesh, ops, C, ice, D.tau_dx_b.data(), D.tau_dy_b.data());
  double L2_uv = 1e9, relax = C.visc_it_relax, eps0 = C.Glens_flow_law_epsilon_sq_0;
  int nit_diverg = 0, it = 0;
  const double n_glen = C.Glens_flow_law_exponent;
  std::vector<double> u_a(nV), v_a(nV), A_vav(nV);
  for (int v = 0; v < nV; ++v) A_vav[v] = vertical_average(mesh.zeta, &ice.A_flow_3D[v], nz, nV);
  bool converged = false;
  while (!converged) {
    ++it;
    spmv(ops.M[OP_ddx_b_a], D.u_vav_b.data(), D.du_dx_a.data());
    spmv(ops.M[OP_ddy_b_a], D.u_vav_b.data(), D.du_dy_a.data());
    spmv(ops.M[OP_ddx_b_a], D.v_vav_b.data(), D.dv_dx_a.data());
    spmv(ops.M[OP_ddy_b_a], D.v_vav_b.data(), D.dv_dy_a.data());
    // calc_effective_viscosity (SSA_main.f90:342-395)
    const double eta_max = 0.5 * std::pow(1e-18, -1.0 / n_glen) * std::pow(eps0, (1.0 - n_glen) / (2.0 * n_glen));
    for (int v = 0; v < nV; ++v) {
      const double eps_sq = D.du_dx_a[v] * D.du_dx_a[v] + D.dv_dy_a[v] * D.dv_dy_a[v] + D.du_dx_a[v] * D.dv_dy_a[v] +
                            0.25 * (D.du_dy_a[v] + D.dv_dx_a[v]) * (D.du_dy_a[v] + D.dv_dx_a[v]) + eps0;
      double eta = 0.5 * std::pow(A_vav[v], -1.0 / n_glen) * std::pow(eps_sq, (1.0 - n_glen) / (2.0 * n_glen));
      eta = std::min(std::max(eta, C.visc_eff_min), eta_max);
      D.eta_vav_a[v] = eta;
      D.N_a[v] = eta * std::max(0.1, ice.Hi[v]);
    }
    spmv(ops.M[OP_map_a_b], D.N_a.data(), D.N_b.data());
    spmv(ops.M[OP_ddx_a_b], D.N_a.data(), D.dN_dx_b.data());
    spmv(ops.M[OP_ddy_a_b], D.N_a.data(), D.dN_dy_b.data());
    // calc_applied_basal_friction_coefficient (:397-445)
    spmv(ops.M[OP_map_b_a], D.u_vav_b.data(), u_a.data());
    spmv(ops.M[OP_map_b_a], D.v_vav_b.data(), v_a.data());
    sliding_law(mesh, C, ice, u_a.data(), v_a.data(), D.basal_friction_coefficient_a.data());
    spmv(ops.M[OP_map_a_b], D.basal_friction_coefficient_a.data(), D.basal_friction_coefficient_b.data());
    if (C.do_GL_subgrid_friction)
      for (int t = 0; t < nTri; ++t)
        D.basal_friction_coefficient_b[t] = D.basal_friction_coefficient_b[t] * std::pow(ice.fraction_gr_b[t], C.subgrid_friction_exponent_on_B_grid);
    int its = 0;
    solve_linearised(mesh, ops, C, D.u_vav_b.data(), D.v_vav_b.data(), D.N_b.data(), D.dN_dx_b.data(), D.dN_dy_b.data(),
                     D.basal_friction_coefficient_b.data(), D.tau_dx_b.data(), D.tau_dy_b.data(), D.u_b_prev.data(), D.v_b_prev.data(),
                     &its, BC_mask, BC_u, BC_v);
    *n_Axb_its += its;
    velocity_limits_relax(C, nTri, D.u_vav_b.data(), D.v_vav_b.data(), D.u_b_prev.data(), D.v_b_prev.data(), relax);
    const double L2_prev = L2_uv;
    L2_uv = L2_norm_uv(nTri, D.u_vav_b.data(), D.v_vav_b.data(), D.u_b_prev.data(), D.v_b_prev.data());
    if (L2_uv > L2_prev) ++nit_diverg; else nit_diverg = 0;
    if (nit_diverg > 2) { nit_diverg = 0; relax *= 0.9; eps0 *= 1.2; }
    if (relax < 0.05 || eps0 > 1e-5) return 4;
    converged = L2_uv < C.visc_it_norm_dUV_tol;
    const int nit = (max_visc_its_override >= 0) ? max_visc_its_override : C.visc_it_nit;
    if (max_visc_its_override >= 0 ? it >= nit : it > nit) break;
  }
  *n_visc_its = it;
  return 0;
}

// ---------------------------------------------------------------------------------------------------
// Analytic solutions (src/UPSY/basic/analytical_solutions/{Schoof,Halfar}_*.f90)
// ---------------------------------------------------------------------------------------------------
void Schoof2006_icestream(double A, double n, double H, double dhdx, double L, double m, double y, double* u, double* tauc) {
  (void)n;
  const double f = -917.0 * 9.81 * H * dhdx;
  const double B = std::pow(A, -1.0 / 3.0);
  const double W = L * std::pow(m + 1.0, 1.0 / m);
  *tauc = f * std::pow(std::fabs(y / L), m);
  const double ua = -2.0 * f * f * f * L * L * L * L / (B * B * B * H * H * H);
  const double ub = (1.0 / 4.0) * (std::pow(y / L, 4.0) - std::pow(m + 1.0, 4.0 / m));
  const double uc = (-3.0 / ((m + 1.0) * (m + 4.0))) * (std::pow(std::fabs(y / L), m + 4.0) - std::pow(m + 1.0, 1.0 + 4.0 / m));
  const double ud = (3.0 / ((m + 1.0) * (m + 1.0) * (2.0 * m + 4.0))) * (std::pow(std::fabs(y / L), 2 * m + 4.0) - std::pow(m + 1.0, 2.0 + 4.0 / m));
  const double ue = (-1.0 / ((m + 1.0) * (m + 1.0) * (m + 1.0) * (3.0 * m + 4.0))) * (std::pow(std::fabs(y / L), 3 * m + 4.0) - std::pow(m + 1.0, 3.0 + 4.0 / m));
  *u = ua * (ub + uc + ud + ue);
  if (std::fabs(y) > W) *u = 0.0;
}

double Halfar_H(double A, double n, double H0, double R0, double x, double y, double t_yr) {
  const double sec_per_year = 31556943.36;
  const double Gamma = (2.0 / 5.0) * (A / sec_per_year) * std::pow(917.0 * 9.81, n);
  const double t0 = 1.0 / ((5.0 * n + 3.0) * Gamma) * std::pow((2.0 * n + 1.0) / (n + 1.0), n) * std::pow(R0, n + 1.0) / std::pow(H0, 2.0 * n + 1.0);
  const double tt = (t0 + t_yr * sec_per_year) / t0;
  const double f1 = std::pow(tt, -2.0 / (5.0 * n + 3.0)), f2 = std::pow(tt, -1.0 / (5.0 * n + 3.0));
  const double f3 = std::sqrt(x * x + y * y) / R0;
  const double G = 1.0 - std::pow(std::min(1.0, f2 * f3), (n + 1.0) / n);
  return H0 * f1 * std::pow(G, n / (2.0 * n + 1.0));
}

}  // namespace orc

// ---------------------------------------------------------------------------------------------------
// apply_grounded_fractions_to_bed_roughness (sliding_laws.f90:568-636) with C%do_subgrid_friction_on_A_grid = .true.
// ---------------------------------------------------------------------------------------------------
void orc::apply_grounded_fractions_to_bed_roughness(int nV, const double* Hi, const double* Hs_slope, const double* fraction_gr, const int* mask_gl_gr,
                                                    const int* mask_cf_gr, const int* mask_gl_fl, const int* mask_grounded_ice,
                                                    const int* mask_floating_ice, const int* mask_icefree_ocean, const double* bed_roughness,
                                                    double* bed_roughness_applied) {
  for (int vi = 0; vi < nV; ++vi) {
    double weight_gr = 1.0;
    const double exponent_hi = std::log10(std::max(1.0, Hi[vi]));
    const double exponent_hs = Hs_slope[vi] / 0.005;
    const double exponent_gr = std::max(0.0, exponent_hi - exponent_hs);
    if (mask_gl_gr[vi]) weight_gr = std::pow(fraction_gr[vi], exponent_gr);
    else if (mask_cf_gr[vi]) weight_gr = std::pow(fraction_gr[vi], exponent_gr);
    else if (mask_gl_fl[vi]) weight_gr = std::pow(fraction_gr[vi], exponent_gr);
    else if (mask_grounded_ice[vi]) weight_gr = 1.0;
    else if (mask_floating_ice[vi]) weight_gr = 0.0;
    else if (mask_icefree_ocean[vi]) weight_gr = 0.0;
    weight_gr = std::min(1.0, std::max(0.0, weight_gr));
    bed_roughness_applied[vi] = bed_roughness[vi] * weight_gr;
  }
}
