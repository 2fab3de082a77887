// SSA / DIVA velocity solve on the device (north_star (a); K3-K7, K10 of SURVEY 2e).
// Reference (paths relative to src/UFEMISM/ice_dynamics/conservation_of_momentum):
//   solve_DIVA_infinite_slab + helpers          SSA_DIVA/DIVA_solver_infinite_slab.f90:52-568
//   solve_SSA + helpers                         SSA_DIVA/SSA_main.f90:88-245, 315-445
//   solve_SSA_DIVA_linearised + row builders    SSA_DIVA/solve_linearised_SSA_DIVA_infinite_slab.f90:23-641
//   driving stress, strain rates, relax, limits, L2 norm   SSA_DIVA/SSA_DIVA_utilities.f90:24-206
//   Glen viscosity                              ../rheology/constitutive_equation.f90:25-80
//   sliding laws                                sliding_laws.f90:25-695
//   zeta integrals                              src/UPSY/mesh/mesh_zeta.f90:175-295
//
// The reference performs ~71 SpMV-equivalents and rebuilds the CSR matrix with per-row heap allocation in every
// viscosity iteration.  Here one viscosity iteration is three fused kernels around the Krylov solve:
//   k_vertex_*    one thread per vertex: b->a gathers (strain rates, shear strain, basal velocity), Glen viscosity
//                 column, vertical average, F-integrals, sliding law, beta_eff
//   k_tri_assemble one thread per triangle: a->b gathers (N, dN/dx, dN/dy, eta column, beta, beta_eff, F2) and the
//                 values of the two stiffness-matrix rows of the triangle, in place on a pattern built once
//   k_epilogue    velocity limits, relaxation, basal velocity / shear stress, L2 norm partial sums
#include "velocity.cuh"

#include <cub/device/device_scan.cuh>
#include <cstring>

#include "csr.cuh"
#include "krylov_impl.cuh"
#include "mesh.cuh"

namespace ufe {

void velocity_free(Velocity* v) { delete v; }

int Velocity::alloc(Ctx* c) {
  const Mesh& M = *c->mesh;
  nz = M.nz;
  UFE_CHECK(nz <= kMaxNz, "nz = %d exceeds the supported maximum of %d layers", nz, kMaxNz);
  const size_t na = M.a.n_ext(), nb = M.b.n_ext(), nzz = nz;
  for (auto* f : {&Hi, &Hs, &A_flow_vav, &till_friction_angle, &beta_sq, &alpha_sq, &effective_pressure, &tauc0, &tauc, &du_dx, &du_dy, &dv_dx,
                  &dv_dy, &eta_vav, &N_a, &beta_a, &beta_eff_a})
    UFE_TRY(f->alloc(na));
  for (auto* f : {&A_flow, &du_dz_3D, &dv_dz_3D, &eta_3D_a, &F1_3D_a, &F2_3D_a}) UFE_TRY(f->alloc(na * nzz));
  for (auto* f : {&fraction_gr_b, &u, &v, &u_base, &v_base, &tau_bx, &tau_by, &u_prev, &v_prev, &tau_dx, &tau_dy, &N_b, &dN_dx_b, &dN_dy_b,
                  &beta_b, &beta_eff_b, &bc_u, &bc_v})
    UFE_TRY(f->alloc(nb));
  for (auto* f : {&eta_3D_b, &u_3D, &v_3D, &F1_3D_b, &F2_3D_b}) UFE_TRY(f->alloc(nb * nzz));
  UFE_TRY(mask_grounded_ice.alloc(na)); UFE_TRY(mask_icefree_land.alloc(na)); UFE_TRY(bc_mask.alloc(nb));
  UFE_TRY(bb.alloc(2 * (size_t)M.b.n_loc())); UFE_TRY(uv.alloc(M.buv.n_ext()));
  UFE_TRY(red.alloc(8)); UFE_TRY(partials.alloc((size_t)4 * c->sm_count * 8)); UFE_TRY(counter.alloc(2));
  cudaStream_t st = c->stream;
  // everything starts from zero, like allocate_DIVA_solver (SSA_DIVA/DIVA_main.f90:298-350)
  for (auto* f : {&Hi, &Hs, &A_flow_vav, &till_friction_angle, &beta_sq, &alpha_sq, &effective_pressure, &tauc0, &tauc, &du_dx, &du_dy, &dv_dx,
                  &dv_dy, &eta_vav, &N_a, &beta_a, &beta_eff_a, &A_flow, &du_dz_3D, &dv_dz_3D, &eta_3D_a, &F1_3D_a, &F2_3D_a, &fraction_gr_b,
                  &u, &v, &u_base, &v_base, &tau_bx, &tau_by, &u_prev, &v_prev, &tau_dx, &tau_dy, &N_b, &dN_dx_b, &dN_dy_b, &beta_b,
                  &beta_eff_b, &bc_u, &bc_v, &eta_3D_b, &u_3D, &v_3D, &F1_3D_b, &F2_3D_b, &bb, &uv})
    UFE_CUDA(cudaMemsetAsync(f->p, 0, sizeof(double) * f->n, st));
  UFE_CUDA(cudaMemsetAsync(mask_grounded_ice.p, 0, sizeof(int) * na, st));
  UFE_CUDA(cudaMemsetAsync(mask_icefree_land.p, 0, sizeof(int) * na, st));
  UFE_CUDA(cudaMemsetAsync(bc_mask.p, 0, sizeof(int) * nb, st));
  UFE_CUDA(cudaMemsetAsync(counter.p, 0, sizeof(unsigned) * 2, st));
  return 0;
}

// ---- the C% scalars the kernels read, by value ---------------------------------------------------------------
struct VelPar {
  double visc_eff_min, vel_max, n_glen, eps0, eta_max, relax;
  double slid_delta_v, slid_Weertman_m, slid_Budd_q, slid_Budd_u, slid_ZI_ut, slid_ZI_p, slid_beta_max, subgrid_exp, Hi_min;
  double A_uniform, is_Hi, is_dhdx, is_L, is_m;
  double ice_density, grav;
  double relax_cfg, ismip_hom_L;   // C%visc_it_relax itself (relax above is the adaptive one) and the period of the ISMIP-HOM copy rows
  int sliding_law, do_subgrid, crossterms, BC_u[4], BC_v[4], diagnostics;
};

static VelPar make_par(const ufe_b200_config* C) {
  VelPar p;
  p.visc_eff_min = C->visc_eff_min; p.vel_max = C->vel_max; p.n_glen = C->Glens_flow_law_exponent; p.eps0 = C->Glens_flow_law_epsilon_sq_0;
  p.eta_max = 0.0; p.relax = C->visc_it_relax;
  p.slid_delta_v = C->slid_delta_v; p.slid_Weertman_m = C->slid_Weertman_m; p.slid_Budd_q = C->slid_Budd_q_plastic;
  p.slid_Budd_u = C->slid_Budd_u_threshold; p.slid_ZI_ut = C->slid_ZI_ut; p.slid_ZI_p = C->slid_ZI_p; p.slid_beta_max = C->slid_beta_max;
  p.subgrid_exp = C->subgrid_friction_exponent_on_B_grid; p.Hi_min = C->Hi_min;
  p.A_uniform = C->uniform_Glens_flow_factor; p.is_Hi = C->refgeo_idealised_SSA_icestream_Hi; p.is_dhdx = C->refgeo_idealised_SSA_icestream_dhdx;
  p.is_L = C->refgeo_idealised_SSA_icestream_L; p.is_m = C->refgeo_idealised_SSA_icestream_m;
  p.ice_density = C->ice_density; p.grav = C->grav;
  p.relax_cfg = C->visc_it_relax; p.ismip_hom_L = C->refgeo_idealised_ISMIP_HOM_L;
  p.sliding_law = C->choice_sliding_law; p.do_subgrid = C->do_GL_subgrid_friction; p.crossterms = C->do_include_SSADIVA_crossterms;
  for (int k = 0; k < 4; ++k) { p.BC_u[k] = C->BC_u[k]; p.BC_v[k] = C->BC_v[k]; }
  p.diagnostics = C->write_diagnostics;
  return p;
}

// ---- device helpers ------------------------------------------------------------------------------------------

// Schoof2006_icestream, till yield stress only (src/UPSY/basic/analytical_solutions/Schoof_SSA_solution.f90:41-51)
__device__ __forceinline__ double schoof_tauc(const VelPar& p, double y) {
  const double f = -p.ice_density * p.grav * p.is_Hi * p.is_dhdx;
  return f * pow(fabs(y / p.is_L), p.is_m);
}

// calc_basal_friction_coefficient and the laws it dispatches to (sliding_laws.f90:25-428), one vertex
__device__ __forceinline__ double sliding_beta(const VelPar& p, double u_a, double v_a, double tauc, double beta_sq, double alpha_sq,
                                               double neff, double y) {
  const double uabs = sqrt(p.slid_delta_v * p.slid_delta_v + u_a * u_a + v_a * v_a);
  double b = 0.0;
  switch (p.sliding_law) {
    case UFE_SLID_NO_SLIDING: b = 0.0; break;
    case UFE_SLID_WEERTMAN: b = beta_sq * pow(uabs, 1.0 / p.slid_Weertman_m - 1.0); break;
    case UFE_SLID_COULOMB: b = tauc / uabs; break;
    case UFE_SLID_BUDD: b = tauc * pow(uabs, p.slid_Budd_q - 1.0) / pow(p.slid_Budd_u, p.slid_Budd_q); break;
    case UFE_SLID_TSAI2015: b = fmin(alpha_sq * neff, beta_sq * pow(uabs, 1.0 / p.slid_Weertman_m)) * pow(uabs, -1.0); break;
    case UFE_SLID_SCHOOF2005: {
      const double m = p.slid_Weertman_m, aN = alpha_sq * neff;
      b = ((beta_sq * pow(uabs, 1.0 / m) * aN) / pow(pow(beta_sq, m) * uabs + pow(aN, m), 1.0 / m)) * pow(uabs, -1.0);
      break;
    }
    case UFE_SLID_ZOET_IVERSON: b = tauc * pow(uabs, 1.0 / p.slid_ZI_p - 1.0) * pow(uabs + p.slid_ZI_ut, -1.0 / p.slid_ZI_p); break;
    case UFE_SLID_IDEALISED_SSA_ICESTREAM: b = schoof_tauc(p, y) / uabs; break;
  }
  return fmin(p.slid_beta_max, b);
}

__device__ __forceinline__ double glen_eta(double A, double eps_sq, double n) {
  return 0.5 * pow(A, -1.0 / n) * pow(eps_sq, (1.0 - n) / (2.0 * n));
}

// ---- till yield stress (Coulomb family; independent of the velocity, so computed once per solve) -----------------
// ice%till_yield_stress = effective_pressure * tan(pi/180) * till_friction_angle  (sliding_laws.f90:158,214,379)
__global__ void k_tauc0(int n, int off, const double* __restrict__ neff, const double* __restrict__ phi, double* __restrict__ tauc0) {
  const double t = tan((4.0 * atan(1.0)) / 180.0);
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x)
    tauc0[off + r] = neff[off + r] * t * phi[off + r];
}
// extend_till_yield_stress_to_neighbours (sliding_laws.f90:638-695)
__global__ void k_tauc_extend(int n, int off, int vi1, int a_lo, int nV, const int* __restrict__ nC, const int* __restrict__ C,
                              const int* __restrict__ grounded, const int* __restrict__ icefree_land, const double* __restrict__ tauc0,
                              double* __restrict__ tauc, double default_max, double default_min) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const int vi = vi1 + (int)r;
    double t = tauc0[off + r];
    if (icefree_land[off + r]) {
      bool found = false;
      double mn = default_max;
      for (int ci = 0; ci < nC[vi - 1]; ++ci) {
        const int vc = C[(size_t)ci * nV + vi - 1] - a_lo;
        if (grounded[vc]) { mn = fmin(mn, tauc0[vc]); found = true; }
      }
      t = found ? mn : default_min;
    }
    tauc[off + r] = t;
  }
}

// vertical_average of A_flow for the SSA (SSA_main.f90:315-340)
__global__ void k_A_vav(int n, int off, int n_ext, int nz, const double* __restrict__ zeta, const double* __restrict__ A_flow,
                        double* __restrict__ A_vav) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    double s = 0.0;
    for (int k = 0; k < nz - 1; ++k) s = s + 0.5 * (A_flow[(size_t)(k + 1) * n_ext + off + r] + A_flow[(size_t)k * n_ext + off + r]) * (zeta[k + 1] - zeta[k]);
    A_vav[off + r] = s;
  }
}

// calc_driving_stress (SSA_DIVA_utilities.f90:24-63): the three a->b operators share one pattern
__global__ void k_driving_stress(int n, int off, const int* __restrict__ ptr, const int* __restrict__ ind, const double* __restrict__ w_map,
                                 const double* __restrict__ w_ddx, const double* __restrict__ w_ddy, const double* __restrict__ Hi,
                                 const double* __restrict__ Hs, double* __restrict__ tau_dx, double* __restrict__ tau_dy, double rho_g) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    double Hi_b = 0.0, dx = 0.0, dy = 0.0;
    for (int k = ptr[r]; k < ptr[r + 1]; ++k) {
      const int j = ind[k];
      Hi_b += w_map[k] * Hi[j]; dx += w_ddx[k] * Hs[j]; dy += w_ddy[k] * Hs[j];
    }
    tau_dx[off + r] = -rho_g * Hi_b * dx;
    tau_dy[off + r] = -rho_g * Hi_b * dy;
  }
}

// ---- a-grid (vertex) kernels --------------------------------------------------------------------------------
// out( i,k) [i-major, k fastest] = in( k,i) [the reference's layout]  (RECIP: zeta_k / max( floor, in)) over a whole extended range, through a
// shared-memory tile so that both the reads and the writes are full segments
template <bool RECIP>
__global__ void __launch_bounds__(128) k_layers_interleave(int n_ext, int nz, const double* __restrict__ in, const double* __restrict__ zeta, double floor_,
                                                           double* __restrict__ out) {
  extern __shared__ double tile[];   // [128][nz + 1] (odd pitch for nz = 12: no bank conflicts on the column-wise fill)
  const int pitch = nz + 1;
  for (long long base = (long long)blockIdx.x * 128; base < n_ext; base += (long long)gridDim.x * 128) {
    const int cnt = (int)min((long long)128, n_ext - base);
    if ((int)threadIdx.x < cnt)
      for (int k = 0; k < nz; ++k) {
        const double v = in[(size_t)k * n_ext + base + threadIdx.x];
        tile[threadIdx.x * pitch + k] = RECIP ? zeta[k] / fmax(floor_, v) : v;
      }
    __syncthreads();
    for (int t = threadIdx.x; t < cnt * nz; t += 128) out[(size_t)base * nz + t] = tile[(t / nz) * pitch + t % nz];
    __syncthreads();
  }
}

struct VertexArgs {
  int n, off, n_ext, nb_ext, nz, vi1, nV;
  const int* ptr; const int* ind;                      // b->a pattern (M_map_b_a = M_ddx_b_a = M_ddy_b_a)
  const double* w_map; const double* w_ddx; const double* w_ddy;
  const double* zeta; const double* Vy;
  const double* u; const double* v; const double* u_base; const double* v_base; const double* tau_bx; const double* tau_by;
  const double* eta_3D_b; const double* rzT;     // rzT != nullptr: the interleaved zeta / eta copy is gathered instead of eta_3D_b
  const double* Hi; const double* A_flow; const double* A_flow_vav;
  const double* tauc; const double* beta_sq; const double* alpha_sq; const double* neff;
  double* du_dx; double* du_dy; double* dv_dx; double* dv_dy; double* du_dz_3D; double* dv_dz_3D;
  double* eta_3D_a; double* eta_vav; double* N_a; double* F1_3D_a; double* F2_3D_a; double* beta_a; double* beta_eff_a;
};

// DIVA: calc_horizontal_strain_rates, calc_vertical_shear_strain_rates, calc_effective_viscosity (a-grid part),
// calc_F_integrals (a-grid part), calc_effective_basal_friction_coefficient (a-grid part)
// (DIVA_solver_infinite_slab.f90:138-151, 235-457)
// NZ > 0: the number of layers is a compile-time constant, every layer loop unrolls and the three per-column arrays live in registers (168
// registers per thread); NZ = 0: run-time nz, the arrays are local memory (128 registers + 784 B of stack).  Same operations, same order.
// Measured at 4 M vertices: 6.4 ms for NZ = 12 against 5.8 ms for NZ = 0 -- the kernel is bound by the latency of the triangle-column
// gathers, and the register version runs fewer warps per SM; so NZ = 0 is what runs (UFE_B200_VERTEX_NZ12 selects the other for profiling).
// SMEM: the three per-column arrays live in shared memory ([k][thread], conflict-free) instead of local memory: the ncu capture of round 2
// (profiles/r02_step_kernels_4M_ncu.md) showed 864 B of stack per thread going through L1 to DRAM -- 4.1 GB written and 6.3 GB read per launch
// against 2.6 GB of algorithmic traffic.
template <bool SMEM, int CAP>
struct ColumnArray {
  double loc[SMEM ? 1 : CAP];
  double* sh;
  __device__ __forceinline__ double& operator[](int k) { return SMEM ? sh[k * 128] : loc[k]; }
};
template <int NZ, int MB = 6, bool SMEM = false>
__global__ void __launch_bounds__(128, MB) k_vertex_diva(VertexArgs a, VelPar p) {
  constexpr int kCap = NZ > 0 ? NZ : kMaxNz;
  extern __shared__ double s_col[];   // SMEM: [3][nz][128]
  ColumnArray<SMEM, kCap> dzu, dzv, eta;
  if (SMEM) { dzu.sh = s_col + threadIdx.x; dzv.sh = dzu.sh + a.nz * 128; eta.sh = dzv.sh + a.nz * 128; }
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < a.n; r += (long long)gridDim.x * blockDim.x) {
    const size_t i = a.off + r;
    const int nz = NZ > 0 ? NZ : a.nz;
    double du_dx = 0, du_dy = 0, dv_dx = 0, dv_dy = 0, ub = 0, vb = 0;
#pragma unroll
    for (int k = 0; k < nz; ++k) { dzu[k] = 0.0; dzv[k] = 0.0; }
    for (int q = a.ptr[r]; q < a.ptr[r + 1]; ++q) {
      const int tj = a.ind[q];
      const double wm = a.w_map[q], wx = a.w_ddx[q], wy = a.w_ddy[q];
      const double uj = a.u[tj], vj = a.v[tj];
      du_dx += wx * uj; du_dy += wy * uj; dv_dx += wx * vj; dv_dy += wy * vj;
      ub += wm * a.u_base[tj]; vb += wm * a.v_base[tj];
      const double tbx = a.tau_bx[tj], tby = a.tau_by[tj];
      if (a.rzT) {
        const double* __restrict__ rz = a.rzT + (size_t)tj * nz;     // zeta_k / eta of this triangle's column: contiguous
#pragma unroll
        for (int k = 0; k < nz; ++k) {
          const double q_k = rz[k];
          dzu[k] += wm * (tbx * q_k);
          dzv[k] += wm * (tby * q_k);
        }
      } else {
#pragma unroll
        for (int k = 0; k < nz; ++k) {
          const double e = fmax(p.visc_eff_min, a.eta_3D_b[(size_t)k * a.nb_ext + tj]);
          dzu[k] += wm * (tbx * a.zeta[k] / e);
          dzv[k] += wm * (tby * a.zeta[k] / e);
        }
      }
    }
    const double hs = du_dx * du_dx + dv_dy * dv_dy + du_dx * dv_dy + 0.25 * (du_dy + dv_dx) * (du_dy + dv_dx);
#pragma unroll
    for (int k = 0; k < nz; ++k) {
      const double eps_sq = hs + 0.25 * (dzu[k] * dzu[k] + dzv[k] * dzv[k]) + p.eps0;
      const double e = fmin(fmax(glen_eta(a.A_flow[(size_t)k * a.n_ext + i], eps_sq, p.n_glen), p.visc_eff_min), p.eta_max);
      eta[k] = e;
      a.eta_3D_a[(size_t)k * a.n_ext + i] = e;
    }
    double vav = 0.0;
#pragma unroll
    for (int k = 0; k < nz - 1; ++k) vav = vav + 0.5 * (eta[k + 1] + eta[k]) * (a.zeta[k + 1] - a.zeta[k]);
    const double Hi = a.Hi[i];
    a.N_a[i] = vav * fmax((double)0.1f, Hi);   // max( 0.1, Hi) with a default-real literal (DIVA_solver_infinite_slab.f90:341)
    // F-integrals, bottom-up trapezoid (mesh_zeta.f90:175-206)
    const double Hc = -fmax(0.1, Hi);
    double i1 = 0.0, i2 = 0.0;
    a.F1_3D_a[(size_t)(nz - 1) * a.n_ext + i] = Hc * 0.0;
    a.F2_3D_a[(size_t)(nz - 1) * a.n_ext + i] = Hc * 0.0;
#pragma unroll
    for (int k = nz - 2; k >= 0; --k) {
      const double z0 = a.zeta[k], z1 = a.zeta[k + 1];
      i1 = i1 - 0.5 * ((z1 / eta[k + 1]) + (z0 / eta[k])) * (z1 - z0);
      i2 = i2 - 0.5 * ((z1 * z1 / eta[k + 1]) + (z0 * z0 / eta[k])) * (z1 - z0);
      a.F1_3D_a[(size_t)k * a.n_ext + i] = Hc * i1;
      a.F2_3D_a[(size_t)k * a.n_ext + i] = Hc * i2;
    }
    const double F2_surf = Hc * i2;
    const double beta = sliding_beta(p, ub, vb, a.tauc[i], a.beta_sq[i], a.alpha_sq[i], a.neff[i], a.Vy[a.vi1 - 1 + r]);
    a.beta_a[i] = beta;
    a.beta_eff_a[i] = (p.sliding_law == UFE_SLID_NO_SLIDING) ? 1.0 / F2_surf : beta / (1.0 + beta * F2_surf);
    // DIVA%du_dx_a ... DIVA%du_dz_3D_a are results, not diagnostics: set_ice_velocities_to_DIVA_results copies them into ice%du_dx_3D ... after every
    // solve (conservation_of_momentum_main.f90:472-511), so they are stored whatever cfg->write_diagnostics says
    a.du_dx[i] = du_dx; a.du_dy[i] = du_dy; a.dv_dx[i] = dv_dx; a.dv_dy[i] = dv_dy;
#pragma unroll
    for (int k = 0; k < nz; ++k) { a.du_dz_3D[(size_t)k * a.n_ext + i] = dzu[k]; a.dv_dz_3D[(size_t)k * a.n_ext + i] = dzv[k]; }
    if (p.diagnostics) a.eta_vav[i] = vav;
  }
}

// SSA: calc_horizontal_strain_rates, calc_effective_viscosity, calc_applied_basal_friction_coefficient (a-grid part)
// (SSA_main.f90:150-165, 342-445)
__global__ void __launch_bounds__(128) k_vertex_ssa(VertexArgs a, VelPar p) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < a.n; r += (long long)gridDim.x * blockDim.x) {
    const size_t i = a.off + r;
    double du_dx = 0, du_dy = 0, dv_dx = 0, dv_dy = 0, ua = 0, va = 0;
    for (int q = a.ptr[r]; q < a.ptr[r + 1]; ++q) {
      const int tj = a.ind[q];
      const double uj = a.u[tj], vj = a.v[tj];
      du_dx += a.w_ddx[q] * uj; du_dy += a.w_ddy[q] * uj; dv_dx += a.w_ddx[q] * vj; dv_dy += a.w_ddy[q] * vj;
      ua += a.w_map[q] * uj; va += a.w_map[q] * vj;
    }
    const double eps_sq = du_dx * du_dx + dv_dy * dv_dy + du_dx * dv_dy + 0.25 * (du_dy + dv_dx) * (du_dy + dv_dx) + p.eps0;
    const double e = fmin(fmax(glen_eta(a.A_flow_vav[i], eps_sq, p.n_glen), p.visc_eff_min), p.eta_max);
    a.eta_vav[i] = e;
    a.N_a[i] = e * fmax(0.1, a.Hi[i]);
    a.beta_a[i] = sliding_beta(p, ua, va, a.tauc[i], a.beta_sq[i], a.alpha_sq[i], a.neff[i], a.Vy[a.vi1 - 1 + r]);
    a.du_dx[i] = du_dx; a.du_dy[i] = du_dy; a.dv_dx[i] = dv_dx; a.dv_dy[i] = dv_dy;   // SSA%du_dx_a ...: results (set_ice_velocities_to_SSA_results)
  }
}

// ---- stiffness matrix: pattern (once per mesh + BC choice) and values (every viscosity iteration) -------------------
struct StiffArgs {
  int n, off, ti1, b_lo, nTri;                 // owned triangles, offset of the first owned triangle in extended b-vectors
  const int* TriBI; const int* TriC;           // global mesh arrays
  const int* bc_mask; const double* bc_u; const double* bc_v;   // extended b-vectors
  const int* m2_ptr; const int* m2_ind;        // pattern shared by the five M2_*_b_b operators (columns extended-local)
  const double* m2_ddx; const double* m2_ddy; const double* m2_dxx; const double* m2_dxy; const double* m2_dyy;
  int* A_ptr; int* A_ind; double* A_val; int* row_kind; double* bb;
  // when the blocked sliced-ELL copy exists the assembly writes it in the same pass (blk.cu: map = CSR position -> slot), which saves the
  // separate scatter pass over both formats (6 ms per viscosity iteration at 4 M vertices)
  double* blk_val; const int* blk_map; const int* blk_slice_ptr;
};

__device__ __forceinline__ void put(const StiffArgs& s, int k, double v) {
  if (s.A_val) s.A_val[k] = v;
  if (s.blk_map) s.blk_val[s.blk_map[k]] = v;
}

__device__ __forceinline__ int bc_side(int tribi) { return (tribi <= 2) ? 0 : (tribi <= 4) ? 1 : (tribi <= 6) ? 2 : 3; }  // N, E, S, W

// row classification + length (solve_linearised_SSA_DIVA_infinite_slab.f90:89-151)
__global__ void k_stiff_count(StiffArgs s, VelPar p) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < s.n; r += (long long)gridDim.x * blockDim.x) {
    const int ti = s.ti1 + (int)r;
    for (int uv = 0; uv < 2; ++uv) {
      int kind, len;
      if (s.bc_mask[s.off + r] == 1) { kind = 1; len = 1; }
      else if (s.TriBI[ti - 1] > 0) {
        const int side = bc_side(s.TriBI[ti - 1]);
        const int choice = (uv == 0) ? p.BC_u[side] : p.BC_v[side];
        if (choice == UFE_BC_INFINITE) {
          int nn = 0;
          for (int n = 0; n < 3; ++n) if (s.TriC[(size_t)n * s.nTri + ti - 1] != 0) ++nn;
          kind = 2; len = nn + 1;
        } else { kind = 1; len = 1; }
      } else { kind = 0; len = 2 * (s.m2_ptr[r + 1] - s.m2_ptr[r]); }
      s.row_kind[2 * r + uv] = kind;
      s.A_ptr[2 * r + uv] = len;
    }
  }
}

// column indices in the reference's construction order (:262-287 free rows, :519-529 'infinite', diagonal-only otherwise)
__global__ void k_stiff_fill_ind(StiffArgs s) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < s.n; r += (long long)gridDim.x * blockDim.x) {
    const int ti = s.ti1 + (int)r;
    for (int uv = 0; uv < 2; ++uv) {
      const int row = 2 * (int)r + uv;
      int k = s.A_ptr[row];
      const int kind = s.row_kind[row];
      const int diag = 2 * (s.off + (int)r) + uv;
      if (kind == 1) s.A_ind[k] = diag;
      else if (kind == 2) {
        for (int n = 0; n < 3; ++n) {
          const int tj = s.TriC[(size_t)n * s.nTri + ti - 1];
          if (tj != 0) s.A_ind[k++] = 2 * (tj - s.b_lo) + uv;
        }
        s.A_ind[k] = diag;
      } else {
        for (int q = s.m2_ptr[r]; q < s.m2_ptr[r + 1]; ++q) {
          const int tj = s.m2_ind[q];
          s.A_ind[k++] = 2 * tj;
          s.A_ind[k++] = 2 * tj + 1;
        }
      }
    }
  }
}

// values of the two rows of one triangle + load vector
// (calc_SSA_DIVA_stiffness_matrix_row_free :180-329, ..._sans_... :331-479, ..._row_BC :481-641)
__device__ __forceinline__ void assemble_triangle(const StiffArgs& s, const VelPar& p, int r, double N, double dNx, double dNy, double beta,
                                                  double tdx, double tdy) {
  const int self = s.off + r;
  for (int uv = 0; uv < 2; ++uv) {
    const int row = 2 * r + uv;
    int k = s.A_ptr[row];
    const int kind = s.row_kind[row];
    if (kind == 1) {
      put(s, k, 1.0);
      s.bb[row] = (s.bc_mask[self] == 1) ? ((uv == 0) ? s.bc_u[self] : s.bc_v[self]) : 0.0;
    } else if (kind == 2) {
      const int kend = s.A_ptr[row + 1];
      const int nn = kend - k - 1;
      for (; k < kend - 1; ++k) put(s, k, 1.0);
      put(s, k, -1.0 * (double)nn);
      s.bb[row] = 0.0;
    } else {
      // both rows of the triangle free (almost all): the blocked copy lists the neighbours in the order of this loop (blk.cu: order of first
      // appearance), so the slot follows from q alone and the 32 triangles of a slice write 32 consecutive doubles -- no map lookup
      const bool direct = s.blk_map && s.row_kind[2 * r] == 0 && s.row_kind[2 * r + 1] == 0;
      double* bslot = direct ? s.blk_val + ((size_t)s.blk_slice_ptr[r >> 5] * 4 + uv * 2) * 32 + (r & 31) : nullptr;
      for (int q = s.m2_ptr[r]; q < s.m2_ptr[r + 1]; ++q) {
        const bool is_self = (s.m2_ind[q] == self);
        const double ddx = s.m2_ddx[q], ddy = s.m2_ddy[q], dxx = s.m2_dxx[q], dxy = s.m2_dxy[q], dyy = s.m2_dyy[q];
        double Au, Av;
        if (p.crossterms) {
          if (uv == 0) {
            Au = 4.0 * N * dxx + 4.0 * dNx * ddx + N * dyy + dNy * ddy;
            if (is_self) Au = Au - beta;
            Av = 3.0 * N * dxy + 2.0 * dNx * ddy + dNy * ddx;
          } else {
            Av = 4.0 * N * dyy + 4.0 * dNy * ddy + N * dxx + dNx * ddx;
            if (is_self) Av = Av - beta;
            Au = 3.0 * N * dxy + 2.0 * dNy * ddx + dNx * ddy;
          }
        } else {
          if (uv == 0) { Au = 4.0 * dxx + dyy; if (is_self) Au = Au - beta / N; Av = 3.0 * dxy; }
          else { Av = 4.0 * dyy + dxx; if (is_self) Av = Av - beta / N; Au = 3.0 * dxy; }
        }
        if (direct) {
          if (s.A_val) { s.A_val[k] = Au; s.A_val[k + 1] = Av; }
          k += 2; bslot[0] = Au; bslot[32] = Av; bslot += 128;
        }
        else { put(s, k++, Au); put(s, k++, Av); }
      }
      const double tau = (uv == 0) ? tdx : tdy;
      s.bb[row] = p.crossterms ? -tau : -tau / N;
    }
  }
}

// ---- copy rows: BC 'periodic_ISMIP-HOM' and 'infinite_SSA_icestream' (calc_SSA_DIVA_stiffness_matrix_row_BC :547-581, :618-636) ----------
// The matrix row is the identity row k_stiff_count already makes of every choice that is not 'infinite'; what differs is the load vector.
__device__ __forceinline__ bool is_copy_choice(int choice) { return choice == UFE_BC_PERIODIC_ISMIP_HOM || choice == UFE_BC_INFINITE_SSA_ICESTREAM; }

// compact list of the owned copy rows (out_row == nullptr: count only)
__global__ void k_copy_bc_list(StiffArgs s, VelPar p, int* __restrict__ count, int* __restrict__ out_row, int* __restrict__ out_choice) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < s.n; r += (long long)gridDim.x * blockDim.x) {
    const int ti = s.ti1 + (int)r;
    const int bi = s.TriBI[ti - 1];
    if (bi <= 0 || s.bc_mask[s.off + r] == 1) continue;
    const int side = bc_side(bi);
    for (int uv = 0; uv < 2; ++uv) {
      const int choice = (uv == 0) ? p.BC_u[side] : p.BC_v[side];
      if (!is_copy_choice(choice)) continue;
      const int e = atomicAdd(count, 1);
      if (out_row) { out_row[e] = 2 * (int)r + uv; out_choice[e] = choice; }
    }
  }
}

// the point the previous solution is copied from (find_ti_copy_ISMIP_HOM_periodic, mesh_utilities.f90:2645-2658;
// find_ti_copy_SSA_icestream_infinite, :2704-2710)
__device__ __forceinline__ void copy_bc_point(int choice, double gcx, double gcy, double L, double xmin, double xmax, double& px, double& py) {
  if (choice == UFE_BC_PERIODIC_ISMIP_HOM) {
    px = (gcx > 0.0) ? gcx - L / 2.0 : gcx + L / 2.0;
    py = (gcy > 0.0) ? gcy - L / 2.0 : gcy + L / 2.0;
  } else {
    py = gcy;
    px = (gcx < 0.0) ? xmin + (xmax - xmin) * 1.0 / 3.0 : xmin + (xmax - xmin) * 2.0 / 3.0;
  }
}

// find_containing_vertex (mesh_utilities.f90:1368-1416) from the reference's start vertex 5: one greedy walk per copy row, once per pattern
__global__ void k_copy_bc_find(int n_copy, const int* __restrict__ row, const int* __restrict__ choice, int* __restrict__ out_vi, int ti1,
                               int nV, int nTri, const double* __restrict__ V, const int* __restrict__ nC, const int* __restrict__ Cn,
                               const double* __restrict__ TriGC, double L, double xmin, double xmax) {
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < n_copy; e += gridDim.x * blockDim.x) {
    const int ti = ti1 + (row[e] >> 1);
    double px, py;
    copy_bc_point(choice[e], TriGC[ti - 1], TriGC[(size_t)nTri + ti - 1], L, xmin, xmax, px, py);
    int vi = 5, vi_prev = 5;
    for (;;) {
      double dx = V[vi - 1] - px, dy = V[(size_t)nV + vi - 1] - py;
      const double d = sqrt(dx * dx + dy * dy);
      double dcmin = d + 10.0; int vcmin = 0;
      const int nc = nC[vi - 1];
      for (int ci = 0; ci < nc; ++ci) {
        const int vc = Cn[(size_t)ci * nV + vi - 1];
        if (vc == vi_prev) continue;
        dx = V[vc - 1] - px; dy = V[(size_t)nV + vc - 1] - py;
        const double dc = sqrt(dx * dx + dy * dy);
        if (dc < dcmin) { dcmin = dc; vcmin = vc; }
      }
      if (dcmin < d) { vi_prev = vi; vi = vcmin; } else break;
    }
    out_vi[e] = vi;
  }
}

// nranks > 1: own slice of the previous solution into the global vectors (zero elsewhere; completed by one all-reduce)
__global__ void k_copy_bc_scatter(int n, int off, int ti1, int nTri, const double* __restrict__ u, const double* __restrict__ v, double* __restrict__ g) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    g[ti1 - 1 + r] = u[off + r]; g[(size_t)nTri + ti1 - 1 + r] = v[off + r];
  }
}

// load vector of the copy rows: inverse-square-distance weights over the triangles of the containing vertex (normalised first, as the
// reference does), then the relaxation with C%visc_it_relax.  src_u/src_v[tj - src_lo] = previous solution on triangle tj.
__global__ void k_copy_bc_rhs(int n_copy, const int* __restrict__ row, const int* __restrict__ choice, const int* __restrict__ cvi, int ti1,
                              int nV, int nTri, const int* __restrict__ niTri, const int* __restrict__ iTri, const double* __restrict__ TriGC,
                              double L, double xmin, double xmax, double relax, const double* __restrict__ src_u,
                              const double* __restrict__ src_v, int src_lo, double* __restrict__ bb) {
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < n_copy; e += gridDim.x * blockDim.x) {
    const int rw = row[e], ti = ti1 + (rw >> 1), vi = cvi[e];
    const double* __restrict__ prev = (rw & 1) ? src_v : src_u;
    double px, py;
    copy_bc_point(choice[e], TriGC[ti - 1], TriGC[(size_t)nTri + ti - 1], L, xmin, xmax, px, py);
    const int nt = niTri[vi - 1];
    double wsum = 0.0;
    for (int it = 0; it < nt; ++it) {
      const int tj = iTri[(size_t)it * nV + vi - 1];
      const double dx = px - TriGC[tj - 1], dy = py - TriGC[(size_t)nTri + tj - 1];
      const double dist = sqrt(dx * dx + dy * dy);
      wsum += 1.0 / (dist * dist);
    }
    double fixed = 0.0;
    for (int it = 0; it < nt; ++it) {
      const int tj = iTri[(size_t)it * nV + vi - 1];
      const double dx = px - TriGC[tj - 1], dy = py - TriGC[(size_t)nTri + tj - 1];
      const double dist = sqrt(dx * dx + dy * dy);
      fixed = fixed + ((1.0 / (dist * dist)) / wsum) * prev[tj - src_lo];
    }
    bb[rw] = (relax * fixed) + ((1.0 - relax) * prev[ti - src_lo]);
  }
}

// standalone assembly from given b-grid coefficients (the solve_SSA_DIVA_linearised entry point)
__global__ void __launch_bounds__(128) k_assemble(StiffArgs s, VelPar p, const double* __restrict__ N_b, const double* __restrict__ dNx,
                                                  const double* __restrict__ dNy, const double* __restrict__ beta,
                                                  const double* __restrict__ tdx, const double* __restrict__ tdy) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < s.n; r += (long long)gridDim.x * blockDim.x) {
    const int i = s.off + (int)r;
    assemble_triangle(s, p, (int)r, N_b[i], dNx[i], dNy[i], beta[i], tdx[i], tdy[i]);
  }
}

// ---- b-grid (triangle) kernel: a->b gathers fused with the stiffness assembly -----------------------------------------
struct TriArgs {
  int na_ext, nb_ext, nz, final_pass;
  const int* ptr; const int* ind;                // a->b pattern (M_map_a_b = M_ddx_a_b = M_ddy_a_b), 3 entries per row
  const double* w_map; const double* w_ddx; const double* w_ddy;
  const double* N_a; const double* beta_a; const double* beta_eff_a; const double* eta_3D_a; const double* F1_3D_a; const double* F2_3D_a;
  const double* etaT;                            // != nullptr: layer-interleaved copy of eta_3D_a
  const double* fraction_gr_b; const double* tau_dx; const double* tau_dy;
  double* N_b; double* dN_dx_b; double* dN_dy_b; double* beta_b; double* beta_eff_b; double* eta_3D_b; double* F1_3D_b; double* F2_3D_b;
};

// DIVA: map_a_b_3D(eta), map/ddx/ddy_a_b_2D(N), map_a_b_3D(F1,F2), map_a_b_2D(beta, beta_eff), sub-grid friction
// (DIVA_solver_infinite_slab.f90:327-348, 391-393, 445-455) + the assembly of solve_SSA_DIVA_linearised
template <bool DIVA>
__global__ void __launch_bounds__(128) k_tri_assemble(TriArgs t, StiffArgs s, VelPar p) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < s.n; r += (long long)gridDim.x * blockDim.x) {
    const size_t i = s.off + r;
    double N = 0, dNx = 0, dNy = 0, beta = 0, beta_eff = 0;
    const int q1 = t.ptr[r], q2 = t.ptr[r + 1];
    for (int q = q1; q < q2; ++q) {
      const int vj = t.ind[q];
      const double Na = t.N_a[vj];
      N += t.w_map[q] * Na; dNx += t.w_ddx[q] * Na; dNy += t.w_ddy[q] * Na;
      beta += t.w_map[q] * t.beta_a[vj];
      if (DIVA) beta_eff += t.w_map[q] * t.beta_eff_a[vj];
    }
    if (DIVA) {
      for (int k = 0; k < t.nz; ++k) {
        double e = 0, f2 = 0, f1 = 0;
        const bool need_f = (k == 0) || t.final_pass || p.diagnostics;
        for (int q = q1; q < q2; ++q) {
          const size_t j = (size_t)k * t.na_ext + t.ind[q];
          e += t.w_map[q] * (t.etaT ? t.etaT[(size_t)t.ind[q] * t.nz + k] : t.eta_3D_a[j]);
          if (need_f) { f2 += t.w_map[q] * t.F2_3D_a[j]; f1 += t.w_map[q] * t.F1_3D_a[j]; }
        }
        t.eta_3D_b[(size_t)k * t.nb_ext + i] = e;
        if (need_f) { t.F2_3D_b[(size_t)k * t.nb_ext + i] = f2; t.F1_3D_b[(size_t)k * t.nb_ext + i] = f1; }
      }
    }
    double applied;  // the friction coefficient that enters the stiffness matrix
    if (DIVA) {
      if (p.do_subgrid) beta_eff = beta_eff * pow(t.fraction_gr_b[i], p.subgrid_exp);
      applied = beta_eff;
      t.beta_eff_b[i] = beta_eff;
    } else {
      if (p.do_subgrid) beta = beta * pow(t.fraction_gr_b[i], p.subgrid_exp);
      applied = beta;
    }
    t.beta_b[i] = beta;
    t.N_b[i] = N; t.dN_dx_b[i] = dNx; t.dN_dy_b[i] = dNy;
    if (!t.final_pass) assemble_triangle(s, p, (int)r, N, dNx, dNy, applied, t.tau_dx[i], t.tau_dy[i]);
  }
}

// ---- epilogue of a viscosity iteration -------------------------------------------------------------------------------
// apply_velocity_limits, relax_viscosity_iterations, calc_L2_norm_uv (SSA_DIVA_utilities.f90:96-206);
// calc_basal_velocities, calc_basal_shear_stress (DIVA_solver_infinite_slab.f90:459-523)
template <bool DIVA>
__global__ void k_epilogue(int n, int off, int nb_ext, const double* __restrict__ uv_own, double* __restrict__ u, double* __restrict__ v,
                           const double* __restrict__ u_prev, const double* __restrict__ v_prev, const double* __restrict__ beta_b,
                           const double* __restrict__ beta_eff_b, const double* __restrict__ F2_3D_b, double* __restrict__ u_base,
                           double* __restrict__ v_base, double* __restrict__ tau_bx, double* __restrict__ tau_by, VelPar p,
                           double* partials, unsigned* counter, double* red) {
  double acc[2] = {0.0, 0.0};
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const size_t i = off + r;
    double un = uv_own[2 * r], vn = uv_own[2 * r + 1];
    const double uabs = sqrt(un * un + vn * vn);
    if (uabs > p.vel_max) { un = un * p.vel_max / uabs; vn = vn * p.vel_max / uabs; }
    const double up = u_prev[i], vp = v_prev[i];
    un = (p.relax * un) + ((1.0 - p.relax) * up);
    vn = (p.relax * vn) + ((1.0 - p.relax) * vp);
    u[i] = un; v[i] = vn;
    if (DIVA) {
      if (p.sliding_law == UFE_SLID_NO_SLIDING) { u_base[i] = 0.0; v_base[i] = 0.0; }
      else {
        const double den = 1.0 + beta_b[i] * F2_3D_b[i];   // layer 1 (surface) of F2_3D_b
        u_base[i] = un / den; v_base[i] = vn / den;
      }
      tau_bx[i] = un * beta_eff_b[i];
      tau_by[i] = vn * beta_eff_b[i];
    }
    if (un == un && vn == vn) {
      acc[0] += (un - up) * (un - up); acc[0] += (vn - vp) * (vn - vp);
      acc[1] += (un + up) * (un + up); acc[1] += (vn + vp) * (vn + vp);
    }
  }
  (void)nb_ext;
  grid_sum<2>(acc, partials, counter, red);
}

// u_prev = u, v_prev = v (gather_to_all( u_b, u_b_prev), solve_linearised_SSA_DIVA_infinite_slab.f90:54-55)
__global__ void k_store_prev(int n, int off, const double* __restrict__ u, const double* __restrict__ v, double* __restrict__ up,
                             double* __restrict__ vp) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    up[off + r] = u[off + r]; vp[off + r] = v[off + r];
  }
}

__global__ void k_interleave(int n, int off, const double* __restrict__ u, const double* __restrict__ v, double* __restrict__ uv_own) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    uv_own[2 * r] = u[off + r]; uv_own[2 * r + 1] = v[off + r];
  }
}

__global__ void k_deinterleave(int n, int off, const double* __restrict__ uv_own, double* __restrict__ u, double* __restrict__ v) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    u[off + r] = uv_own[2 * r]; v[off + r] = uv_own[2 * r + 1];
  }
}

// calc_3D_velocities (DIVA_solver_infinite_slab.f90:525-568)
__global__ void k_3D_velocities(int n, int off, int nb_ext, int nz, const double* __restrict__ u_base, const double* __restrict__ v_base,
                                const double* __restrict__ tau_bx, const double* __restrict__ tau_by, const double* __restrict__ beta_b,
                                const double* __restrict__ F1_3D_b, double* __restrict__ u_3D, double* __restrict__ v_3D, int no_sliding) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const size_t i = off + r;
    for (int k = 0; k < nz; ++k) {
      const size_t j = (size_t)k * nb_ext + i;
      if (no_sliding) { u_3D[j] = tau_bx[i] * F1_3D_b[j]; v_3D[j] = tau_by[i] * F1_3D_b[j]; }
      else {
        u_3D[j] = u_base[i] * (1.0 + beta_b[i] * F1_3D_b[j]);
        v_3D[j] = v_base[i] * (1.0 + beta_b[i] * F1_3D_b[j]);
      }
    }
  }
}

// ---- host side ---------------------------------------------------------------------------------------------------------

int ensure_velocity(Ctx* c) {
  UFE_CHECK(c->mesh, "no mesh: call ufe_b200_mesh_create first");
  UFE_CHECK(c->mesh->have_ops, "no mesh operators: call ufe_b200_calc_matrix_operators first");
  if (!c->vel) {
    c->vel = new Velocity();
    int rc = c->vel->alloc(c);
    if (rc) { delete c->vel; c->vel = nullptr; return rc; }
  }
  return 0;
}

// skip_csr: only the blocked copy is wanted (FP64 multigrid solve); the CSR values are re-formed on demand (ensure_csr_values)
static StiffArgs stiff_args(Ctx* c, bool skip_csr = false) {
  Mesh& M = *c->mesh; Velocity& V = *c->vel;
  StiffArgs s;
  s.n = M.b.n_loc(); s.off = M.b.off(); s.ti1 = M.b.i1; s.b_lo = M.b.lo; s.nTri = M.nTri;
  s.TriBI = M.TriBI.p; s.TriC = M.TriC.p; s.bc_mask = V.bc_mask.p; s.bc_u = V.bc_u.p; s.bc_v = V.bc_v.p;
  const Mesh::OpGroup& g = M.grp[Mesh::G_M2];
  s.m2_ptr = g.ptr.p; s.m2_ind = g.ind.p;
  s.m2_ddx = g.val[0].p; s.m2_ddy = g.val[1].p; s.m2_dxx = g.val[2].p; s.m2_dxy = g.val[3].p; s.m2_dyy = g.val[4].p;
  s.A_ptr = V.A.ptr; s.A_ind = V.A.ind; s.A_val = V.A.val; s.row_kind = V.row_kind.p; s.bb = V.bb.p;
  const bool direct = V.blk.valid && V.blk.map.p && V.blk.val.p && !getenv("UFE_B200_BLK_SCATTER");
  s.blk_val = direct ? V.blk.val.p : nullptr; s.blk_map = direct ? V.blk.map.p : nullptr; s.blk_slice_ptr = V.blk.slice_ptr.p;
  if (skip_csr && direct) s.A_val = nullptr;
  return s;
}

// The CSR values are only read by the CSR / Jacobi solves, the FP32 multigrid storage and ufe_b200_get_stiffness; the default FP64 multigrid
// solve streams the blocked copy alone, so the assembly then leaves the (uncoalesced) CSR value writes out.
static bool csr_values_unneeded(Ctx* c, const ufe_b200_config* cfg) {
  return cfg->pc_type == UFE_PC_MG && cfg->stiffness_format != 1 && !c->mg_params.fp32 && !getenv("UFE_B200_BLK_SCATTER");
}
__global__ void k_blk_gather(long long nnz, const int* __restrict__ map, const double* __restrict__ bv, double* __restrict__ v) {
  for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += (long long)gridDim.x * blockDim.x) v[k] = bv[map[k]];
}
static int ensure_csr_values(Ctx* c) {
  Velocity& V = *c->vel;
  if (!V.csr_stale) return 0;
  UFE_LAUNCH(c, TC_ASSEMBLY);
  k_blk_gather<<<grid_for(c, V.A.nnz, 256), 256, 0, c->stream>>>(V.A.nnz, V.blk.map.p, V.blk.val.p, V.A.val);
  UFE_CUDA(cudaGetLastError());
  V.csr_stale = false;
  return 0;
}

// The sparsity pattern is a pure function of the mesh, the BC choices and the prescribed-velocity mask:
// build it once and only rebuild when one of those changes.
static int ensure_pattern(Ctx* c, const VelPar& p, unsigned long long bc_version) {
  Mesh& M = *c->mesh; Velocity& V = *c->vel;
  unsigned long long key = 1469598103934665603ull;
  auto mix = [&](unsigned long long x) { key = (key ^ x) * 1099511628211ull; };
  bool any_copy = false;
  for (int k = 0; k < 4; ++k) {
    mix(p.BC_u[k] + 1); mix(p.BC_v[k] + 7);
    // crash('unknown choice_BC_u/v ...') in the reference (solve_linearised_SSA_DIVA_infinite_slab.f90:510, 586); 'infinite_SSA_icestream' exists for u only
    UFE_CHECK(p.BC_u[k] >= UFE_BC_INFINITE && p.BC_u[k] <= UFE_BC_INFINITE_SSA_ICESTREAM, "unknown choice_BC_u %d", p.BC_u[k]);
    UFE_CHECK(p.BC_v[k] >= UFE_BC_INFINITE && p.BC_v[k] <= UFE_BC_PERIODIC_ISMIP_HOM, "unknown choice_BC_v %d", p.BC_v[k]);
    any_copy = any_copy || p.BC_u[k] >= UFE_BC_PERIODIC_ISMIP_HOM || p.BC_v[k] >= UFE_BC_PERIODIC_ISMIP_HOM;
  }
  mix(bc_version);
  if (any_copy) { unsigned long long lb; memcpy(&lb, &p.ismip_hom_L, sizeof(lb)); mix(lb); }
  if (V.pattern_valid && V.pattern_key == key) return 0;
  const int n = M.b.n_loc(), rows = 2 * n;
  UFE_TRY(V.A.ptr_buf.alloc((size_t)rows + 1)); UFE_TRY(V.row_kind.alloc(rows));
  V.A.ptr = V.A.ptr_buf.p;
  UFE_CUDA(cudaMemsetAsync(V.A.ptr, 0, sizeof(int) * ((size_t)rows + 1), c->stream));
  StiffArgs s = stiff_args(c);
  const int grid = grid_for(c, n, 128, 16);
  { UFE_LAUNCH(c, TC_ASSEMBLY); k_stiff_count<<<grid, 128, 0, c->stream>>>(s, p); }
  size_t tmp_bytes = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, V.A.ptr, V.A.ptr, rows + 1, c->stream);
  DBuf<char> tmp;
  UFE_TRY(tmp.alloc(tmp_bytes));
  c->launches++;
  UFE_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, tmp_bytes, V.A.ptr, V.A.ptr, rows + 1, c->stream));
  int nnz = 0;
  UFE_CUDA(cudaMemcpyAsync(&nnz, V.A.ptr + rows, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  UFE_TRY(V.A.ind_buf.alloc(nnz)); UFE_TRY(V.A.val_buf.alloc(nnz));
  V.A.ind = V.A.ind_buf.p; V.A.val = V.A.val_buf.p;
  V.A.m_loc = rows; V.A.nnz = nnz; V.A.rows = M.buv; V.A.cols = M.buv;
  const double avg = rows > 0 ? (double)nnz / rows : 0.0;
  V.A.vec_width = avg <= 6.0 ? 4 : avg <= 14.0 ? 8 : 16;
  V.A.valid = true;
  s = stiff_args(c);
  { UFE_LAUNCH(c, TC_ASSEMBLY); k_stiff_fill_ind<<<grid, 128, 0, c->stream>>>(s); }
  UFE_CUDA(cudaGetLastError());
  UFE_TRY(blk_build(c, V.blk, V.A, n, M.b.off()));
  V.mg.valid = false;
  V.n_copy = 0; V.copy_any = any_copy;
  if (any_copy) {   // the copy rows and, once per pattern, the vertex each one copies around
    UFE_CHECK(M.nV >= 5, "copy boundary conditions need a mesh of at least 5 vertices");
    UFE_TRY(V.copy_count.alloc(1));
    UFE_CUDA(cudaMemsetAsync(V.copy_count.p, 0, sizeof(int), c->stream));
    { UFE_LAUNCH(c, TC_ASSEMBLY); k_copy_bc_list<<<grid, 128, 0, c->stream>>>(s, p, V.copy_count.p, nullptr, nullptr); }
    int nc = 0;
    UFE_CUDA(cudaMemcpyAsync(&nc, V.copy_count.p, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    UFE_CUDA(cudaStreamSynchronize(c->stream));
    if (nc > 0) {
      UFE_TRY(V.copy_row.alloc(nc)); UFE_TRY(V.copy_choice.alloc(nc)); UFE_TRY(V.copy_vi.alloc(nc));
      UFE_CUDA(cudaMemsetAsync(V.copy_count.p, 0, sizeof(int), c->stream));
      { UFE_LAUNCH(c, TC_ASSEMBLY); k_copy_bc_list<<<grid, 128, 0, c->stream>>>(s, p, V.copy_count.p, V.copy_row.p, V.copy_choice.p); }
      UFE_LAUNCH(c, TC_ASSEMBLY);
      k_copy_bc_find<<<grid_for(c, nc, 128), 128, 0, c->stream>>>(nc, V.copy_row.p, V.copy_choice.p, V.copy_vi.p, M.b.i1, M.nV, M.nTri, M.V.p,
                                                                  M.nC.p, M.C.p, M.TriGC.p, p.ismip_hom_L, M.xmin, M.xmax);
      UFE_CUDA(cudaGetLastError());
    }
    V.n_copy = nc;
    if (c->nranks > 1) UFE_TRY(V.uv_glob.alloc(2 * (size_t)M.nTri));
  }
  V.pattern_key = key; V.pattern_valid = true;
  return 0;
}

// Load vector of the copy rows from the previous solution (= V.u, V.v at assembly time); call after the assembly kernel.
// Collective when nranks > 1 (every rank of a pattern with copy choices calls it, whether it owns copy rows or not).
static int copy_bc_rhs(Ctx* c, const VelPar& p) {
  Mesh& M = *c->mesh; Velocity& V = *c->vel;
  if (!V.copy_any) return 0;
  const double* src_u = V.u.p; const double* src_v = V.v.p; int src_lo = M.b.lo;
  if (c->nranks > 1) {
    UFE_CUDA(cudaMemsetAsync(V.uv_glob.p, 0, sizeof(double) * 2 * (size_t)M.nTri, c->stream));
    { UFE_LAUNCH(c, TC_ASSEMBLY);
      k_copy_bc_scatter<<<grid_for(c, M.b.n_loc(), 256), 256, 0, c->stream>>>(M.b.n_loc(), M.b.off(), M.b.i1, M.nTri, V.u.p, V.v.p, V.uv_glob.p); }
    UFE_TRY(allreduce_sum_large(c, V.uv_glob.p, 2 * (size_t)M.nTri, false));
    src_u = V.uv_glob.p; src_v = V.uv_glob.p + M.nTri; src_lo = 1;
  }
  if (V.n_copy > 0) {
    UFE_LAUNCH(c, TC_ASSEMBLY);
    k_copy_bc_rhs<<<grid_for(c, V.n_copy, 128), 128, 0, c->stream>>>(V.n_copy, V.copy_row.p, V.copy_choice.p, V.copy_vi.p, M.b.i1, M.nV, M.nTri,
                                                                    M.niTri.p, M.iTri.p, M.TriGC.p, p.ismip_hom_L, M.xmin, M.xmax, p.relax_cfg,
                                                                    src_u, src_v, src_lo, V.bb.p);
    UFE_CUDA(cudaGetLastError());
  }
  return 0;
}

static int solve_linear_system(Ctx* c, const ufe_b200_config* cfg, double rtol, double abstol, int* its) {
  Velocity& V = *c->vel;
  KrylovParams kp{rtol, abstol, cfg->ksp_maxit > 0 ? cfg->ksp_maxit : 2000, cfg->ksp_type, cfg->pc_type};
  if (cfg->ksp_initial_guess_nonzero) {  // uv_buv( tiuv2n( ti,1:2)) = u_b, v_b (solve_linearised_SSA_DIVA_infinite_slab.f90:74-84)
    Mesh& M = *c->mesh;
    kp.use_guess = 1;
    UFE_LAUNCH(c, TC_EPILOGUE);
    k_interleave<<<grid_for(c, M.b.n_loc(), 128, 16), 128, 0, c->stream>>>(M.b.n_loc(), M.b.off(), V.u.p, V.v.p, V.uv.p + M.buv.off());
  }
  if (cfg->stiffness_format == 1 || kp.pc_type != UFE_PC_MG || c->mg_params.fp32) UFE_TRY(ensure_csr_values(c));   // no-op unless the assembly skipped them
  if (cfg->stiffness_format == 1) return csr_ksp_solve(c, V.A, V.bb.p, V.uv.p, kp, TC_SPMV_STIFFNESS, its);
  UFE_CHECK(kp.ksp_type == UFE_KSP_BICGSTAB, "only BiCGStab is available for the (non-symmetric) stiffness matrix");
  if (kp.pc_type == UFE_PC_MG) {
    Mesh& M = *c->mesh;
    if (!V.mg.valid) { V.mg.p = c->mg_params; UFE_TRY(mg_build(c, V.mg, V.blk, M.b)); }
    if (!V.blk_filled) UFE_TRY(blk_update_values(c, V.blk, V.A));
    UFE_TRY(mg_update_values(c, V.mg, V.blk, V.A));
    return bicgstab_solve_mg(c, V.mg, BlkLauncher{&V.blk, M.b.off()}, V.A.cols, V.bb.p, V.uv.p, kp, its);
  }
  UFE_CHECK(kp.pc_type == UFE_PC_JACOBI || kp.pc_type == UFE_PC_NONE, "stiffness solve supports pc = jacobi | none | multigrid");
  if (!c->krylov) c->krylov = new Krylov();
  UFE_TRY(c->krylov->dinv.alloc(V.A.m_loc));
  UFE_TRY(csr_make_dinv(c, V.A, c->krylov->dinv.p, kp.pc_type));
  if (!V.blk_filled) UFE_TRY(blk_update_values(c, V.blk, V.A));
  return bicgstab_solve_mat(c, BlkLauncher{&V.blk, c->mesh->b.off()}, V.A.cols, V.bb.p, c->krylov->dinv.p, V.uv.p, kp, TC_SPMV_STIFFNESS, its);
}

template <typename T>
static int h2d_field(Ctx* c, T* d_ext, const Range& r, const T* h, int ncols = 1) {
  if (!h) return 0;
  if (r.n_ext() == r.n_loc()) {   // no halo: one contiguous copy
    UFE_CUDA(cudaMemcpyAsync(d_ext, h, sizeof(T) * (size_t)r.n_loc() * ncols, cudaMemcpyHostToDevice, c->stream));
    return 0;
  }
  UFE_CUDA(cudaMemcpy2DAsync(d_ext + r.off(), sizeof(T) * r.n_ext(), h, sizeof(T) * r.n_loc(), sizeof(T) * r.n_loc(), ncols,
                             cudaMemcpyHostToDevice, c->stream));
  return 0;
}
static int d2h_field(Ctx* c, double* h, const double* d_ext, const Range& r, int ncols = 1) {
  if (!h) return 0;
  if (r.n_ext() == r.n_loc()) {
    UFE_CUDA(cudaMemcpyAsync(h, d_ext, sizeof(double) * (size_t)r.n_loc() * ncols, cudaMemcpyDeviceToHost, c->stream));
    return 0;
  }
  UFE_CUDA(cudaMemcpy2DAsync(h, sizeof(double) * r.n_loc(), d_ext + r.off(), sizeof(double) * r.n_ext(), sizeof(double) * r.n_loc(), ncols,
                             cudaMemcpyDeviceToHost, c->stream));
  return 0;
}

static int upload_bc(Ctx* c, const int32_t* mask, const double* bu, const double* bv) {
  Mesh& M = *c->mesh; Velocity& V = *c->vel;
  if (mask || bu || bv) {
    UFE_CHECK(mask && bu && bv, "need to provide prescribed u,v fields and mask!");  // DIVA_solver_infinite_slab.f90:104-107
    UFE_TRY(h2d_field(c, V.bc_mask.p, M.b, mask)); UFE_TRY(h2d_field(c, V.bc_u.p, M.b, bu)); UFE_TRY(h2d_field(c, V.bc_v.p, M.b, bv));
    // The sparsity pattern (and the multigrid hierarchy on it) depends on the mask only, not on the prescribed values: a caller that passes the same
    // mask every time step keeps its pattern.  With several ranks every rank would have to take the same decision (the multigrid set-up holds
    // collectives), so there the pattern is rebuilt whenever a mask is passed.
    const size_t nb = (size_t)M.b.n_loc();
    const bool same = c->nranks == 1 && V.have_bc && V.bc_mask_host.size() == nb && memcmp(V.bc_mask_host.data(), mask, nb * sizeof(int32_t)) == 0;
    V.have_bc = true;
    if (!same) {
      V.bc_mask_host.assign(mask, mask + nb);
      V.pattern_valid = false;
      ++c->bc_version;
    }
  } else if (V.have_bc) {
    UFE_CUDA(cudaMemsetAsync(V.bc_mask.p, 0, sizeof(int) * V.bc_mask.n, c->stream));
    V.have_bc = false; V.pattern_valid = false; ++c->bc_version; V.bc_mask_host.clear();
  }
  return 0;
}

int velocity_upload(Ctx* c, const ufe_b200_ice_inputs* ice, const ufe_b200_velocity_state* st, int flags) {
  UFE_TRY(ensure_velocity(c));
  Mesh& M = *c->mesh; Velocity& V = *c->vel;
  const int nz = M.nz;
  UFE_REGION(c, TC_H2D);
  if (ice) {
    UFE_CHECK(((ice->Hi && ice->Hs) || (flags & VU_SKIP_GEOMETRY)) && ice->A_flow && ice->fraction_gr_b && ice->mask_grounded_ice,
              "ice inputs: Hi, Hs, A_flow, fraction_gr_b, mask_grounded_ice are required");
    if (!(flags & VU_SKIP_GEOMETRY)) { UFE_TRY(h2d_field(c, V.Hi.p, M.a, ice->Hi)); UFE_TRY(h2d_field(c, V.Hs.p, M.a, ice->Hs)); }
    UFE_TRY(h2d_field(c, V.A_flow.p, M.a, ice->A_flow, nz)); UFE_TRY(h2d_field(c, V.fraction_gr_b.p, M.b, ice->fraction_gr_b));
    UFE_TRY(h2d_field(c, V.till_friction_angle.p, M.a, ice->till_friction_angle)); UFE_TRY(h2d_field(c, V.beta_sq.p, M.a, ice->beta_sq));
    UFE_TRY(h2d_field(c, V.alpha_sq.p, M.a, ice->alpha_sq)); UFE_TRY(h2d_field(c, V.effective_pressure.p, M.a, ice->effective_pressure));
    UFE_TRY(h2d_field(c, V.mask_grounded_ice.p, M.a, ice->mask_grounded_ice));
    UFE_TRY(h2d_field(c, V.mask_icefree_land.p, M.a, ice->mask_icefree_land));
    V.have_tfa = ice->till_friction_angle != nullptr; V.have_beta_sq = ice->beta_sq != nullptr;
    V.have_alpha_sq = ice->alpha_sq != nullptr; V.have_neff = ice->effective_pressure != nullptr;
    int any = 0;
    for (int i = 0; i < M.a.n_loc(); ++i) if (ice->mask_grounded_ice[i]) { any = 1; break; }
    V.grounded_ice_exists = any;
    if (c->nranks > 1) {  // MPI_ALLREDUCE( ..., MPI_LOR, ...) (DIVA_solver_infinite_slab.f90:88)
      double f = any;
      UFE_CUDA(cudaMemcpyAsync(V.red.p + 4, &f, sizeof(double), cudaMemcpyHostToDevice, c->stream));
      UFE_TRY(allreduce_sum(c, V.red.p + 4, 1));
      UFE_CUDA(cudaMemcpyAsync(&f, V.red.p + 4, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
      UFE_CUDA(cudaStreamSynchronize(c->stream));
      V.grounded_ice_exists = f > 0.0;
    }
  }
  if (st && (flags & VU_STATE_UV)) {
    UFE_TRY(h2d_field(c, V.u.p, M.b, (const double*)st->u_vav_b)); UFE_TRY(h2d_field(c, V.v.p, M.b, (const double*)st->v_vav_b));
  }
  if (st && (flags & VU_STATE_REST)) {
    UFE_TRY(h2d_field(c, V.u_base.p, M.b, (const double*)st->u_base_b)); UFE_TRY(h2d_field(c, V.v_base.p, M.b, (const double*)st->v_base_b));
    UFE_TRY(h2d_field(c, V.tau_bx.p, M.b, (const double*)st->tau_bx_b)); UFE_TRY(h2d_field(c, V.tau_by.p, M.b, (const double*)st->tau_by_b));
    UFE_TRY(h2d_field(c, V.eta_3D_b.p, M.b, (const double*)st->eta_3D_b, nz));
  }
  V.uploaded = true;
  return 0;
}

int velocity_download(Ctx* c, ufe_b200_velocity_state* st, bool sync) {
  UFE_CHECK(c->vel && c->vel->uploaded, "velocity_download: nothing on the device");
  Mesh& M = *c->mesh; Velocity& V = *c->vel;
  const int nz = M.nz;
  {
  UFE_REGION(c, TC_D2H);
  UFE_TRY(d2h_field(c, st->u_vav_b, V.u.p, M.b)); UFE_TRY(d2h_field(c, st->v_vav_b, V.v.p, M.b));
  UFE_TRY(d2h_field(c, st->u_base_b, V.u_base.p, M.b)); UFE_TRY(d2h_field(c, st->v_base_b, V.v_base.p, M.b));
  UFE_TRY(d2h_field(c, st->tau_bx_b, V.tau_bx.p, M.b)); UFE_TRY(d2h_field(c, st->tau_by_b, V.tau_by.p, M.b));
  UFE_TRY(d2h_field(c, st->eta_3D_b, V.eta_3D_b.p, M.b, nz));
  UFE_TRY(d2h_field(c, st->u_b_prev, V.u_prev.p, M.b)); UFE_TRY(d2h_field(c, st->v_b_prev, V.v_prev.p, M.b));
  UFE_TRY(d2h_field(c, st->tau_dx_b, V.tau_dx.p, M.b)); UFE_TRY(d2h_field(c, st->tau_dy_b, V.tau_dy.p, M.b));
  UFE_TRY(d2h_field(c, st->u_3D_b, V.u_3D.p, M.b, nz)); UFE_TRY(d2h_field(c, st->v_3D_b, V.v_3D.p, M.b, nz));
  UFE_TRY(d2h_field(c, st->du_dx_a, V.du_dx.p, M.a)); UFE_TRY(d2h_field(c, st->du_dy_a, V.du_dy.p, M.a));
  UFE_TRY(d2h_field(c, st->dv_dx_a, V.dv_dx.p, M.a)); UFE_TRY(d2h_field(c, st->dv_dy_a, V.dv_dy.p, M.a));
  UFE_TRY(d2h_field(c, st->du_dz_3D_a, V.du_dz_3D.p, M.a, nz)); UFE_TRY(d2h_field(c, st->dv_dz_3D_a, V.dv_dz_3D.p, M.a, nz));
  UFE_TRY(d2h_field(c, st->eta_3D_a, V.eta_3D_a.p, M.a, nz)); UFE_TRY(d2h_field(c, st->eta_vav_a, V.eta_vav.p, M.a));
  UFE_TRY(d2h_field(c, st->N_a, V.N_a.p, M.a)); UFE_TRY(d2h_field(c, st->N_b, V.N_b.p, M.b));
  UFE_TRY(d2h_field(c, st->dN_dx_b, V.dN_dx_b.p, M.b)); UFE_TRY(d2h_field(c, st->dN_dy_b, V.dN_dy_b.p, M.b));
  UFE_TRY(d2h_field(c, st->F1_3D_a, V.F1_3D_a.p, M.a, nz)); UFE_TRY(d2h_field(c, st->F2_3D_a, V.F2_3D_a.p, M.a, nz));
  UFE_TRY(d2h_field(c, st->F1_3D_b, V.F1_3D_b.p, M.b, nz)); UFE_TRY(d2h_field(c, st->F2_3D_b, V.F2_3D_b.p, M.b, nz));
  UFE_TRY(d2h_field(c, st->basal_friction_coefficient_a, V.beta_a.p, M.a));
  UFE_TRY(d2h_field(c, st->basal_friction_coefficient_b, V.beta_b.p, M.b));
  UFE_TRY(d2h_field(c, st->beta_eff_a, V.beta_eff_a.p, M.a)); UFE_TRY(d2h_field(c, st->beta_eff_b, V.beta_eff_b.p, M.b));
  }
  if (sync) UFE_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

// solve_DIVA_infinite_slab (approximation = 1) / solve_SSA (approximation = 0) on device-resident fields
int velocity_solve(Ctx* c, const ufe_b200_config* cfg, int approximation, int* n_visc_its, int* n_Axb_its) {
  UFE_CHECK(c->vel && c->vel->uploaded, "velocity solve: call ufe_b200_velocity_upload first");
  Mesh& M = *c->mesh; Velocity& V = *c->vel;
  cudaStream_t st = c->stream;
  const bool DIVA = approximation == 1;
  const int nz = M.nz, na = M.a.n_loc(), nb = M.b.n_loc();
  *n_visc_its = 0; *n_Axb_its = 0;
  VelPar p = make_par(cfg);
  const int law = cfg->choice_sliding_law;
  const bool coulomb_family = law == UFE_SLID_COULOMB || law == UFE_SLID_BUDD || law == UFE_SLID_ZOET_IVERSON;
  if (coulomb_family) UFE_CHECK(V.have_tfa && V.have_neff, "this sliding law needs till_friction_angle and effective_pressure");
  if (law == UFE_SLID_WEERTMAN || law == UFE_SLID_TSAI2015 || law == UFE_SLID_SCHOOF2005) UFE_CHECK(V.have_beta_sq, "this sliding law needs beta_sq");
  if (law == UFE_SLID_TSAI2015 || law == UFE_SLID_SCHOOF2005) UFE_CHECK(V.have_alpha_sq && V.have_neff, "this sliding law needs alpha_sq and effective_pressure");

  // no grounded ice (or, for the SSA, no sliding): nothing to solve (DIVA_solver_infinite_slab.f90:87-98, SSA_main.f90:116-122)
  if (!V.grounded_ice_exists || (!DIVA && law == UFE_SLID_NO_SLIDING)) {
    UFE_CUDA(cudaMemsetAsync(V.u.p, 0, sizeof(double) * V.u.n, st)); UFE_CUDA(cudaMemsetAsync(V.v.p, 0, sizeof(double) * V.v.n, st));
    if (DIVA) {
      UFE_CUDA(cudaMemsetAsync(V.u_base.p, 0, sizeof(double) * V.u_base.n, st)); UFE_CUDA(cudaMemsetAsync(V.v_base.p, 0, sizeof(double) * V.v_base.n, st));
      UFE_CUDA(cudaMemsetAsync(V.u_3D.p, 0, sizeof(double) * V.u_3D.n, st)); UFE_CUDA(cudaMemsetAsync(V.v_3D.p, 0, sizeof(double) * V.v_3D.n, st));
    }
    // the solver object's strain rates stay at their initial zeros in the reference (allocate_SSA_solver / allocate_DIVA_solver)
    for (DBuf<double>* b : {&V.du_dx, &V.du_dy, &V.dv_dx, &V.dv_dy}) UFE_CUDA(cudaMemsetAsync(b->p, 0, sizeof(double) * b->n, st));
    if (DIVA) { UFE_CUDA(cudaMemsetAsync(V.du_dz_3D.p, 0, sizeof(double) * V.du_dz_3D.n, st)); UFE_CUDA(cudaMemsetAsync(V.dv_dz_3D.p, 0, sizeof(double) * V.dv_dz_3D.n, st)); V.have_dudz_3d = true; }
    V.last_approximation = DIVA ? UFE_APPROX_DIVA : UFE_APPROX_SSA; V.have_strain_2d = true;
    UFE_CUDA(cudaStreamSynchronize(st));
    return 0;
  }
  UFE_TRY(ensure_pattern(c, p, c->bc_version));
  const int g_a = grid_for(c, na, 128, 16), g_b = grid_for(c, nb, 128, 16);
  const Mesh::OpGroup& gab = M.grp[Mesh::G_AB];
  const Mesh::OpGroup& gba = M.grp[Mesh::G_BA];

  // --- once per solve: halos of the geometry, till yield stress, driving stress, vertically averaged flow factor
  UFE_TRY(halo_exchange(c, M.a, V.Hi.p)); UFE_TRY(halo_exchange(c, M.a, V.Hs.p));
  if (coulomb_family) {
    { UFE_LAUNCH(c, TC_COLUMN); k_tauc0<<<g_a, 128, 0, st>>>(na, M.a.off(), V.effective_pressure.p, V.till_friction_angle.p, V.tauc0.p); }
    UFE_TRY(halo_exchange(c, M.a, V.tauc0.p));
    UFE_TRY(halo_exchange_int(c, M.a, V.mask_grounded_ice.p));  // the neighbour test reads mask_grounded_ice of halo vertices
    UFE_LAUNCH(c, TC_COLUMN);
    k_tauc_extend<<<g_a, 128, 0, st>>>(na, M.a.off(), M.a.i1, M.a.lo, M.nV, M.nC.p, M.C.p, V.mask_grounded_ice.p, V.mask_icefree_land.p,
                                       V.tauc0.p, V.tauc.p, 1000.0 * cfg->ice_density * cfg->grav, cfg->Hi_min * cfg->ice_density * cfg->grav);
  }
  {
    UFE_LAUNCH(c, TC_TRI);
    k_driving_stress<<<g_b, 128, 0, st>>>(nb, M.b.off(), gab.ptr.p, gab.ind.p, gab.val[0].p, gab.val[1].p, gab.val[2].p, V.Hi.p, V.Hs.p,
                                          V.tau_dx.p, V.tau_dy.p, cfg->ice_density * cfg->grav);
  }
  if (!DIVA) {
    UFE_LAUNCH(c, TC_COLUMN);
    k_A_vav<<<g_a, 128, 0, st>>>(na, M.a.off(), M.a.n_ext(), nz, M.zeta.p, V.A_flow.p, V.A_flow_vav.p);
  }

  VertexArgs va;
  va.n = na; va.off = M.a.off(); va.n_ext = M.a.n_ext(); va.nb_ext = M.b.n_ext(); va.nz = nz; va.vi1 = M.a.i1; va.nV = M.nV;
  va.ptr = gba.ptr.p; va.ind = gba.ind.p; va.w_map = gba.val[0].p; va.w_ddx = gba.val[1].p; va.w_ddy = gba.val[2].p;
  va.zeta = M.zeta.p; va.Vy = M.V.p + M.nV;
  va.u = V.u.p; va.v = V.v.p; va.u_base = V.u_base.p; va.v_base = V.v_base.p; va.tau_bx = V.tau_bx.p; va.tau_by = V.tau_by.p;
  va.eta_3D_b = V.eta_3D_b.p; va.Hi = V.Hi.p; va.A_flow = V.A_flow.p; va.A_flow_vav = V.A_flow_vav.p;
  va.tauc = V.tauc.p; va.beta_sq = V.beta_sq.p; va.alpha_sq = V.alpha_sq.p; va.neff = V.effective_pressure.p;
  va.du_dx = V.du_dx.p; va.du_dy = V.du_dy.p; va.dv_dx = V.dv_dx.p; va.dv_dy = V.dv_dy.p; va.du_dz_3D = V.du_dz_3D.p; va.dv_dz_3D = V.dv_dz_3D.p;
  va.eta_3D_a = V.eta_3D_a.p; va.eta_vav = V.eta_vav.p; va.N_a = V.N_a.p; va.F1_3D_a = V.F1_3D_a.p; va.F2_3D_a = V.F2_3D_a.p;
  va.beta_a = V.beta_a.p; va.beta_eff_a = V.beta_eff_a.p;
  TriArgs ta;
  ta.na_ext = M.a.n_ext(); ta.nb_ext = M.b.n_ext(); ta.nz = nz; ta.final_pass = 0;
  ta.ptr = gab.ptr.p; ta.ind = gab.ind.p; ta.w_map = gab.val[0].p; ta.w_ddx = gab.val[1].p; ta.w_ddy = gab.val[2].p;
  ta.N_a = V.N_a.p; ta.beta_a = V.beta_a.p; ta.beta_eff_a = V.beta_eff_a.p; ta.eta_3D_a = V.eta_3D_a.p; ta.F1_3D_a = V.F1_3D_a.p; ta.F2_3D_a = V.F2_3D_a.p;
  ta.fraction_gr_b = V.fraction_gr_b.p; ta.tau_dx = V.tau_dx.p; ta.tau_dy = V.tau_dy.p;
  ta.N_b = V.N_b.p; ta.dN_dx_b = V.dN_dx_b.p; ta.dN_dy_b = V.dN_dy_b.p; ta.beta_b = V.beta_b.p; ta.beta_eff_b = V.beta_eff_b.p;
  ta.eta_3D_b = V.eta_3D_b.p; ta.F1_3D_b = V.F1_3D_b.p; ta.F2_3D_b = V.F2_3D_b.p;
  StiffArgs sa = stiff_args(c, csr_values_unneeded(c, cfg));
  // layer-interleaved work copies of the two gathered 3-D fields (UFE_B200_INTERLEAVE=0: gather the reference layout directly, as before)
  // Measured at 4 M vertices (profiles/r02_perf_probe_variants_4M.log): the b-grid copy takes the column class from 6.1 to 3.7 ms per viscosity
  // iteration; the a-grid copy does nothing for k_tri_assemble (6.96 against 6.88 ms: its time is the operator rows and the 40 values it writes per
  // triangle, not the 36 eta gathers), so it is off unless UFE_B200_INTERLEAVE_A=1.  Read per solve: scripts/perf_probe.py toggles them.
  const bool interleave = DIVA && !(getenv("UFE_B200_INTERLEAVE") && atoi(getenv("UFE_B200_INTERLEAVE")) == 0);
  const bool interleave_a = DIVA && getenv("UFE_B200_INTERLEAVE_A") && atoi(getenv("UFE_B200_INTERLEAVE_A")) == 1;
  const size_t tile_bytes = sizeof(double) * 128 * (size_t)(nz + 1);
  va.rzT = nullptr; ta.etaT = nullptr;
  if (interleave) { UFE_TRY(V.rzT_b.alloc((size_t)M.b.n_ext() * nz)); va.rzT = V.rzT_b.p; }
  if (interleave_a) { UFE_TRY(V.etaT_a.alloc((size_t)M.a.n_ext() * nz)); ta.etaT = V.etaT_a.p; }

  // adaptive relaxation (DIVA_solver_infinite_slab.f90:121-131)
  double L2_uv = 1e9, relax = cfg->visc_it_relax, eps0 = cfg->Glens_flow_law_epsilon_sq_0;
  int nit_diverg = 0, it = 0;
  const double n_glen = cfg->Glens_flow_law_exponent;
  bool converged = false;
  while (!converged) {
    ++it;
    p.relax = relax; p.eps0 = eps0;
    p.eta_max = 0.5 * pow(1e-18, -1.0 / n_glen) * pow(eps0, (1.0 - n_glen) / (2.0 * n_glen));
    // halos of everything the vertex kernel gathers from the b-grid
    UFE_TRY(halo_exchange(c, M.b, V.u.p)); UFE_TRY(halo_exchange(c, M.b, V.v.p));
    if (DIVA) {
      UFE_TRY(halo_exchange(c, M.b, V.u_base.p)); UFE_TRY(halo_exchange(c, M.b, V.v_base.p));
      UFE_TRY(halo_exchange(c, M.b, V.tau_bx.p)); UFE_TRY(halo_exchange(c, M.b, V.tau_by.p));
      UFE_TRY(halo_exchange(c, M.b, V.eta_3D_b.p, nz));
      if (interleave) {
        UFE_LAUNCH(c, TC_COLUMN);
        k_layers_interleave<true><<<grid_for(c, M.b.n_ext(), 128), 128, tile_bytes, st>>>(M.b.n_ext(), nz, V.eta_3D_b.p, M.zeta.p, p.visc_eff_min, V.rzT_b.p);
      }
    }
    {
      UFE_LAUNCH(c, TC_COLUMN);
      if (!DIVA) k_vertex_ssa<<<g_a, 128, 0, st>>>(va, p);
      else if (nz == 12 && getenv("UFE_B200_VERTEX_NZ12")) k_vertex_diva<12><<<g_a, 128, 0, st>>>(va, p);
      else {
        // registers capped for 6 resident blocks per SM (80 registers): the kernel is bound by gather latency, more warps beat fewer spills
        // (column class per step at 4 M vertices: 6.5 ms with 4 blocks / 128 registers, 5.25 ms with 6, 5.4 ms with 8)
        static const int mb = getenv("UFE_B200_VERTEX_MB") ? atoi(getenv("UFE_B200_VERTEX_MB")) : 6;
        const bool col_smem = !(getenv("UFE_B200_VERTEX_SMEM") && atoi(getenv("UFE_B200_VERTEX_SMEM")) == 0);
        const size_t col_bytes = sizeof(double) * 3 * 128 * (size_t)nz;    // 36 KB at nz = 12: 4 blocks per SM leave ~100 KB of L1 for the gathers
        if (col_smem) {
          static bool attr_set = false;
          if (!attr_set) {
            UFE_CUDA(cudaFuncSetAttribute(k_vertex_diva<0, 4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(sizeof(double) * 3 * 128 * kMaxNz)));
            attr_set = true;
          }
          k_vertex_diva<0, 4, true><<<g_a, 128, col_bytes, st>>>(va, p);
        }
        else if (mb >= 8) k_vertex_diva<0, 8><<<g_a, 128, 0, st>>>(va, p);
        else if (mb >= 6) k_vertex_diva<0, 6><<<g_a, 128, 0, st>>>(va, p);
        else k_vertex_diva<0, 4><<<g_a, 128, 0, st>>>(va, p);
      }
    }
    UFE_TRY(halo_exchange(c, M.a, V.N_a.p)); UFE_TRY(halo_exchange(c, M.a, V.beta_a.p));
    if (DIVA) {
      UFE_TRY(halo_exchange(c, M.a, V.beta_eff_a.p)); UFE_TRY(halo_exchange(c, M.a, V.eta_3D_a.p, nz));
      UFE_TRY(halo_exchange(c, M.a, V.F2_3D_a.p, nz)); UFE_TRY(halo_exchange(c, M.a, V.F1_3D_a.p, nz));
      if (interleave_a) {
        UFE_LAUNCH(c, TC_ASSEMBLY);
        k_layers_interleave<false><<<grid_for(c, M.a.n_ext(), 128), 128, tile_bytes, st>>>(M.a.n_ext(), nz, V.eta_3D_a.p, M.zeta.p, 0.0, V.etaT_a.p);
      }
    }
    {
      UFE_LAUNCH(c, TC_ASSEMBLY);
      if (DIVA) k_tri_assemble<true><<<g_b, 128, 0, st>>>(ta, sa, p); else k_tri_assemble<false><<<g_b, 128, 0, st>>>(ta, sa, p);
      V.blk_filled = sa.blk_map != nullptr; V.csr_stale = sa.A_val == nullptr;
    }
    UFE_TRY(copy_bc_rhs(c, p));   // u, v still hold the previous solution here (k_store_prev below copies the same values)
    { UFE_LAUNCH(c, TC_EPILOGUE); k_store_prev<<<g_b, 128, 0, st>>>(nb, M.b.off(), V.u.p, V.v.p, V.u_prev.p, V.v_prev.p); }
    int its = 0;
    UFE_TRY(solve_linear_system(c, cfg, cfg->stress_balance_PETSc_rtol, cfg->stress_balance_PETSc_abstol, &its));
    *n_Axb_its += its;
    {
      UFE_LAUNCH(c, TC_EPILOGUE);
      const double* uv_own = V.uv.p + M.buv.off();
      if (DIVA) k_epilogue<true><<<g_b, kBlock, 0, st>>>(nb, M.b.off(), M.b.n_ext(), uv_own, V.u.p, V.v.p, V.u_prev.p, V.v_prev.p, V.beta_b.p,
                                                         V.beta_eff_b.p, V.F2_3D_b.p, V.u_base.p, V.v_base.p, V.tau_bx.p, V.tau_by.p, p,
                                                         V.partials.p, V.counter.p, V.red.p);
      else k_epilogue<false><<<g_b, kBlock, 0, st>>>(nb, M.b.off(), M.b.n_ext(), uv_own, V.u.p, V.v.p, V.u_prev.p, V.v_prev.p, V.beta_b.p,
                                                     V.beta_eff_b.p, V.F2_3D_b.p, V.u_base.p, V.v_base.p, V.tau_bx.p, V.tau_by.p, p,
                                                     V.partials.p, V.counter.p, V.red.p);
    }
    UFE_TRY(allreduce_sum(c, V.red.p, 2));
    UFE_CUDA(cudaMemcpyAsync(c->h_pinned + 8, V.red.p, 2 * sizeof(double), cudaMemcpyDeviceToHost, st));
    UFE_CUDA(cudaStreamSynchronize(st));
    const double res1 = c->h_pinned[8], res2 = c->h_pinned[9];
    const double L2_prev = L2_uv;
    L2_uv = 2.0 * res1 / fmax(res2, 1e-8);
    // divergence control (DIVA_solver_infinite_slab.f90:180-196)
    if (L2_uv > L2_prev) ++nit_diverg; else nit_diverg = 0;
    if (nit_diverg > 2) { nit_diverg = 0; relax *= 0.9; eps0 *= 1.2; }
    UFE_CHECK(!(relax < 0.05), "viscosity iteration still diverges even with very low relaxation factor!");
    UFE_CHECK(!(eps0 > 1e-5), "viscosity iteration still diverges even with very high effective strain rate regularisation!");
    converged = L2_uv < cfg->visc_it_norm_dUV_tol;
    if (it > cfg->visc_it_nit) break;  // warning in the reference (:211-214); not an error
  }
  if (DIVA) {
    // F1_3D_b / F2_3D_b of all layers are only needed now (calc_3D_velocities) and by restart / checksum consumers
    if (!p.diagnostics) {
      ta.final_pass = 1;
      UFE_LAUNCH(c, TC_TRI);
      k_tri_assemble<true><<<g_b, 128, 0, st>>>(ta, sa, p);
    }
    UFE_LAUNCH(c, TC_TRI);
    k_3D_velocities<<<g_b, 128, 0, st>>>(nb, M.b.off(), M.b.n_ext(), nz, V.u_base.p, V.v_base.p, V.tau_bx.p, V.tau_by.p, V.beta_b.p, V.F1_3D_b.p,
                                         V.u_3D.p, V.v_3D.p, law == UFE_SLID_NO_SLIDING ? 1 : 0);
  }
  UFE_CUDA(cudaGetLastError());
  UFE_CUDA(cudaStreamSynchronize(st));
  *n_visc_its = it;
  V.last_approximation = DIVA ? UFE_APPROX_DIVA : UFE_APPROX_SSA;
  V.have_strain_2d = true;
  if (DIVA) { V.have_dudz_3d = true; V.have_sia = false; }   // DIVA%du_dz_3D_a is part of every viscosity iteration; u_3D_b is DIVA's now
  return 0;
}

}  // namespace ufe

using namespace ufe;

extern "C" {

int ufe_b200_velocity_upload(ufe_b200_handle h, const ufe_b200_ice_inputs* ice, const ufe_b200_velocity_state* st) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c, "NULL handle");
  UFE_CUDA(cudaSetDevice(c->device));
  UFE_TRY(velocity_upload(c, ice, st));
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

int ufe_b200_velocity_solve_resident(ufe_b200_handle h, const ufe_b200_config* cfg, int approximation, int* n_visc_its, int* n_Axb_its) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && cfg && n_visc_its && n_Axb_its, "NULL argument");
  UFE_CUDA(cudaSetDevice(c->device));
  return velocity_solve(c, cfg, approximation, n_visc_its, n_Axb_its);
}

int ufe_b200_velocity_download(ufe_b200_handle h, ufe_b200_velocity_state* st) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && st, "NULL argument");
  UFE_CUDA(cudaSetDevice(c->device));
  return velocity_download(c, st);
}

static int solve_velocity_host(ufe_b200_handle h, const ufe_b200_config* cfg, int approximation, const ufe_b200_ice_inputs* ice,
                               ufe_b200_velocity_state* st, int* n_visc_its, int* n_Axb_its, const int32_t* bm, const double* bu,
                               const double* bv) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && cfg && ice && st && n_visc_its && n_Axb_its, "NULL argument");
  UFE_CUDA(cudaSetDevice(c->device));
  UFE_TRY(velocity_upload(c, ice, st));
  UFE_TRY(upload_bc(c, bm, bu, bv));
  UFE_TRY(velocity_solve(c, cfg, approximation, n_visc_its, n_Axb_its));
  return velocity_download(c, st);
}

int ufe_b200_solve_DIVA(ufe_b200_handle h, const ufe_b200_config* cfg, const ufe_b200_ice_inputs* ice, ufe_b200_velocity_state* DIVA,
                        int* n_visc_its, int* n_Axb_its, const int32_t* bm, const double* bu, const double* bv) {
  return solve_velocity_host(h, cfg, 1, ice, DIVA, n_visc_its, n_Axb_its, bm, bu, bv);
}
int ufe_b200_solve_SSA(ufe_b200_handle h, const ufe_b200_config* cfg, const ufe_b200_ice_inputs* ice, ufe_b200_velocity_state* SSA,
                       int* n_visc_its, int* n_Axb_its, const int32_t* bm, const double* bu, const double* bv) {
  return solve_velocity_host(h, cfg, 0, ice, SSA, n_visc_its, n_Axb_its, bm, bu, bv);
}

int ufe_b200_solve_SSA_DIVA_linearised(ufe_b200_handle h, const ufe_b200_config* cfg, double* u_b, double* v_b, const double* N_b,
                                       const double* dN_dx_b, const double* dN_dy_b, const double* beta_b, const double* tau_dx_b,
                                       const double* tau_dy_b, double* u_b_prev, double* v_b_prev, double rtol, double abstol,
                                       int* n_Axb_its, const int32_t* bm, const double* bu, const double* bv) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && cfg && u_b && v_b && N_b && dN_dx_b && dN_dy_b && beta_b && tau_dx_b && tau_dy_b, "NULL argument");
  UFE_CUDA(cudaSetDevice(c->device));
  UFE_TRY(ensure_velocity(c));
  Mesh& M = *c->mesh; Velocity& V = *c->vel;
  cudaStream_t st = c->stream;
  const int nb = M.b.n_loc();
  UFE_TRY(h2d_field(c, V.u.p, M.b, (const double*)u_b)); UFE_TRY(h2d_field(c, V.v.p, M.b, (const double*)v_b));
  UFE_TRY(h2d_field(c, V.N_b.p, M.b, N_b)); UFE_TRY(h2d_field(c, V.dN_dx_b.p, M.b, dN_dx_b)); UFE_TRY(h2d_field(c, V.dN_dy_b.p, M.b, dN_dy_b));
  UFE_TRY(h2d_field(c, V.beta_eff_b.p, M.b, beta_b)); UFE_TRY(h2d_field(c, V.tau_dx.p, M.b, tau_dx_b)); UFE_TRY(h2d_field(c, V.tau_dy.p, M.b, tau_dy_b));
  UFE_TRY(upload_bc(c, bm, bu, bv));
  V.uploaded = true;
  VelPar p = make_par(cfg);
  UFE_TRY(ensure_pattern(c, p, c->bc_version));
  StiffArgs sa = stiff_args(c, csr_values_unneeded(c, cfg));
  const int g_b = grid_for(c, nb, 128, 16);
  { UFE_LAUNCH(c, TC_EPILOGUE); k_store_prev<<<g_b, 128, 0, st>>>(nb, M.b.off(), V.u.p, V.v.p, V.u_prev.p, V.v_prev.p); }
  { UFE_LAUNCH(c, TC_ASSEMBLY); k_assemble<<<g_b, 128, 0, st>>>(sa, p, V.N_b.p, V.dN_dx_b.p, V.dN_dy_b.p, V.beta_eff_b.p, V.tau_dx.p, V.tau_dy.p); }
  V.blk_filled = sa.blk_map != nullptr; V.csr_stale = sa.A_val == nullptr;
  UFE_TRY(copy_bc_rhs(c, p));
  int its = 0;
  UFE_TRY(solve_linear_system(c, cfg, rtol, abstol, &its));
  { UFE_LAUNCH(c, TC_EPILOGUE); k_deinterleave<<<g_b, 128, 0, st>>>(nb, M.b.off(), V.uv.p + M.buv.off(), V.u.p, V.v.p); }
  UFE_CUDA(cudaGetLastError());
  UFE_TRY(d2h_field(c, u_b, V.u.p, M.b)); UFE_TRY(d2h_field(c, v_b, V.v.p, M.b));
  UFE_TRY(d2h_field(c, u_b_prev, V.u_prev.p, M.b)); UFE_TRY(d2h_field(c, v_b_prev, V.v_prev.p, M.b));
  UFE_CUDA(cudaStreamSynchronize(st));
  if (n_Axb_its) *n_Axb_its = its;
  return 0;
}

int ufe_b200_get_stiffness_size(ufe_b200_handle h, int* m_loc, int* nnz_loc) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && c->vel && c->vel->pattern_valid, "no stiffness matrix has been assembled yet");
  *m_loc = c->vel->A.m_loc; *nnz_loc = (int)c->vel->A.nnz;
  return 0;
}

// reference layout: ptr 1-based offsets, ind 1-based global columns (n = tiuv2n(ti,uv) = 2(ti-1)+uv, mesh_translation_tables.f90:190-198)
int ufe_b200_get_stiffness(ufe_b200_handle h, int32_t* ptr, int32_t* ind, double* val, double* bb) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && c->vel && c->vel->pattern_valid, "no stiffness matrix has been assembled yet");
  UFE_CUDA(cudaSetDevice(c->device));
  UFE_TRY(ensure_csr_values(c));
  const DevCsr& A = c->vel->A;
  UFE_CUDA(cudaMemcpyAsync(ptr, A.ptr, sizeof(int) * (A.m_loc + 1), cudaMemcpyDeviceToHost, c->stream));
  UFE_CUDA(cudaMemcpyAsync(ind, A.ind, sizeof(int) * A.nnz, cudaMemcpyDeviceToHost, c->stream));
  UFE_CUDA(cudaMemcpyAsync(val, A.val, sizeof(double) * A.nnz, cudaMemcpyDeviceToHost, c->stream));
  UFE_CUDA(cudaMemcpyAsync(bb, c->vel->bb.p, sizeof(double) * A.m_loc, cudaMemcpyDeviceToHost, c->stream));
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  for (int r = 0; r <= A.m_loc; ++r) ptr[r] += 1;
  for (long long k = 0; k < A.nnz; ++k) ind[k] += A.cols.lo;
  return 0;
}

int ufe_b200_set_mg_params(ufe_b200_handle h, int nu_pre, int nu_post, double omega, double overcorr, int w_levels, int fp32, int dense_max) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c != nullptr, "ufe_b200_set_mg_params: null handle");
  MgParams& p = c->mg_params;
  const int old_fp32 = p.fp32, old_dense = p.dense_max;
  if (nu_pre >= 1) p.nu_pre = nu_pre;
  if (nu_post >= 1) p.nu_post = nu_post;
  if (omega > 0.0) p.omega = omega;
  if (overcorr > 0.0) p.overcorr = overcorr;
  if (w_levels >= 0) p.w_levels = w_levels;
  if (fp32 >= 0) p.fp32 = fp32 != 0;
  if (dense_max >= 1) p.dense_max = dense_max < 128 ? dense_max : 128;
  if (c->vel) {
    Mg& G = c->vel->mg;
    G.drop_graph();                                   // the captured launch sequence depends on the cycle shape
    if (p.fp32 != old_fp32 || p.dense_max != old_dense) G.valid = false;   // storage precision / depth are fixed at build time
    else if (G.valid) { const int ug = G.p.use_graph; G.p = p; G.p.use_graph = ug; }
  }
  return 0;
}

int ufe_b200_set_mg_option(ufe_b200_handle h, const char* name, double value) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c != nullptr && name != nullptr, "ufe_b200_set_mg_option: NULL argument");
  MgParams& p = c->mg_params;
  bool rebuild = false;
  if (!strcmp(name, "adaptive")) p.adaptive = (int)value;
  else if (!strcmp(name, "adaptive_max")) p.adaptive_max = value;
  else if (!strcmp(name, "tail_max")) { p.tail_max = (int)value; rebuild = true; }
  else if (!strcmp(name, "tail_cluster")) { p.tail_cluster = (int)value; rebuild = true; }
  else if (!strcmp(name, "coarse_sweeps")) p.coarse_sweeps = (int)value;
  else if (!strcmp(name, "max_levels")) { p.max_levels = (int)value; rebuild = true; }     // a shorter hierarchy ends in coarse_sweeps sweeps
  else if (!strcmp(name, "dense_max")) { p.dense_max = (int)value; rebuild = true; }
  else if (!strcmp(name, "stop_nodes")) { p.stop_nodes = (int)value; rebuild = true; }
  else if (!strcmp(name, "coarsest_cluster")) { p.coarsest_cluster = (int)value; rebuild = true; }
  else if (!strcmp(name, "coarsest_dsmem")) p.coarsest_dsmem = (int)value;
  else if (!strcmp(name, "coarsest_threads")) p.coarsest_threads = (int)value;
  else if (!strcmp(name, "coarsest_cheb")) p.coarsest_cheb = (int)value;
  else if (!strcmp(name, "cheb_ratio")) p.cheb_ratio = value;
  else if (!strcmp(name, "w_levels")) p.w_levels = (int)value;
  else if (!strcmp(name, "fuse")) p.fuse = (int)value;
  else if (!strcmp(name, "use_graph")) p.use_graph = (int)value;
  else if (!strcmp(name, "nu_pre")) p.nu_pre = (int)value;          // 0 allowed here: no pre-smoothing (the right-hand side itself is restricted)
  else if (!strcmp(name, "nu_post")) p.nu_post = (int)value;
  else UFE_CHECK(false, "ufe_b200_set_mg_option: unknown option '%s'", name);
  if (c->vel) {
    Mg& G = c->vel->mg;
    G.drop_graph();
    if (rebuild) G.valid = false;
    else if (G.valid) G.p = p;
  }
  return 0;
}

int ufe_b200_get_mg_info(ufe_b200_handle h, int* n_levels, int* nodes_per_level, long long* fine_bytes, long long* coarse_bytes, double* setup_s) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && c->vel && c->vel->mg.valid, "ufe_b200_get_mg_info: no multigrid hierarchy has been built");
  const Mg& G = c->vel->mg;
  const long long vb = G.p.fp32 ? 4 : 8;
  if (n_levels) *n_levels = (int)G.lv.size();
  long long cb = 0;
  for (size_t l = 0; l < G.lv.size(); ++l) {
    if (nodes_per_level && l < 16) nodes_per_level[l] = G.lv[l].n;
    const long long b = G.lv[l].n_slots * (4 + 4 * vb) + 4LL * (G.lv[l].n_slices + 1);
    if (l == 0) { if (fine_bytes) *fine_bytes = b; } else cb += b;
  }
  if (coarse_bytes) *coarse_bytes = cb;
  if (setup_s) *setup_s = G.setup_host_s;
  return 0;
}

int ufe_b200_get_stiffness_storage_bytes(ufe_b200_handle h, long long* csr_bytes, long long* blocked_bytes) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && c->vel && c->vel->pattern_valid, "no stiffness matrix has been assembled yet");
  const DevCsr& A = c->vel->A;
  *csr_bytes = 12LL * A.nnz + 4LL * (A.m_loc + 1);
  *blocked_bytes = 36LL * c->vel->blk.n_slots + 4LL * (c->vel->blk.n_slices + 1);
  return 0;
}

}  // extern "C"
