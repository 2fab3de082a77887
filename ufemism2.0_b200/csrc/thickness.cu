// Ice-thickness (mass conservation) step on the device (north_star (c); K11/K12 of SURVEY 2e).
// Reference (paths relative to src/UFEMISM/ice_dynamics):
//   calc_dHi_dt                              conservation_of_mass/conservation_of_mass_main.f90:27-119
//   calc_dHi_dt_explicit + BC                conservation_of_mass/conservation_of_mass_explicit.f90:24-297
//   calc_dHi_dt_semiimplicit + BC            conservation_of_mass/conservation_of_mass_semiimplicit.f90:25-322
//   calc_ice_flux_divergence_matrix_upwind, calc_flux_limited_timestep, calc_n_interior_neighbours
//                                            conservation_of_mass/conservation_of_mass_utilities.f90:23-239
//   map_velocities_from_b_to_c_2D            utilities/map_velocities_to_c_grid.f90:20-77
//   ice_surface_elevation, Hi_from_Hb_Hs_and_SL   src/UPSY/basic/math_utilities/ice_geometry_basics.f90:33-94
//
// The reference allgathers u_b, v_b, u_c, v_c and fraction_margin and rebuilds M_divQ with add_entry twice per call.
// Here the b->c velocity map is fused into the flux stencil (no c-grid vector exists), M_divQ lives on a pattern built
// once per mesh ([vi, C(vi,1..nC)] per row, the reference's construction order) and only halos are exchanged.
#include "thickness.cuh"

#include <cub/device/device_scan.cuh>

#include "csr.cuh"
#include "mesh.cuh"
#include "reduce.cuh"
#include "velocity.cuh"

namespace ufe {

void thickness_free(Thickness* t) { delete t; }

namespace {

struct ThPar {
  double rho_i, rho_w, Hi_min, dt_ice_max;
  int BC_H[4];
};

__device__ __forceinline__ double surf_elev(const ThPar& p, double Hi, double Hb, double SL) {
  return Hi + fmax(SL - p.rho_i / p.rho_w * Hi, Hb);
}
__device__ __forceinline__ double Hi_from_Hb_Hs_SL(const ThPar& p, double Hb, double Hs, double SL) {
  const double Hi_float = fmax(0.0, (SL - Hb) * (p.rho_w / p.rho_i));
  const double Hs_float = Hb + Hi_float;
  if (Hs > Hs_float) return Hs - Hb;
  return fmin(Hi_float, (Hs - SL) / (1.0 - (p.rho_i / p.rho_w)));
}
__device__ __forceinline__ int side_of(int bi) { return (bi <= 2) ? 0 : (bi <= 4) ? 1 : (bi <= 6) ? 2 : 3; }

struct MeshV {
  int nV, nE, a_lo, b_lo, vi1, n, off;
  const int *nC, *C, *VE, *ETri, *VBI;
  const double *A, *Cw, *D_x, *D_y, *D;
};

__global__ void k_th_count(MeshV m, int* __restrict__ ptr) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < m.n; r += (long long)gridDim.x * blockDim.x)
    ptr[r] = m.nC[m.vi1 - 1 + r] + 1;
}
__global__ void k_th_fill_ind(MeshV m, const int* __restrict__ ptr, int* __restrict__ ind) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < m.n; r += (long long)gridDim.x * blockDim.x) {
    const int vi = m.vi1 + (int)r;
    int k = ptr[r];
    ind[k++] = m.off + (int)r;
    for (int ci = 0; ci < m.nC[vi - 1]; ++ci) ind[k++] = m.C[(size_t)ci * m.nV + vi - 1] - m.a_lo;
  }
}

// M_divQ values (conservation_of_mass_utilities.f90:78-132) with the b->c map fused in (map_velocities_to_c_grid.f90:51-70),
// divQ = M_divQ Hi and the explicit rate dHi_dt = -divQ + f_margin (SMB + BMB - target) + LMB (explicit.f90:102-113)
__global__ void __launch_bounds__(128) k_th_flux(MeshV m, ThPar p, const int* __restrict__ ptr, double* __restrict__ val,
                                                 const double* __restrict__ u_b, const double* __restrict__ v_b,
                                                 const double* __restrict__ fm, const double* __restrict__ Hi,
                                                 const double* __restrict__ SMB, const double* __restrict__ BMB,
                                                 const double* __restrict__ LMB, const double* __restrict__ target,
                                                 double* __restrict__ divQ, double* __restrict__ dHi_dt, double* __restrict__ AMB,
                                                 unsigned long long* dtbits) {
  double dt_lim = p.dt_ice_max;
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < m.n; r += (long long)gridDim.x * blockDim.x) {
    const int vi = m.vi1 + (int)r;
    const size_t i = m.off + r;
    const double A_i = m.A[vi - 1];
    const bool full_i = fm[i] >= 1.0;
    const int k0 = ptr[r];
    double c0 = 0.0, dq_nb = 0.0;
    for (int ci = 0; ci < m.nC[vi - 1]; ++ci) {
      const size_t k = (size_t)ci * m.nV + vi - 1;
      const int ei = m.VE[k], vj = m.C[k] - m.a_lo;
      const int til = m.ETri[ei - 1], tir = m.ETri[(size_t)m.nE + ei - 1];
      double uc, vc;
      if (til == 0) { uc = u_b[tir - m.b_lo]; vc = v_b[tir - m.b_lo]; }
      else if (tir == 0) { uc = u_b[til - m.b_lo]; vc = v_b[til - m.b_lo]; }
      else { uc = (u_b[til - m.b_lo] + u_b[tir - m.b_lo]) / 2.0; vc = (v_b[til - m.b_lo] + v_b[tir - m.b_lo]) / 2.0; }
      const double L_c = m.Cw[k];
      const double u_perp = uc * m.D_x[k] / m.D[k] + vc * m.D_y[k] / m.D[k];
      if (full_i) c0 = c0 + L_c * fmax(0.0, u_perp) / A_i;
      double cc = 0.0;
      if (fm[vj] >= 1.0) cc = L_c * fmin(0.0, u_perp) / A_i;
      val[k0 + 1 + ci] = cc;
    }
    val[k0] = c0;
    // divQ in the row order of the matrix: diagonal first, then the neighbours
    double dq = c0 * Hi[i];
    for (int ci = 0; ci < m.nC[vi - 1]; ++ci) dq += val[k0 + 1 + ci] * Hi[m.C[(size_t)ci * m.nV + vi - 1] - m.a_lo];
    (void)dq_nb;
    divQ[i] = dq;
    const double dh = -dq + fm[i] * (SMB[i] + BMB[i] - target[i]) + LMB[i];
    dHi_dt[i] = dh;
    AMB[i] = dh;
    // calc_flux_limited_timestep (utilities.f90:160-205), verbatim including max( dHi_dt, 1e-9)
    if (Hi[i] > p.Hi_min && dh < 0.0) dt_lim = fmin(dt_lim, Hi[i] / fmax(dh, 1e-9));
  }
  for (int o = 16; o > 0; o >>= 1) dt_lim = fmin(dt_lim, __shfl_xor_sync(0xffffffffu, dt_lim, o));
  if ((threadIdx.x & 31) == 0) atomicMin(dtbits, (unsigned long long)__double_as_longlong(dt_lim));  // dt_lim > 0: bit order = value order
}

// Hi( t+dt) = max( 0, Hi + dHi_dt dt) and its surface elevation (explicit.f90:125; BC :169-171)
__global__ void k_th_explicit(int n, int off, ThPar p, double dt, const double* __restrict__ Hi, const double* __restrict__ dHi_dt,
                              const double* __restrict__ Hb, const double* __restrict__ SL, double* __restrict__ Hi_new,
                              double* __restrict__ Hs, double* __restrict__ Hs_tot) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const size_t i = off + r;
    const double h = fmax(0.0, Hi[i] + dHi_dt[i] * dt);
    Hi_new[i] = h;
    const double s = surf_elev(p, h, Hb[i], SL[i]);
    Hs[i] = s; Hs_tot[i] = s;
  }
}
__global__ void k_th_hs(int n, int off, ThPar p, const double* __restrict__ Hi_new, const double* __restrict__ Hb,
                        const double* __restrict__ SL, double* __restrict__ Hs, double* __restrict__ Hs_tot) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const size_t i = off + r;
    const double s = surf_elev(p, Hi_new[i], Hb[i], SL[i]);
    Hs[i] = s; Hs_tot[i] = s;
  }
}

// calc_n_interior_neighbours (utilities.f90:207-237)
__global__ void k_th_nint(MeshV m, const int* __restrict__ mask_noice, int* __restrict__ n_int) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < m.n; r += (long long)gridDim.x * blockDim.x) {
    const int vi = m.vi1 + (int)r;
    int cnt = 0;
    if (m.VBI[vi - 1] > 0)
      for (int ci = 0; ci < m.nC[vi - 1]; ++ci) {
        const int vj = m.C[(size_t)ci * m.nV + vi - 1];
        if (m.VBI[vj - 1] == 0 && !mask_noice[vj - m.a_lo]) ++cnt;
      }
    n_int[r] = cnt;
  }
}

// apply_ice_thickness_BC_explicit, one pass (explicit.f90:181-296): pass 1 = border vertices with interior neighbours,
// pass 2 = border vertices without.  Reads the gathered snapshot Hs_tot, writes Hs and Hi_new of border vertices only.
__global__ void k_th_bc_pass(MeshV m, ThPar p, int pass, const int* __restrict__ mask_noice, const int* __restrict__ n_int,
                             const double* __restrict__ Hs_tot, const double* __restrict__ Hb, const double* __restrict__ SL,
                             double* __restrict__ Hs, double* __restrict__ Hi_new) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < m.n; r += (long long)gridDim.x * blockDim.x) {
    const int vi = m.vi1 + (int)r;
    const int bi = m.VBI[vi - 1];
    if (bi == 0) continue;
    const size_t i = m.off + r;
    if (p.BC_H[side_of(bi)] == UFE_BC_ZERO) { Hi_new[i] = 0.0; continue; }
    const bool has_int = n_int[r] > 0;
    if ((pass == 1) != has_int) continue;
    double Hs_sum = 0.0;
    for (int ci = 0; ci < m.nC[vi - 1]; ++ci) {
      const int vj = m.C[(size_t)ci * m.nV + vi - 1];
      if (pass == 2 || (m.VBI[vj - 1] == 0 && !mask_noice[vj - m.a_lo])) Hs_sum += Hs_tot[vj - m.a_lo];
    }
    const double Hs_av = Hs_sum / (double)(pass == 1 ? n_int[r] : m.nC[vi - 1]);
    const double s = fmax(Hb[i], Hs_av);
    Hs[i] = s;
    Hi_new[i] = Hi_from_Hb_Hs_SL(p, Hb[i], s, SL[i]);
  }
}

// prescribed thickness, no-ice mask, dHi_dt and AMB of the explicit scheme (explicit.f90:130-155)
__global__ void k_th_explicit_finish(int n, int off, double dt, const int* __restrict__ bc_mask, const double* __restrict__ bc_Hi,
                                     int have_bc, const int* __restrict__ mask_noice, const double* __restrict__ Hi,
                                     double* __restrict__ Hi_new, double* __restrict__ dHi_dt, double* __restrict__ AMB) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const size_t i = off + r;
    double h = Hi_new[i];
    if (have_bc && bc_mask[i] == 1) h = fmax(0.0, bc_Hi[i]);
    if (mask_noice[i]) h = 0.0;
    Hi_new[i] = h;
    const double d = (h - Hi[i]) / dt;
    dHi_dt[i] = d;
    AMB[i] = d - AMB[i];
  }
}

// AA = 1 + dt f_s M_divQ, bb, and the identity rows of the three BC routines (semiimplicit.f90:125-155, 184-322)
__global__ void k_th_semiimplicit(MeshV m, double dt, double fs, const int* __restrict__ ptr, const double* __restrict__ M0,
                                  double* __restrict__ AA, double* __restrict__ bb, const double* __restrict__ Hi,
                                  const double* __restrict__ divQ, const double* __restrict__ fm, const double* __restrict__ SMB,
                                  const double* __restrict__ BMB, const double* __restrict__ LMB, const double* __restrict__ target,
                                  const double* __restrict__ Hi_ex, const int* __restrict__ bc_mask, const double* __restrict__ bc_Hi,
                                  int have_bc, const int* __restrict__ mask_noice) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < m.n; r += (long long)gridDim.x * blockDim.x) {
    const int vi = m.vi1 + (int)r;
    const size_t i = m.off + r;
    const int k0 = ptr[r], k1 = ptr[r + 1];
    double b = Hi[i] - (dt * (1.0 - fs) * divQ[i]) + fmax(-1.0 * Hi[i], dt * (fm[i] * (SMB[i] + BMB[i] - target[i]) + LMB[i]));
    bool identity = false;
    if (m.VBI[vi - 1] > 0) { identity = true; b = Hi_ex[i]; }
    if (have_bc && bc_mask[i] == 1) { identity = true; b = bc_Hi[i]; }
    if (mask_noice[i]) { identity = true; b = 0.0; }
    if (identity) {
      AA[k0] = 1.0;
      for (int k = k0 + 1; k < k1; ++k) AA[k] = 0.0;
    } else {
      AA[k0] = M0[k0] * dt * fs + 1.0;
      for (int k = k0 + 1; k < k1; ++k) AA[k] = M0[k] * dt * fs;
    }
    bb[r] = b;
  }
}

// dHi_dt and AMB of the semi-implicit scheme (semiimplicit.f90:161-174) followed, when asked, by the tail of calc_dHi_dt
// (main.f90:83-109): clamp to zero, AMB += correction, dHi_dt recomputed
__global__ void k_th_rates(int n, int off, double dt_scheme, double dt_final, int scheme_rates, int finish, const double* __restrict__ Hi,
                           double* __restrict__ Hi_new, double* __restrict__ dHi_dt, double* __restrict__ AMB) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const size_t i = off + r;
    double h = Hi_new[i], d = dHi_dt[i], a = AMB[i];
    if (scheme_rates) {
      const double d0 = (h - Hi[i]) / dt_scheme;
      a = d0 - d0;   // AMB = dHi_dt - AMB with AMB = (Hi_tplusdt - Hi) / dt just before (semiimplicit.f90:162-174)
      d = d0;
    }
    if (finish) {
      if (h < 0.0) h = 0.0;
      a = a + (h - Hi[i]) / dt_final - d;
      d = (h - Hi[i]) / dt_final;
      Hi_new[i] = h;
    }
    dHi_dt[i] = d; AMB[i] = a;
  }
}

}  // namespace

int ensure_thickness(Ctx* c) {
  UFE_CHECK(c->mesh && c->mesh->have_ops, "no mesh / operators: call ufe_b200_mesh_create and ufe_b200_calc_matrix_operators first");
  if (c->thick) return 0;
  Mesh& M = *c->mesh;
  Thickness* T = new Thickness();
  c->thick = T;
  const size_t na = M.a.n_ext();
  for (auto* f : {&T->Hi, &T->Hb, &T->SL, &T->SMB, &T->BMB, &T->LMB, &T->fraction_margin, &T->target, &T->bc_Hi, &T->divQ, &T->dHi_dt, &T->AMB,
                  &T->Hi_new, &T->Hi_ex, &T->Hs, &T->Hs_tot}) {
    UFE_TRY(f->alloc(na));
    UFE_CUDA(cudaMemsetAsync(f->p, 0, sizeof(double) * na, c->stream));
  }
  UFE_TRY(T->bb.alloc(M.a.n_loc())); UFE_TRY(T->mask_noice.alloc(na)); UFE_TRY(T->bc_mask.alloc(na)); UFE_TRY(T->n_int.alloc(M.a.n_loc()));
  UFE_CUDA(cudaMemsetAsync(T->mask_noice.p, 0, sizeof(int) * na, c->stream));
  UFE_CUDA(cudaMemsetAsync(T->bc_mask.p, 0, sizeof(int) * na, c->stream));
  UFE_TRY(T->u_b.alloc(M.b.n_ext())); UFE_TRY(T->v_b.alloc(M.b.n_ext())); UFE_TRY(T->dtbits.alloc(1));
  UFE_CUDA(cudaMemsetAsync(T->u_b.p, 0, sizeof(double) * M.b.n_ext(), c->stream));
  UFE_CUDA(cudaMemsetAsync(T->v_b.p, 0, sizeof(double) * M.b.n_ext(), c->stream));
  return 0;
}

static MeshV mesh_view(const Mesh& M) {
  MeshV m;
  m.nV = M.nV; m.nE = M.nE; m.a_lo = M.a.lo; m.b_lo = M.b.lo; m.vi1 = M.a.i1; m.n = M.a.n_loc(); m.off = M.a.off();
  m.nC = M.nC.p; m.C = M.C.p; m.VE = M.VE.p; m.ETri = M.ETri.p; m.VBI = M.VBI.p;
  m.A = M.A.p; m.Cw = M.Cw.p; m.D_x = M.D_x.p; m.D_y = M.D_y.p; m.D = M.D.p;
  return m;
}

static int ensure_th_pattern(Ctx* c) {
  Mesh& M = *c->mesh; Thickness& T = *c->thick;
  if (T.pattern) return 0;
  const int n = M.a.n_loc();
  MeshV mv = mesh_view(M);
  UFE_TRY(T.M.ptr_buf.alloc((size_t)n + 1));
  T.M.ptr = T.M.ptr_buf.p;
  UFE_CUDA(cudaMemsetAsync(T.M.ptr, 0, sizeof(int) * ((size_t)n + 1), c->stream));
  const int grid = grid_for(c, n, 128, 16);
  { UFE_LAUNCH(c, TC_THICKNESS); k_th_count<<<grid, 128, 0, c->stream>>>(mv, T.M.ptr); }
  size_t tmp_bytes = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, T.M.ptr, T.M.ptr, n + 1, c->stream);
  DBuf<char> tmp;
  UFE_TRY(tmp.alloc(tmp_bytes));
  c->launches++;
  UFE_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, tmp_bytes, T.M.ptr, T.M.ptr, n + 1, c->stream));
  int nnz = 0;
  UFE_CUDA(cudaMemcpyAsync(&nnz, T.M.ptr + n, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  UFE_TRY(T.M.ind_buf.alloc(nnz)); UFE_TRY(T.M.val_buf.alloc(nnz)); UFE_TRY(T.Mval0.alloc(nnz));
  T.M.ind = T.M.ind_buf.p; T.M.val = T.M.val_buf.p;
  T.M.m_loc = n; T.M.nnz = nnz; T.M.rows = M.a; T.M.cols = M.a; T.M.vec_width = 8; T.M.valid = true;
  T.M.sell_src_ptr = nullptr;   // a new pattern: the sliced-ELL copy of the previous one is void
  { UFE_LAUNCH(c, TC_THICKNESS); k_th_fill_ind<<<grid, 128, 0, c->stream>>>(mv, T.M.ptr, T.M.ind); }
  UFE_CUDA(cudaGetLastError());
  T.pattern = true;
  return 0;
}

template <typename T>
static int th_h2d(Ctx* c, T* d_ext, const Range& r, const T* h) {
  if (!h) return 0;
  UFE_CUDA(cudaMemcpyAsync(d_ext + r.off(), h, sizeof(T) * r.n_loc(), cudaMemcpyHostToDevice, c->stream));
  return 0;
}
static int th_d2h(Ctx* c, double* h, const double* d_ext, const Range& r) {
  if (!h) return 0;
  UFE_CUDA(cudaMemcpyAsync(h, d_ext + r.off(), sizeof(double) * r.n_loc(), cudaMemcpyDeviceToHost, c->stream));
  return 0;
}

// apply_ice_thickness_BC_explicit on Hi_new (device)
static int th_apply_bc(Ctx* c, const MeshV& mv, const ThPar& p, double* Hi_new) {
  Mesh& M = *c->mesh; Thickness& T = *c->thick;
  const int grid = grid_for(c, mv.n, 128, 16);
  cudaStream_t st = c->stream;
  { UFE_LAUNCH(c, TC_THICKNESS); k_th_hs<<<grid, 128, 0, st>>>(mv.n, mv.off, p, Hi_new, T.Hb.p, T.SL.p, T.Hs.p, T.Hs_tot.p); }
  UFE_TRY(halo_exchange(c, M.a, T.Hs_tot.p));
  { UFE_LAUNCH(c, TC_THICKNESS); k_th_bc_pass<<<grid, 128, 0, st>>>(mv, p, 1, T.mask_noice.p, T.n_int.p, T.Hs_tot.p, T.Hb.p, T.SL.p, T.Hs.p, Hi_new); }
  UFE_CUDA(cudaMemcpyAsync(T.Hs_tot.p + mv.off, T.Hs.p + mv.off, sizeof(double) * mv.n, cudaMemcpyDeviceToDevice, st));  // second gather_to_all
  UFE_TRY(halo_exchange(c, M.a, T.Hs_tot.p));
  { UFE_LAUNCH(c, TC_THICKNESS); k_th_bc_pass<<<grid, 128, 0, st>>>(mv, p, 2, T.mask_noice.p, T.n_int.p, T.Hs_tot.p, T.Hb.p, T.SL.p, T.Hs.p, Hi_new); }
  return 0;
}

// calc_dHi_dt on device-resident inputs; u_b/v_b: extended b-grid vectors with valid halos
int thickness_step(Ctx* c, const ufe_b200_config* cfg, const double* u_b, const double* v_b, double* dt, int finish, int* n_Axb_its) {
  Mesh& M = *c->mesh; Thickness& T = *c->thick;
  cudaStream_t st = c->stream;
  UFE_TRY(ensure_th_pattern(c));
  MeshV mv = mesh_view(M);
  ThPar p;
  p.rho_i = cfg->ice_density; p.rho_w = cfg->seawater_density; p.Hi_min = cfg->Hi_min; p.dt_ice_max = cfg->dt_ice_max;
  for (int k = 0; k < 4; ++k) p.BC_H[k] = cfg->BC_H[k];
  const int n = mv.n, grid = grid_for(c, n, 128, 16);
  *n_Axb_its = 0;
  const double dt_in = *dt;
  // halos of what the stencil gathers
  UFE_TRY(halo_exchange(c, M.a, T.Hi.p)); UFE_TRY(halo_exchange(c, M.a, T.fraction_margin.p));
  UFE_TRY(halo_exchange_int(c, M.a, T.mask_noice.p));
  const unsigned long long big = (unsigned long long)__builtin_bit_cast(long long, cfg->dt_ice_max);
  UFE_CUDA(cudaMemcpyAsync(T.dtbits.p, &big, sizeof(big), cudaMemcpyHostToDevice, st));
  {
    UFE_LAUNCH(c, TC_THICKNESS);
    k_th_flux<<<grid, 128, 0, st>>>(mv, p, T.M.ptr, T.Mval0.p, u_b, v_b, T.fraction_margin.p, T.Hi.p, T.SMB.p, T.BMB.p, T.LMB.p, T.target.p,
                                    T.divQ.p, T.dHi_dt.p, T.AMB.p, T.dtbits.p);
  }
  { UFE_LAUNCH(c, TC_THICKNESS); k_th_nint<<<grid, 128, 0, st>>>(mv, T.mask_noice.p, T.n_int.p); }
  unsigned long long bits = 0;
  UFE_CUDA(cudaMemcpyAsync(&bits, T.dtbits.p, sizeof(bits), cudaMemcpyDeviceToHost, st));
  UFE_CUDA(cudaStreamSynchronize(st));
  double dt_max = __builtin_bit_cast(double, bits);
  if (c->nranks > 1) {  // MPI_ALLREDUCE( ..., MPI_MIN, ...) (utilities.f90:198)
    UFE_CUDA(cudaMemcpyAsync(T.bb.p, &dt_max, sizeof(double), cudaMemcpyHostToDevice, st));
    UFE_TRY(allreduce_min(c, T.bb.p, 1));
    UFE_CUDA(cudaMemcpyAsync(&dt_max, T.bb.p, sizeof(double), cudaMemcpyDeviceToHost, st));
    UFE_CUDA(cudaStreamSynchronize(st));
  }
  dt_max = dt_max > cfg->dt_ice_min ? dt_max : cfg->dt_ice_min;
  const double dt_ex = dt_in < dt_max ? dt_in : dt_max;
  const bool semi = cfg->choice_ice_integration_method == UFE_INTEGRATION_SEMI_IMPLICIT;
  double* Hi_ex = semi ? T.Hi_ex.p : T.Hi_new.p;
  { UFE_LAUNCH(c, TC_THICKNESS); k_th_explicit<<<grid, 128, 0, st>>>(n, mv.off, p, dt_ex, T.Hi.p, T.dHi_dt.p, T.Hb.p, T.SL.p, Hi_ex, T.Hs.p, T.Hs_tot.p); }
  UFE_TRY(th_apply_bc(c, mv, p, Hi_ex));
  {
    UFE_LAUNCH(c, TC_THICKNESS);
    k_th_explicit_finish<<<grid, 128, 0, st>>>(n, mv.off, dt_ex, T.bc_mask.p, T.bc_Hi.p, T.have_bc ? 1 : 0, T.mask_noice.p, T.Hi.p, Hi_ex,
                                               T.dHi_dt.p, T.AMB.p);
  }
  double dt_final;
  if (!semi) {
    *dt = dt_ex; dt_final = dt_ex;
    if (finish) { UFE_LAUNCH(c, TC_THICKNESS); k_th_rates<<<grid, 128, 0, st>>>(n, mv.off, dt_ex, dt_final, 0, 1, T.Hi.p, T.Hi_new.p, T.dHi_dt.p, T.AMB.p); }
  } else {
    // the semi-implicit scheme keeps the caller's dt (semiimplicit.f90:104-113: only dt_ex is limited)
    dt_final = dt_in;
    UFE_TRY(th_apply_bc(c, mv, p, T.Hi_ex.p));   // apply_ice_thickness_BC_matrix_domain_border re-applies the explicit BC (:236)
    {
      UFE_LAUNCH(c, TC_THICKNESS);
      k_th_semiimplicit<<<grid, 128, 0, st>>>(mv, dt_in, cfg->dHi_semiimplicit_fs, T.M.ptr, T.Mval0.p, T.M.val, T.bb.p, T.Hi.p, T.divQ.p,
                                              T.fraction_margin.p, T.SMB.p, T.BMB.p, T.LMB.p, T.target.p, T.Hi_ex.p, T.bc_mask.p, T.bc_Hi.p,
                                              T.have_bc ? 1 : 0, T.mask_noice.p);
    }
    KrylovParams kp{cfg->dHi_PETSc_rtol, cfg->dHi_PETSc_abstol, cfg->ksp_maxit > 0 ? cfg->ksp_maxit : 2000, UFE_KSP_BICGSTAB,
                    cfg->pc_type == UFE_PC_NONE ? UFE_PC_NONE : UFE_PC_JACOBI};
    if (cfg->ksp_initial_guess_nonzero) {  // start from the explicit solution (the reference passes Hi + dt dHi_dt, semiimplicit.f90:152)
      kp.use_guess = 1;
      UFE_CUDA(cudaMemcpyAsync(T.Hi_new.p + mv.off, T.Hi_ex.p + mv.off, sizeof(double) * n, cudaMemcpyDeviceToDevice, st));
    }
    UFE_TRY(csr_ksp_solve(c, T.M, T.bb.p, T.Hi_new.p, kp, TC_THICKNESS, n_Axb_its));
    { UFE_LAUNCH(c, TC_THICKNESS); k_th_rates<<<grid, 128, 0, st>>>(n, mv.off, dt_in, dt_final, 1, finish, T.Hi.p, T.Hi_new.p, T.dHi_dt.p, T.AMB.p); }
  }
  UFE_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace ufe

using namespace ufe;

extern "C" {

int ufe_b200_calc_dHi_dt(ufe_b200_handle h, const ufe_b200_config* cfg, ufe_b200_thickness_io* io, double* dt, int finish, int* n_Axb_its) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && cfg && io && dt, "NULL argument");
  UFE_CUDA(cudaSetDevice(c->device));
  UFE_TRY(ensure_thickness(c));
  Mesh& M = *c->mesh; Thickness& T = *c->thick;
  UFE_CHECK(io->Hi && io->Hb && io->SL && io->SMB && io->BMB && io->LMB && io->fraction_margin && io->mask_noice && io->dHi_dt_target,
            "ufe_b200_calc_dHi_dt: Hi, Hb, SL, SMB, BMB, LMB, fraction_margin, mask_noice, dHi_dt_target are required");
  const double *u_b, *v_b;
  {
  UFE_REGION(c, TC_H2D);
  UFE_TRY(th_h2d(c, T.Hi.p, M.a, io->Hi)); UFE_TRY(th_h2d(c, T.Hb.p, M.a, io->Hb)); UFE_TRY(th_h2d(c, T.SL.p, M.a, io->SL));
  UFE_TRY(th_h2d(c, T.SMB.p, M.a, io->SMB)); UFE_TRY(th_h2d(c, T.BMB.p, M.a, io->BMB)); UFE_TRY(th_h2d(c, T.LMB.p, M.a, io->LMB));
  UFE_TRY(th_h2d(c, T.fraction_margin.p, M.a, io->fraction_margin)); UFE_TRY(th_h2d(c, T.target.p, M.a, io->dHi_dt_target));
  UFE_TRY(th_h2d(c, T.mask_noice.p, M.a, io->mask_noice));
  if (io->BC_prescr_mask || io->BC_prescr_Hi) {
    UFE_CHECK(io->BC_prescr_mask && io->BC_prescr_Hi, "need to provide prescribed both Hi and mask!");  // explicit.f90:134-136
    UFE_TRY(th_h2d(c, T.bc_mask.p, M.a, io->BC_prescr_mask)); UFE_TRY(th_h2d(c, T.bc_Hi.p, M.a, io->BC_prescr_Hi));
    T.have_bc = true;
  } else T.have_bc = false;
  if (io->u_vav_b && io->v_vav_b) { UFE_TRY(th_h2d(c, T.u_b.p, M.b, io->u_vav_b)); UFE_TRY(th_h2d(c, T.v_b.p, M.b, io->v_vav_b)); }
  }
  if (io->u_vav_b && io->v_vav_b) {
    UFE_TRY(halo_exchange(c, M.b, T.u_b.p)); UFE_TRY(halo_exchange(c, M.b, T.v_b.p));
    u_b = T.u_b.p; v_b = T.v_b.p;
  } else {
    UFE_CHECK(c->vel && c->vel->uploaded, "ufe_b200_calc_dHi_dt: no velocities given and none resident on the device");
    UFE_TRY(halo_exchange(c, M.b, c->vel->u.p)); UFE_TRY(halo_exchange(c, M.b, c->vel->v.p));
    u_b = c->vel->u.p; v_b = c->vel->v.p;
  }
  int its = 0;
  UFE_TRY(thickness_step(c, cfg, u_b, v_b, dt, finish, &its));
  {
    UFE_REGION(c, TC_D2H);
    UFE_TRY(th_d2h(c, io->AMB, T.AMB.p, M.a)); UFE_TRY(th_d2h(c, io->dHi_dt, T.dHi_dt.p, M.a));
    UFE_TRY(th_d2h(c, io->Hi_tplusdt, T.Hi_new.p, M.a)); UFE_TRY(th_d2h(c, io->divQ, T.divQ.p, M.a));
  }
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  if (n_Axb_its) *n_Axb_its = its;
  return 0;
}

// tail of calc_dHi_dt (conservation_of_mass_main.f90:83-109) for callers that ran the spill-over flux on the host
int ufe_b200_dHi_dt_finish(ufe_b200_handle h, const double* Hi, double* Hi_tplusdt, double dt, double* AMB, double* dHi_dt) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && Hi && Hi_tplusdt && AMB && dHi_dt, "NULL argument");
  UFE_CUDA(cudaSetDevice(c->device));
  UFE_TRY(ensure_thickness(c));
  Mesh& M = *c->mesh; Thickness& T = *c->thick;
  UFE_TRY(th_h2d(c, T.Hi.p, M.a, Hi)); UFE_TRY(th_h2d(c, T.Hi_new.p, M.a, (const double*)Hi_tplusdt));
  UFE_TRY(th_h2d(c, T.AMB.p, M.a, (const double*)AMB)); UFE_TRY(th_h2d(c, T.dHi_dt.p, M.a, (const double*)dHi_dt));
  {
    UFE_LAUNCH(c, TC_THICKNESS);
    k_th_rates<<<grid_for(c, M.a.n_loc(), 128, 16), 128, 0, c->stream>>>(M.a.n_loc(), M.a.off(), dt, dt, 0, 1, T.Hi.p, T.Hi_new.p, T.dHi_dt.p, T.AMB.p);
  }
  UFE_CUDA(cudaGetLastError());
  UFE_TRY(th_d2h(c, Hi_tplusdt, T.Hi_new.p, M.a)); UFE_TRY(th_d2h(c, AMB, T.AMB.p, M.a)); UFE_TRY(th_d2h(c, dHi_dt, T.dHi_dt.p, M.a));
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

}  // extern "C"
