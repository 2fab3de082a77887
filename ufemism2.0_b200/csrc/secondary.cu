// calc_secondary_velocities on the device: surface / basal / vertically averaged velocities on the b-grid, their a-grid
// maps (incl. the full 3-D field), absolute values and the slide/shear ratio.
// Reference: src/UFEMISM/ice_dynamics/conservation_of_momentum/conservation_of_momentum_main.f90:177-246 -- one loop over triangles,
// 2 map_b_a_3D + 6 map_b_a_2D SpMVs (the matrix re-read 2 nz + 6 times), two loops over vertices.  Here: one triangle kernel (a column read
// of u_3D_b, v_3D_b) and one vertex kernel that gathers each neighbouring triangle once per layer for all outputs.
#include <cstring>

#include "mesh.cuh"
#include "velocity.cuh"

namespace ufe {

struct Secondary {
  DBuf<double> b2d[9];        // u_surf_b, v_surf_b, uabs_surf_b, u_base_b, v_base_b, uabs_base_b, u_vav_b, v_vav_b, uabs_vav_b (extended b range)
  DBuf<double> u_3D, v_3D;    // (nz, a range)
  DBuf<double> a2d[10];       // u_surf, v_surf, u_base, v_base, u_vav, v_vav, uabs_surf, uabs_base, uabs_vav, R_shear
  bool have_3D = false;
  // calc_vertical_velocities
  DBuf<double> in[6];         // Hi, Hs, Hib, dHb_dt, dHi_dt, BMB (extended a range)
  DBuf<int> mask[2];          // mask_grounded_ice, mask_floating_ice
  DBuf<double> w_3D, dw_dz_3D, dz_coef;
  DBuf<int> dz_nb;
};
void secondary_free(Secondary* s) { delete s; }

namespace {

struct SecB { double* p[9]; };
struct SecA { double* p[10]; };

__global__ void k_sec_tri(int n, int off, int nb_ext, int nz, const double* __restrict__ zeta, const double* __restrict__ u_3D_b,
                          const double* __restrict__ v_3D_b, SecB o) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const size_t i = off + r;
    double us = 0.0, vs = 0.0, ub = 0.0, vb = 0.0, ua = 0.0, va = 0.0, up = 0.0, vp = 0.0;
    for (int k = 0; k < nz; ++k) {
      const double u = u_3D_b[(size_t)k * nb_ext + i], v = v_3D_b[(size_t)k * nb_ext + i];
      if (k == 0) { us = u; vs = v; }
      else { ua = ua + 0.5 * (u + up) * (zeta[k] - zeta[k - 1]); va = va + 0.5 * (v + vp) * (zeta[k] - zeta[k - 1]); }   // vertical_average
      if (k == nz - 1) { ub = u; vb = v; }
      up = u; vp = v;
    }
    o.p[0][i] = us; o.p[1][i] = vs; o.p[2][i] = sqrt(us * us + vs * vs);
    o.p[3][i] = ub; o.p[4][i] = vb; o.p[5][i] = sqrt(ub * ub + vb * vb);
    o.p[6][i] = ua; o.p[7][i] = va; o.p[8][i] = sqrt(ua * ua + va * va);
  }
}

// map_b_a of the 3-D fields and of the six 2-D fields through the shared b->a pattern
__global__ void k_sec_vertex(int n, int off, int na_ext, int nb_ext, int nz, const int* __restrict__ ptr, const int* __restrict__ ind,
                             const double* __restrict__ w_map, const double* __restrict__ u_3D_b, const double* __restrict__ v_3D_b, SecB b,
                             double* __restrict__ u_3D, double* __restrict__ v_3D, SecA o) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const int k1 = ptr[r], k2 = ptr[r + 1];
    const size_t i = off + r;
    double us = 0.0, vs = 0.0, ub = 0.0, vb = 0.0;
    for (int l = 0; l < nz; ++l) {
      double u = 0.0, v = 0.0;
      for (int k = k1; k < k2; ++k) {
        const size_t j = (size_t)l * nb_ext + ind[k];
        u += w_map[k] * u_3D_b[j]; v += w_map[k] * v_3D_b[j];
      }
      u_3D[(size_t)l * na_ext + i] = u; v_3D[(size_t)l * na_ext + i] = v;
      if (l == 0) { us = u; vs = v; }           // map_b_a_2D( u_surf_b) is the same sum as layer 1 of map_b_a_3D( u_3D_b)
      if (l == nz - 1) { ub = u; vb = v; }
    }
    double ua = 0.0, va = 0.0;
    for (int k = k1; k < k2; ++k) { ua += w_map[k] * b.p[6][ind[k]]; va += w_map[k] * b.p[7][ind[k]]; }
    o.p[0][i] = us; o.p[1][i] = vs; o.p[2][i] = ub; o.p[3][i] = vb; o.p[4][i] = ua; o.p[5][i] = va;
    const double as = sqrt(us * us + vs * vs), ab = sqrt(ub * ub + vb * vb);
    o.p[6][i] = as; o.p[7][i] = ab; o.p[8][i] = sqrt(ua * ua + va * va);
    o.p[9][i] = (ab + 0.1) / (as + 0.1);        // slide/shear ratio
  }
}

// ---- calc_vertical_velocities (conservation_of_mass/vertical_velocities.f90:23-243) -----------------------------------------------------
// One thread per vertex column.  The reference materialises u_3D_c, v_3D_c (nE x nz, allgathered) and the zeta gradients (6 SpMVs into
// nV x nz fields); here the b->c map is folded into the Voronoi-boundary flux (each edge's two triangles are read once per layer and
// carried to the next), the six a->a gradients are one pass over the shared ddx/ddy pattern, and dw/dz uses the column just written.
struct VvMesh {
  int nV, nE, a_lo, b_lo, vi1;
  const int *nC, *C, *VE, *ETri;
  const double *V, *A, *Cw, *zeta;
};
struct VvIn { const double* p[6]; const int* mask[2]; };

// calc_shape_functions_1D_reg_2nd_order (src/UPSY/basic/math_utilities/shape_functions.f90:17-79) with two neighbours -> the rows of
// M_ddzeta_k_k_1D (calc_vertical_operators_reg_1D, src/UPSY/mesh/mesh_zeta.f90:300-376); nz rows, done on the host
void ddzeta_rows(const std::vector<double>& zeta, std::vector<int>& nb, std::vector<double>& coef) {
  const int nz = (int)zeta.size();
  nb.assign((size_t)nz * 2, 0); coef.assign((size_t)nz * 3, 0.0);
  for (int k = 0; k < nz; ++k) {
    const int i0 = (k == 0) ? 1 : (k == nz - 1) ? nz - 3 : k - 1, i1 = (k == 0) ? 2 : (k == nz - 1) ? nz - 2 : k + 1;
    const double dx[2] = {zeta[i0] - zeta[k], zeta[i1] - zeta[k]};
    double w2[2], a11 = 0, a12 = 0, a21 = 0, a22 = 0;
    for (int ci = 0; ci < 2; ++ci) {
      const double w = 1.0 / std::pow(std::fabs(dx[ci]), 1.5);
      w2[ci] = w * w;
      a11 += w2[ci] * dx[ci] * dx[ci];
      a12 += w2[ci] * dx[ci] * 1 / 2 * (dx[ci] * dx[ci]);
      a21 += w2[ci] * 1 / 2 * (dx[ci] * dx[ci]) * dx[ci];
      a22 += w2[ci] * 1 / 2 * (dx[ci] * dx[ci]) * 1 / 2 * (dx[ci] * dx[ci]);
    }
    const double det = a11 * a22 - a12 * a21, m11 = a22 / det, m12 = -a12 / det;
    double nc[2];
    for (int ci = 0; ci < 2; ++ci) nc[ci] = w2[ci] * ((m11 * dx[ci]) + (m12 * 1 / 2 * (dx[ci] * dx[ci])));
    nb[2 * k] = i0; nb[2 * k + 1] = i1;
    coef[3 * k] = -(nc[0] + nc[1]); coef[3 * k + 1] = nc[0]; coef[3 * k + 2] = nc[1];
  }
}

__global__ void __launch_bounds__(128) k_vv_column(VvMesh m, int n, int off, int na_ext, int nb_ext, int nz, double rho_ratio,
                                                   const int* __restrict__ ptr, const int* __restrict__ ind, const double* __restrict__ ddx,
                                                   const double* __restrict__ ddy, VvIn in, const double* __restrict__ u_b,
                                                   const double* __restrict__ v_b, const double* __restrict__ u_a, const double* __restrict__ v_a,
                                                   const int* __restrict__ dz_nb, const double* __restrict__ dz_coef, double* __restrict__ w,
                                                   double* __restrict__ dw_dz) {
  const double* __restrict__ Hi = in.p[0]; const double* __restrict__ Hs = in.p[1]; const double* __restrict__ Hib = in.p[2];
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const size_t i = off + r;
    const int vi = m.vi1 + (int)r;
    const bool grounded = in.mask[0][i] != 0, floating = in.mask[1][i] != 0;
    if (!(grounded || floating)) {                                     // :150-153, calc_dw_dz :232-233
      for (int k = 0; k < nz; ++k) { w[(size_t)k * na_ext + i] = 0.0; dw_dz[(size_t)k * na_ext + i] = nan(""); }
      continue;
    }
    double dHib_dx = 0.0, dHib_dy = 0.0, dHs_dx = 0.0, dHs_dy = 0.0, dHi_dx = 0.0, dHi_dy = 0.0;
    for (int k = ptr[r]; k < ptr[r + 1]; ++k) {
      const int j = ind[k];
      const double ax = ddx[k], ay = ddy[k];
      dHib_dx += ax * Hib[j]; dHib_dy += ay * Hib[j];
      dHs_dx += ax * Hs[j]; dHs_dy += ay * Hs[j];
      dHi_dx += ax * Hi[j]; dHi_dy += ay * Hi[j];
    }
    const double dHib_dt = grounded ? in.p[3][i] : -in.p[4][i] * rho_ratio;   // :93-107
    const size_t ib = (size_t)(nz - 1) * na_ext + i;
    const double wb = (u_a[ib] * dHib_dx) + (v_a[ib] * dHib_dy) + dHib_dt + fmin(0.0, in.p[5][i]);   // :125-133
    w[ib] = wb;
    if (Hi[i] < 10.0) {                                                 // :136-139
      for (int k = 0; k < nz; ++k) w[(size_t)k * na_ext + i] = wb;
    } else {
      double cint[kMaxNz];
      for (int k = 0; k < nz - 1; ++k) cint[k] = 0.0;
      for (int ci = 0; ci < m.nC[vi - 1]; ++ci) {                       // :153-166, edge velocities as map_velocities_from_b_to_c_3D
        const size_t kk = (size_t)ci * m.nV + vi - 1;
        const int vj = m.C[kk], ei = m.VE[kk];
        int til = m.ETri[ei - 1], tir = m.ETri[(size_t)m.nE + ei - 1];
        const bool one = (til == 0 || tir == 0);
        if (til == 0) til = tir;
        if (tir == 0) tir = til;
        const double dS = m.Cw[kk];
        double nx = m.V[vj - 1] - m.V[vi - 1], ny = m.V[(size_t)m.nV + vj - 1] - m.V[(size_t)m.nV + vi - 1];
        const double nn = sqrt(nx * nx + ny * ny);
        nx = nx / nn; ny = ny / nn;
        double up = 0.0, vp = 0.0;
        for (int k = nz - 1; k >= 0; --k) {
          const size_t o = (size_t)k * nb_ext;
          const double ul = u_b[o + til - m.b_lo], vl = v_b[o + til - m.b_lo];
          double uc = ul, vc = vl;
          if (!one) { uc = (ul + u_b[o + tir - m.b_lo]) / 2.0; vc = (vl + v_b[o + tir - m.b_lo]) / 2.0; }
          if (k < nz - 1) {
            const double u_ks = 0.5 * (uc + up), v_ks = 0.5 * (vc + vp);
            cint[k] = cint[k] + (u_ks * nx + v_ks * ny) * dS;
          }
          up = uc; vp = vc;
        }
      }
      const double Hc = fmax(10.0, Hi[i]), A_i = m.A[vi - 1];
      double wk = wb, ua_p = u_a[ib], va_p = v_a[ib];
      for (int ks = nz - 2; ks >= 0; --ks) {                            // :143-188
        const double z0 = m.zeta[ks], z1 = m.zeta[ks + 1], dzeta = z1 - z0;
        const double grad_uv = cint[ks] / A_i;
        const double ua = u_a[(size_t)ks * na_ext + i], va = v_a[(size_t)ks * na_ext + i];
        const double du_dzeta = (ua_p - ua) / dzeta, dv_dzeta = (va_p - va) / dzeta;
        // zeta gradients (zeta_gradients.f90:96-113) at the two layers, then their staggered mean (:174-176)
        const double zx = 0.5 * ((1.0 / Hc) * (dHs_dx - z0 * dHi_dx) + (1.0 / Hc) * (dHs_dx - z1 * dHi_dx));
        const double zy = 0.5 * ((1.0 / Hc) * (dHs_dy - z0 * dHi_dy) + (1.0 / Hc) * (dHs_dy - z1 * dHi_dy));
        const double zz = 0.5 * ((-1.0 / Hc) + (-1.0 / Hc));
        const double dw_dzeta = -1.0 / zz * (grad_uv + zx * du_dzeta + zy * dv_dzeta);
        wk = wk - dzeta * dw_dzeta;
        w[(size_t)ks * na_ext + i] = wk;
        ua_p = ua; va_p = va;
      }
    }
    for (int k = 0; k < nz; ++k) {                                      // calc_dw_dz :226-239
      const double d = dz_coef[3 * k] * w[(size_t)k * na_ext + i] + dz_coef[3 * k + 1] * w[(size_t)dz_nb[2 * k] * na_ext + i] +
                       dz_coef[3 * k + 2] * w[(size_t)dz_nb[2 * k + 1] * na_ext + i];
      dw_dz[(size_t)k * na_ext + i] = -1.0 / Hi[i] * d;
    }
  }
}

}  // namespace

}  // namespace ufe

using namespace ufe;

extern "C" int ufe_b200_calc_secondary_velocities(ufe_b200_handle h, const double* u_3D_b, const double* v_3D_b, const ufe_b200_secondary_velocities* out) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && out, "ufe_b200_calc_secondary_velocities: NULL argument");
  UFE_CHECK((u_3D_b == nullptr) == (v_3D_b == nullptr), "ufe_b200_calc_secondary_velocities: give both u_3D_b and v_3D_b or neither");
  UFE_CUDA(cudaSetDevice(c->device));
  UFE_TRY(ensure_velocity(c));
  Mesh& M = *c->mesh; Velocity& V = *c->vel;
  UFE_CHECK(u_3D_b || V.uploaded || (c->sec && c->sec->have_3D), "ufe_b200_calc_secondary_velocities: no 3-D velocities given and none resident on the device");
  cudaStream_t st = c->stream;
  const int nz = M.nz, na = M.a.n_loc(), nb = M.b.n_loc();
  const size_t na_ext = M.a.n_ext(), nb_ext = M.b.n_ext();
  if (!c->sec) c->sec = new Secondary();
  Secondary& S = *c->sec;
  for (auto& b : S.b2d) UFE_TRY(b.alloc(nb_ext));
  for (auto& a : S.a2d) UFE_TRY(a.alloc(na_ext));
  UFE_TRY(S.u_3D.alloc(na_ext * nz)); UFE_TRY(S.v_3D.alloc(na_ext * nz));
  if (u_3D_b) {
    UFE_REGION(c, TC_H2D);
    UFE_CUDA(cudaMemcpy2DAsync(V.u_3D.p + M.b.off(), sizeof(double) * nb_ext, u_3D_b, sizeof(double) * nb, sizeof(double) * nb, nz, cudaMemcpyHostToDevice, st));
    UFE_CUDA(cudaMemcpy2DAsync(V.v_3D.p + M.b.off(), sizeof(double) * nb_ext, v_3D_b, sizeof(double) * nb, sizeof(double) * nb, nz, cudaMemcpyHostToDevice, st));
  }
  SecB sb; SecA sa;
  for (int q = 0; q < 9; ++q) sb.p[q] = S.b2d[q].p;
  for (int q = 0; q < 10; ++q) sa.p[q] = S.a2d[q].p;
  {
    UFE_LAUNCH(c, TC_COLUMN);
    k_sec_tri<<<grid_for(c, nb, 128, 16), 128, 0, st>>>(nb, M.b.off(), (int)nb_ext, nz, M.zeta.p, V.u_3D.p, V.v_3D.p, sb);
  }
  // the vertex kernel reads the triangles around each owned vertex: halos of the 3-D fields and of the vertical averages
  UFE_TRY(halo_exchange(c, M.b, V.u_3D.p, nz)); UFE_TRY(halo_exchange(c, M.b, V.v_3D.p, nz));
  UFE_TRY(halo_exchange(c, M.b, S.b2d[6].p)); UFE_TRY(halo_exchange(c, M.b, S.b2d[7].p));
  const Mesh::OpGroup& gba = M.grp[Mesh::G_BA];
  {
    UFE_LAUNCH(c, TC_COLUMN);
    k_sec_vertex<<<grid_for(c, na, 128, 16), 128, 0, st>>>(na, M.a.off(), (int)na_ext, (int)nb_ext, nz, gba.ptr.p, gba.ind.p, gba.val[0].p, V.u_3D.p, V.v_3D.p,
                                                           sb, S.u_3D.p, S.v_3D.p, sa);
  }
  UFE_CUDA(cudaGetLastError());
  {
    UFE_REGION(c, TC_D2H);
    double* const hb[9] = {out->u_surf_b, out->v_surf_b, out->uabs_surf_b, out->u_base_b, out->v_base_b, out->uabs_base_b, out->u_vav_b, out->v_vav_b, out->uabs_vav_b};
    double* const ha[10] = {out->u_surf, out->v_surf, out->u_base, out->v_base, out->u_vav, out->v_vav, out->uabs_surf, out->uabs_base, out->uabs_vav, out->R_shear};
    for (int q = 0; q < 9; ++q)
      if (hb[q]) UFE_CUDA(cudaMemcpyAsync(hb[q], S.b2d[q].p + M.b.off(), sizeof(double) * nb, cudaMemcpyDeviceToHost, st));
    for (int q = 0; q < 10; ++q)
      if (ha[q]) UFE_CUDA(cudaMemcpyAsync(ha[q], S.a2d[q].p + M.a.off(), sizeof(double) * na, cudaMemcpyDeviceToHost, st));
    if (out->u_3D) UFE_CUDA(cudaMemcpy2DAsync(out->u_3D, sizeof(double) * na, S.u_3D.p + M.a.off(), sizeof(double) * na_ext, sizeof(double) * na, nz, cudaMemcpyDeviceToHost, st));
    if (out->v_3D) UFE_CUDA(cudaMemcpy2DAsync(out->v_3D, sizeof(double) * na, S.v_3D.p + M.a.off(), sizeof(double) * na_ext, sizeof(double) * na, nz, cudaMemcpyDeviceToHost, st));
  }
  UFE_CUDA(cudaStreamSynchronize(st));
  S.have_3D = true;
  return 0;
}

// ---- set_ice_velocities_to_{SIA,SSA,SIASSA,DIVA}_results (conservation_of_momentum_main.f90:311-511) -------------------------------------
namespace ufe { namespace {
// ice%u_3D_b( ti,:) from the solver results, in place in the device's u_3D_b / v_3D_b:
//   mode 0 (SSA)                  u_3D_b( ti,k) = SSA%u_b( ti)                                              (:361-364)
//   mode 1 (SIA/SSA, 'add')       u_3D_b( ti,k) = SIA%u_3D_b( ti,k) + SSA%u_b( ti)                          (:405-408)
//   mode 2 ('add_SIA_reduced')    u_3D_b( ti,k) = w_sia_u SIA%u_3D_b( ti,k) + SSA%u_b( ti),
//                                 w_sia_u = 1 - (2/pi) atan( |u_b|^2 / 30^2), likewise v                     (:430-435)
__global__ void k_set_velocities_b(int mode, int n, int off, size_t nb_ext, int nz, const double* __restrict__ u, const double* __restrict__ v,
                                   double* __restrict__ u_3D, double* __restrict__ v_3D) {
  const double pi = 3.141592653589793238462643383279502884197;   // parameters.f90:17
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const size_t i = off + r;
    const double ub = u[i], vb = v[i];
    double wu = 1.0, wv = 1.0;
    if (mode == 0) { wu = 0.0; wv = 0.0; }
    else if (mode == 2) {
      wu = 1.0 - (2.0 / pi) * atan((fabs(ub) * fabs(ub)) / (30.0 * 30.0));
      wv = 1.0 - (2.0 / pi) * atan((fabs(vb) * fabs(vb)) / (30.0 * 30.0));
    }
    for (int k = 0; k < nz; ++k) {
      const size_t j = (size_t)k * nb_ext + i;
      if (mode == 0) { u_3D[j] = ub; v_3D[j] = vb; }
      else { u_3D[j] = wu * u_3D[j] + ub; v_3D[j] = wv * v_3D[j] + vb; }
    }
  }
}
// ice%du_dx_3D( vi,:) = SSA%du_dx_a( vi) and friends: one 2-D a-grid field into every layer
__global__ void k_layers_from_2D(int n, int off, size_t na_ext, int nz, const double* __restrict__ a2d, double* __restrict__ a3d) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const double x = a2d[off + r];
    for (int k = 0; k < nz; ++k) a3d[(size_t)k * na_ext + off + r] = x;
  }
}
} }  // namespace ufe::<anon>

extern "C" int ufe_b200_set_ice_velocities(ufe_b200_handle h, int approximation, int hybrid_scheme, const ufe_b200_ice_velocities* out) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && out, "ufe_b200_set_ice_velocities: NULL argument");
  UFE_CHECK(approximation >= UFE_APPROX_SSA && approximation <= UFE_APPROX_SIASSA, "unknown choice_stress_balance_approximation %d", approximation);
  UFE_CUDA(cudaSetDevice(c->device));
  UFE_CHECK(c->vel && c->vel->uploaded, "ufe_b200_set_ice_velocities: no solver results on the device");
  Mesh& M = *c->mesh; Velocity& V = *c->vel;
  cudaStream_t st = c->stream;
  const int nz = M.nz, na = M.a.n_loc(), nb = M.b.n_loc();
  const size_t na_ext = M.a.n_ext(), nb_ext = M.b.n_ext();
  int mode = -1;   // what happens to the device's u_3D_b, v_3D_b
  switch (approximation) {
    case UFE_APPROX_SIA:  UFE_CHECK(V.last_approximation == UFE_APPROX_SIA, "set_ice_velocities_to_SIA_results: the last solve was not solve_SIA"); break;
    case UFE_APPROX_DIVA: UFE_CHECK(V.last_approximation == UFE_APPROX_DIVA, "set_ice_velocities_to_DIVA_results: the last solve was not solve_DIVA"); break;
    case UFE_APPROX_SSA:  UFE_CHECK(V.last_approximation == UFE_APPROX_SSA, "set_ice_velocities_to_SSA_results: the last solve was not solve_SSA"); mode = 0; break;
    default:
      UFE_CHECK(V.have_sia && V.last_approximation == UFE_APPROX_SSA, "set_ice_velocities_to_SIASSA_results: call ufe_b200_solve_SIA, then ufe_b200_solve_SSA");
      UFE_CHECK(hybrid_scheme == UFE_HYBRID_ADD || hybrid_scheme == UFE_HYBRID_ADD_SIA_REDUCED, "unknown choice_hybrid_SIASSA_scheme_config %d", hybrid_scheme);
      mode = hybrid_scheme == UFE_HYBRID_ADD ? 1 : 2;
      V.have_sia = false;   // u_3D_b, v_3D_b now hold the hybrid velocities; the next hybrid step starts from a new solve_SIA
  }
  if (mode >= 0) {
    UFE_LAUNCH(c, TC_TRI);
    k_set_velocities_b<<<grid_for(c, nb, 128, 16), 128, 0, st>>>(mode, nb, M.b.off(), nb_ext, nz, V.u.p, V.v.p, V.u_3D.p, V.v_3D.p);
    UFE_CUDA(cudaGetLastError());
  }
  auto d2h3 = [&](double* dst, const double* src, size_t n_ext, int off, int n_loc) {
    return dst ? cudaMemcpy2DAsync(dst, sizeof(double) * n_loc, src + off, sizeof(double) * n_ext, sizeof(double) * n_loc, nz, cudaMemcpyDeviceToHost, st)
               : cudaSuccess;
  };
  { UFE_REGION(c, TC_D2H); UFE_CUDA(d2h3(out->u_3D_b, V.u_3D.p, nb_ext, M.b.off(), nb)); UFE_CUDA(d2h3(out->v_3D_b, V.v_3D.p, nb_ext, M.b.off(), nb)); }
  // horizontal strain rates: zero in the SIA, the solver's a-grid fields in every layer otherwise
  double* const h2d_out[4] = {out->du_dx_3D, out->du_dy_3D, out->dv_dx_3D, out->dv_dy_3D};
  const double* const src2d[4] = {V.du_dx.p, V.du_dy.p, V.dv_dx.p, V.dv_dy.p};
  DBuf<double> layers;
  for (int q = 0; q < 4; ++q) {
    if (!h2d_out[q]) continue;
    if (approximation == UFE_APPROX_SIA) { memset(h2d_out[q], 0, sizeof(double) * (size_t)na * nz); continue; }
    UFE_CHECK(V.have_strain_2d, "ufe_b200_set_ice_velocities: no SSA / DIVA solve has produced strain rates yet");
    UFE_TRY(layers.alloc(na_ext * nz));
    { UFE_LAUNCH(c, TC_COLUMN); k_layers_from_2D<<<grid_for(c, na, 128, 16), 128, 0, st>>>(na, M.a.off(), na_ext, nz, src2d[q], layers.p); }
    UFE_CUDA(cudaGetLastError());
    UFE_REGION(c, TC_D2H);
    UFE_CUDA(d2h3(h2d_out[q], layers.p, na_ext, M.a.off(), na));
    UFE_CUDA(cudaStreamSynchronize(st));   // the scratch buffer is reused by the next field
  }
  // vertical shear: the SIA's (also in the hybrid), DIVA's own, zero in the SSA
  if (out->du_dz_3D || out->dv_dz_3D) {
    if (approximation == UFE_APPROX_SSA) {
      if (out->du_dz_3D) memset(out->du_dz_3D, 0, sizeof(double) * (size_t)na * nz);
      if (out->dv_dz_3D) memset(out->dv_dz_3D, 0, sizeof(double) * (size_t)na * nz);
    } else {
      UFE_CHECK(V.have_dudz_3d, "ufe_b200_set_ice_velocities: du_dz_3D is not on the device (give du_dz_3D_a, dv_dz_3D_a to ufe_b200_solve_SIA)");
      UFE_REGION(c, TC_D2H);
      UFE_CUDA(d2h3(out->du_dz_3D, V.du_dz_3D.p, na_ext, M.a.off(), na)); UFE_CUDA(d2h3(out->dv_dz_3D, V.dv_dz_3D.p, na_ext, M.a.off(), na));
    }
  }
  if (out->dw_dx_3D) memset(out->dw_dx_3D, 0, sizeof(double) * (size_t)na * nz);   // all four routines neglect the gradients of w
  if (out->dw_dy_3D) memset(out->dw_dy_3D, 0, sizeof(double) * (size_t)na * nz);
  UFE_CUDA(cudaStreamSynchronize(st));
  return 0;
}

extern "C" int ufe_b200_calc_vertical_velocities(ufe_b200_handle h, const ufe_b200_config* cfg, const ufe_b200_vertical_velocity_io* io) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && cfg && io, "ufe_b200_calc_vertical_velocities: NULL argument");
  UFE_CHECK(io->Hi && io->Hs && io->Hib && io->dHb_dt && io->dHi_dt && io->BMB && io->mask_grounded_ice && io->mask_floating_ice && io->w_3D,
            "ufe_b200_calc_vertical_velocities: NULL field");
  UFE_CUDA(cudaSetDevice(c->device));
  UFE_TRY(ensure_velocity(c));
  UFE_CHECK(c->sec && c->sec->have_3D, "ufe_b200_calc_vertical_velocities: call ufe_b200_calc_secondary_velocities first (u_3D, v_3D on the a-grid)");
  Mesh& M = *c->mesh; Velocity& V = *c->vel; Secondary& S = *c->sec;
  cudaStream_t st = c->stream;
  const int nz = M.nz, na = M.a.n_loc();
  const size_t na_ext = M.a.n_ext(), nb_ext = M.b.n_ext();
  UFE_CHECK(nz >= 3 && nz <= kMaxNz, "ufe_b200_calc_vertical_velocities: nz out of range");
  for (auto& b : S.in) UFE_TRY(b.alloc(na_ext));
  for (auto& b : S.mask) UFE_TRY(b.alloc(na_ext));
  UFE_TRY(S.w_3D.alloc(na_ext * nz)); UFE_TRY(S.dw_dz_3D.alloc(na_ext * nz));
  if (!S.dz_nb.p) {
    std::vector<int> nb; std::vector<double> coef;
    ddzeta_rows(M.h_zeta, nb, coef);
    UFE_TRY(S.dz_nb.alloc(nb.size())); UFE_TRY(S.dz_coef.alloc(coef.size()));
    UFE_CUDA(cudaMemcpyAsync(S.dz_nb.p, nb.data(), sizeof(int) * nb.size(), cudaMemcpyHostToDevice, st));
    UFE_CUDA(cudaMemcpyAsync(S.dz_coef.p, coef.data(), sizeof(double) * coef.size(), cudaMemcpyHostToDevice, st));
    UFE_CUDA(cudaStreamSynchronize(st));   // nb, coef are stack vectors
  }
  {
    UFE_REGION(c, TC_H2D);
    const double* const hin[6] = {io->Hi, io->Hs, io->Hib, io->dHb_dt, io->dHi_dt, io->BMB};
    for (int q = 0; q < 6; ++q) UFE_CUDA(cudaMemcpyAsync(S.in[q].p + M.a.off(), hin[q], sizeof(double) * na, cudaMemcpyHostToDevice, st));
    UFE_CUDA(cudaMemcpyAsync(S.mask[0].p + M.a.off(), io->mask_grounded_ice, sizeof(int) * na, cudaMemcpyHostToDevice, st));
    UFE_CUDA(cudaMemcpyAsync(S.mask[1].p + M.a.off(), io->mask_floating_ice, sizeof(int) * na, cudaMemcpyHostToDevice, st));
  }
  for (int q = 0; q < 3; ++q) UFE_TRY(halo_exchange(c, M.a, S.in[q].p));     // the a->a gradients read neighbouring vertices
  VvMesh m{M.nV, M.nE, M.a.lo, M.b.lo, M.a.i1, M.nC.p, M.C.p, M.VE.p, M.ETri.p, M.V.p, M.A.p, M.Cw.p, M.zeta.p};
  VvIn in;
  for (int q = 0; q < 6; ++q) in.p[q] = S.in[q].p;
  in.mask[0] = S.mask[0].p; in.mask[1] = S.mask[1].p;
  const Mesh::OpGroup& gaa = M.grp[Mesh::G_AA];
  {
    UFE_LAUNCH(c, TC_COLUMN);
    k_vv_column<<<grid_for(c, na, 128, 16), 128, 0, st>>>(m, na, M.a.off(), (int)na_ext, (int)nb_ext, nz, cfg->ice_density / cfg->seawater_density, gaa.ptr.p,
                                                          gaa.ind.p, gaa.val[0].p, gaa.val[1].p, in, V.u_3D.p, V.v_3D.p, S.u_3D.p, S.v_3D.p, S.dz_nb.p,
                                                          S.dz_coef.p, S.w_3D.p, S.dw_dz_3D.p);
  }
  UFE_CUDA(cudaGetLastError());
  {
    UFE_REGION(c, TC_D2H);
    UFE_CUDA(cudaMemcpy2DAsync(io->w_3D, sizeof(double) * na, S.w_3D.p + M.a.off(), sizeof(double) * na_ext, sizeof(double) * na, nz, cudaMemcpyDeviceToHost, st));
    if (io->dw_dz_3D)
      UFE_CUDA(cudaMemcpy2DAsync(io->dw_dz_3D, sizeof(double) * na, S.dw_dz_3D.p + M.a.off(), sizeof(double) * na_ext, sizeof(double) * na, nz, cudaMemcpyDeviceToHost, st));
  }
  UFE_CUDA(cudaStreamSynchronize(st));
  return 0;
}
