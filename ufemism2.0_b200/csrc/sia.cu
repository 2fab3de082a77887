// solve_SIA on the device: the analytical shallow-ice solution with Glen's flow law (config 4, Halfar dome).
// Reference: src/UFEMISM/ice_dynamics/conservation_of_momentum/SIA/SIA_main.f90:65-158 (Bueler and Brown, 2009, Eqs. 12-13):
//   D(z) = -2 (rho g)^n |grad h|^(n-1) int_b^z A(T*) (h - zeta)^n dzeta,   u = D dh/dx,   v = D dh/dy
// The reference runs 7 SpMV-equivalents (map_a_b x2, ddx/ddy_a_a, ddx/ddy_a_b, map_a_b_3D over nz layers) plus a column loop; here one
// triangle kernel gathers its three vertices through the shared a->b pattern (map, ddx, ddy values) for every layer and integrates the
// column bottom-up in registers, and one vertex kernel does the a->a gradients and the vertical shear strain rates.  HBM-bound:
// ~ 8 (3 nz + 2) B gathered + 8 * 3 nz B written per triangle.
#include "mesh.cuh"
#include "velocity.cuh"

namespace ufe {

namespace {

__global__ void k_sia_tri(int n, int off, int na_ext, int nb_ext, int nz, const int* __restrict__ ptr, const int* __restrict__ ind,
                          const double* __restrict__ w_map, const double* __restrict__ w_ddx, const double* __restrict__ w_ddy,
                          const double* __restrict__ zeta, const double* __restrict__ Hi, const double* __restrict__ Hs,
                          const double* __restrict__ A_flow, double glen_n, double rg_n, double max_diff, double* __restrict__ u_3D,
                          double* __restrict__ v_3D, double* __restrict__ D_3D) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const int k1 = ptr[r], k2 = ptr[r + 1];
    double Hi_b = 0.0, Hs_b = 0.0, dx = 0.0, dy = 0.0;
    for (int k = k1; k < k2; ++k) {
      const int j = ind[k];
      Hi_b += w_map[k] * Hi[j]; Hs_b += w_map[k] * Hs[j]; dx += w_ddx[k] * Hs[j]; dy += w_ddy[k] * Hs[j];
    }
    const double fac = -2.0 * rg_n * pow(sqrt(dx * dx + dy * dy), glen_n - 1.0);
    // bottom-up trapezoid (integrate_from_zeta_is_one_to_zeta_is_zetap with z in the place of zeta, mesh_zeta.f90:175-206)
    double integral = 0.0, f_below = 0.0, z_below = 0.0;
    for (int l = nz - 1; l >= 0; --l) {
      double A_b = 0.0;
      for (int k = k1; k < k2; ++k) A_b += w_map[k] * A_flow[(size_t)l * na_ext + ind[k]];
      const double z = Hs_b - zeta[l] * Hi_b;
      const double f = A_b * pow(Hs_b - z, glen_n);
      if (l < nz - 1) integral -= 0.5 * (f_below + f) * (z_below - z);
      f_below = f; z_below = z;
      const double D = fmax(-max_diff, fac * integral);
      const size_t o = (size_t)l * nb_ext + off + r;
      if (D_3D) D_3D[o] = D;
      u_3D[o] = D * dx;
      v_3D[o] = D * dy;
    }
  }
}

__global__ void k_sia_vertex(int n, int off, int na_ext, int nz, const int* __restrict__ ptr, const int* __restrict__ ind,
                             const double* __restrict__ w_ddx, const double* __restrict__ w_ddy, const double* __restrict__ zeta,
                             const double* __restrict__ Hi, const double* __restrict__ Hs, const double* __restrict__ A_flow, double glen_n,
                             double rg_n, double* __restrict__ du_dz, double* __restrict__ dv_dz) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    double dx = 0.0, dy = 0.0;
    for (int k = ptr[r]; k < ptr[r + 1]; ++k) { dx += w_ddx[k] * Hs[ind[k]]; dy += w_ddy[k] * Hs[ind[k]]; }
    const size_t i = off + r;
    const double fac = -2.0 * rg_n * pow(sqrt(dx * dx + dy * dy), glen_n - 1.0);
    const double hs = Hs[i], hi = Hi[i];
    for (int l = 0; l < nz; ++l) {
      const double z = hs - zeta[l] * hi;
      const double cf = fac * A_flow[(size_t)l * na_ext + i] * pow(hs - z, glen_n);
      du_dz[(size_t)l * na_ext + i] = cf * dx;
      dv_dz[(size_t)l * na_ext + i] = cf * dy;
    }
  }
}

}  // namespace

}  // namespace ufe

using namespace ufe;

extern "C" int ufe_b200_solve_SIA(ufe_b200_handle h, const ufe_b200_config* cfg, double SIA_maximum_diffusivity, const double* Hi, const double* Hs,
                                  const double* A_flow, double* u_3D_b, double* v_3D_b, double* D_3D_b, double* du_dz_3D_a, double* dv_dz_3D_a) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && cfg && Hi && Hs && A_flow && u_3D_b && v_3D_b, "ufe_b200_solve_SIA: Hi, Hs, A_flow, u_3D_b and v_3D_b are required");
  UFE_CHECK((du_dz_3D_a == nullptr) == (dv_dz_3D_a == nullptr), "ufe_b200_solve_SIA: give both du_dz_3D_a and dv_dz_3D_a or neither");
  UFE_CUDA(cudaSetDevice(c->device));
  UFE_TRY(ensure_velocity(c));
  Mesh& M = *c->mesh; Velocity& V = *c->vel;
  cudaStream_t st = c->stream;
  const int nz = M.nz, na = M.a.n_loc(), nb = M.b.n_loc();
  const size_t na_ext = M.a.n_ext(), nb_ext = M.b.n_ext();
  auto h2d = [&](double* d, const double* s, int ncols) {
    return cudaMemcpy2DAsync(d + M.a.off(), sizeof(double) * na_ext, s, sizeof(double) * na, sizeof(double) * na, ncols, cudaMemcpyHostToDevice, st);
  };
  { UFE_REGION(c, TC_H2D); UFE_CUDA(h2d(V.Hi.p, Hi, 1)); UFE_CUDA(h2d(V.Hs.p, Hs, 1)); UFE_CUDA(h2d(V.A_flow.p, A_flow, nz)); }
  UFE_TRY(halo_exchange(c, M.a, V.Hi.p)); UFE_TRY(halo_exchange(c, M.a, V.Hs.p)); UFE_TRY(halo_exchange(c, M.a, V.A_flow.p, nz));
  const double n = cfg->Glens_flow_law_exponent, rg_n = pow(cfg->ice_density * cfg->grav, n);
  const Mesh::OpGroup& gab = M.grp[Mesh::G_AB];
  const Mesh::OpGroup& gaa = M.grp[Mesh::G_AA];
  DBuf<double> D;   // D_3D_b is not part of the DIVA state: a scratch buffer, only when asked for
  if (D_3D_b) UFE_TRY(D.alloc(nb_ext * nz));
  {
    UFE_LAUNCH(c, TC_COLUMN);
    k_sia_tri<<<grid_for(c, nb, 128, 16), 128, 0, st>>>(nb, M.b.off(), (int)na_ext, (int)nb_ext, nz, gab.ptr.p, gab.ind.p, gab.val[0].p, gab.val[1].p,
                                                        gab.val[2].p, M.zeta.p, V.Hi.p, V.Hs.p, V.A_flow.p, n, rg_n, SIA_maximum_diffusivity,
                                                        V.u_3D.p, V.v_3D.p, D.p);
  }
  if (du_dz_3D_a) {
    UFE_LAUNCH(c, TC_COLUMN);
    k_sia_vertex<<<grid_for(c, na, 128, 16), 128, 0, st>>>(na, M.a.off(), (int)na_ext, nz, gaa.ptr.p, gaa.ind.p, gaa.val[0].p, gaa.val[1].p, M.zeta.p,
                                                           V.Hi.p, V.Hs.p, V.A_flow.p, n, rg_n, V.du_dz_3D.p, V.dv_dz_3D.p);
  }
  UFE_CUDA(cudaGetLastError());
  auto d2h = [&](double* dst, const double* src, size_t n_ext, int off, int n_loc) {
    return dst ? cudaMemcpy2DAsync(dst, sizeof(double) * n_loc, src + off, sizeof(double) * n_ext, sizeof(double) * n_loc, nz, cudaMemcpyDeviceToHost, st)
               : cudaSuccess;
  };
  {
    UFE_REGION(c, TC_D2H);
    UFE_CUDA(d2h(u_3D_b, V.u_3D.p, nb_ext, M.b.off(), nb)); UFE_CUDA(d2h(v_3D_b, V.v_3D.p, nb_ext, M.b.off(), nb));
    UFE_CUDA(d2h(D_3D_b, D.p, nb_ext, M.b.off(), nb));
    UFE_CUDA(d2h(du_dz_3D_a, V.du_dz_3D.p, na_ext, M.a.off(), na)); UFE_CUDA(d2h(dv_dz_3D_a, V.dv_dz_3D.p, na_ext, M.a.off(), na));
  }
  UFE_CUDA(cudaStreamSynchronize(st));
  V.uploaded = true;   // Hi, Hs, A_flow, u_3D, v_3D are resident now
  V.have_sia = true; V.last_approximation = UFE_APPROX_SIA; V.have_dudz_3d = du_dz_3D_a != nullptr;
  return 0;
}
