// Geometry / mask / sub-grid stage between the thickness and velocity solves, on the device.
// Reference (src/UFEMISM/ice_dynamics/utilities/):
//   determine_masks                               masks_mod.f90:25-217
//   calc_effective_thickness                      subgrid_ice_margin.f90:19-158
//   calc_grounded_fractions                       subgrid_grounded_fractions_main.f90:22-120
//   ..._bilin_interp_TAF_{a,b}, calc_grounded_area_triangle*    subgrid_grounded_fractions_bilin_TAF.f90:20-369
//   ..._bedrock_CDF_{a,b}                         subgrid_grounded_fractions_bedrock_CDF.f90:22-181
//   calc_critical_timestep_{SIA,adv}              time_step_criteria.f90:20-141
//   is_floating, thickness_above_floatation       src/UPSY/basic/math_utilities/ice_geometry_basics.f90:20-64
// The reference gathers every field to every process (15 gather_to_all over these routines); here one halo exchange per field that a
// neighbour stencil reads, and each routine is one or two passes over the vertices / triangles.
#include "geometry.cuh"
#include "mesh.cuh"
#include "velocity.cuh"

namespace ufe {

struct Geometry {
  DBuf<double> a[10];      // vertex fields over the extended a range: Hi, Hb, SL, dHb, TAF, out0, out1, scratch, ...
  DBuf<double> b[3];       // triangle fields over the extended b range: TAF_b, fraction_gr_b, scratch
  DBuf<int> ia[3];         // kind (1..4) / margin flag / floating flag over the extended a range
  DBuf<int> masks;         // [11][n_loc] results of determine_masks
  DBuf<double> cdf, cdf_b; // (n_loc, nbins)
  DBuf<double> u, v;       // host-given velocities (extended b range)
  DBuf<double> D3;         // SIA diffusivity (nz, extended b range)
  DBuf<unsigned long long> dtbits;
  // ufe_b200_geometry_resident: inputs that stay on the device for the resident time step
  bool dynamic = false;
  int choice = 0, nbins = 0;
  DBuf<double> dHb_res, cdf_res, cdf_b_res;
  DBuf<int> flag;
};
void geometry_free(Geometry* g) { delete g; }

namespace {

struct GeoMesh {
  int nV, nTri, nE, a_lo, b_lo, vi1, ti1;
  const int *nC, *C, *niTri, *iTri, *VBI, *Tri, *TriBI, *VE, *ETri;
  const double *V, *Tricc;
};
GeoMesh geo_mesh(const Mesh& M) {
  return GeoMesh{M.nV, M.nTri, M.nE, M.a.lo, M.b.lo, M.a.i1, M.b.i1, M.nC.p, M.C.p, M.niTri.p, M.iTri.p, M.VBI.p, M.Tri.p, M.TriBI.p, M.VE.p,
                 M.ETri.p, M.V.p, M.Tricc.p};
}

__device__ __forceinline__ bool is_floating(double rw_over_ri, double Hi, double Hb, double SL) { return Hi < (SL - Hb) * rw_over_ri; }

// basic masks as one code per vertex: 1 ice-free land, 2 ice-free ocean, 3 grounded ice, 4 floating ice (masks_mod.f90:63-103), and the TAF
__global__ void k_geo_kind(int n, int off, double rw_over_ri, const double* __restrict__ Hi, const double* __restrict__ Hb,
                           const double* __restrict__ SL, int* __restrict__ kind, double* __restrict__ TAF) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const size_t i = off + r;
    const bool fl = is_floating(rw_over_ri, Hi[i], Hb[i], SL[i]);
    kind[i] = fl ? (Hi[i] > 0.0 ? 4 : 2) : (Hi[i] > 0.0 ? 3 : 1);
    if (TAF) TAF[i] = Hi[i] - fmax(0.0, (SL[i] - Hb[i]) * rw_over_ri);
  }
}

// transitional masks (masks_mod.f90:121-203); out = [11][n]: mask, land, ocean, grounded, floating, margin, gl_gr, gl_fl, cf_gr, cf_fl, coastline
__global__ void k_geo_masks(GeoMesh m, int n, int off, const int* __restrict__ kind, int* __restrict__ out) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const int vi = m.vi1 + (int)r;
    const int k = kind[off + r];
    bool nb_noice = false, nb_fl = false, nb_gr = false, nb_oc = false;
    for (int ci = 0; ci < m.nC[vi - 1]; ++ci) {
      const int kj = kind[m.C[(size_t)ci * m.nV + vi - 1] - m.a_lo];
      nb_noice |= kj <= 2; nb_fl |= kj == 4; nb_gr |= kj == 3; nb_oc |= kj == 2;
    }
    const bool margin = k >= 3 && nb_noice, gl_gr = k == 3 && nb_fl, gl_fl = k == 4 && nb_gr, cf_gr = k == 3 && nb_oc, cf_fl = k == 4 && nb_oc,
               coast = k == 1 && nb_oc;
    int mask = k;                          // the reference overwrites in this order, so the last true test wins
    if (margin) mask = 9;
    if (gl_gr) mask = 5;
    if (gl_fl) mask = 6;
    if (cf_gr) mask = 7;
    if (cf_fl) mask = 8;
    if (coast) mask = 10;
    const int v[11] = {mask, k == 1, k == 2, k == 3, k == 4, margin, gl_gr, gl_fl, cf_gr, cf_fl, coast};
#pragma unroll
    for (int q = 0; q < 11; ++q) out[(size_t)q * n + r] = v[q];
  }
}

// calc_effective_thickness, first pass: margin flag (subgrid_ice_margin.f90:53-61)
__global__ void k_geo_margin(GeoMesh m, int n, int off, const double* __restrict__ Hi, int* __restrict__ margin) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const int vi = m.vi1 + (int)r;
    bool mg = false;
    if (Hi[off + r] > 0.0)
      for (int ci = 0; ci < m.nC[vi - 1]; ++ci) mg |= Hi[m.C[(size_t)ci * m.nV + vi - 1] - m.a_lo] == 0.0;
    margin[off + r] = mg;
  }
}
// second pass (:87-146)
__global__ void k_geo_hi_eff(GeoMesh m, int n, int off, double rw_over_ri, const double* __restrict__ Hi, const double* __restrict__ Hb,
                             const double* __restrict__ SL, const int* __restrict__ margin, double* __restrict__ Hi_eff,
                             double* __restrict__ fraction_margin) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const int vi = m.vi1 + (int)r;
    const size_t i = off + r;
    const double H = Hi[i];
    const bool fl = is_floating(rw_over_ri, H, Hb[i], SL[i]);
    double he = H, fm = 1.0;
    if (fl && !(H > 0.0)) { he = 0.0; fm = 0.0; }
    if (margin[i] && fl) {
      double hmax = 0.0;
      for (int ci = 0; ci < m.nC[vi - 1]; ++ci) {
        const int j = m.C[(size_t)ci * m.nV + vi - 1] - m.a_lo;
        if (!margin[j]) hmax = fmax(hmax, Hi[j]);
      }
      if (hmax > H) { he = hmax; fm = H / hmax; }
    }
    Hi_eff[i] = he; fraction_margin[i] = fm;
  }
}

struct P2 { double x, y; };
__device__ __forceinline__ double tri_area(P2 p, P2 q, P2 r) { return fabs((q.x - p.x) * (r.y - p.y) - (q.y - p.y) * (r.x - p.x)) / 2.0; }
__device__ __forceinline__ double corner_area(P2 va, P2 vb, P2 vc, double ta, double tb, double tc) {
  const double lab = ta / (ta - tb), lac = ta / (ta - tc);
  const P2 pab{(va.x * (1.0 - lab)) + (vb.x * lab), (va.y * (1.0 - lab)) + (vb.y * lab)};
  const P2 pac{(va.x * (1.0 - lac)) + (vc.x * lac), (va.y * (1.0 - lac)) + (vc.y * lac)};
  return tri_area(va, pab, pac);
}
// calc_grounded_area_triangle (bilin_TAF.f90:244-321)
__device__ void grounded_area(P2 va, P2 vb, P2 vc, double TAFa, double TAFb, double TAFc, double& A_tot, double& A_grnd) {
  const double tol = 1e-9;
  A_tot = tri_area(va, vb, vc);
  const double a = TAFa >= 0.0 ? fmax(tol, TAFa) : fmin(-tol, TAFa);
  const double b = TAFb >= 0.0 ? fmax(tol, TAFb) : fmin(-tol, TAFb);
  const double c = TAFc >= 0.0 ? fmax(tol, TAFc) : fmin(-tol, TAFc);
  if (a >= 0 && b >= 0 && c >= 0) A_grnd = A_tot;
  else if (a <= 0 && b <= 0 && c <= 0) A_grnd = 0.0;
  else if (a >= 0 && b <= 0 && c <= 0) A_grnd = corner_area(va, vb, vc, a, b, c);
  else if (a <= 0 && b >= 0 && c <= 0) A_grnd = corner_area(vb, vc, va, b, c, a);
  else if (a <= 0 && b <= 0 && c >= 0) A_grnd = corner_area(vc, va, vb, c, a, b);
  else if (a <= 0 && b >= 0 && c >= 0) A_grnd = A_tot - corner_area(va, vb, vc, a, b, c);
  else if (a >= 0 && b <= 0 && c >= 0) A_grnd = A_tot - corner_area(vb, vc, va, b, c, a);
  else A_grnd = A_tot - corner_area(vc, va, vb, c, a, b);
}
__device__ __forceinline__ P2 vert(const GeoMesh& m, int vi) { return P2{m.V[vi - 1], m.V[(size_t)m.nV + vi - 1]}; }
__device__ __forceinline__ P2 tricc(const GeoMesh& m, int ti) { return P2{m.Tricc[ti - 1], m.Tricc[(size_t)m.nTri + ti - 1]}; }

// bedrock CDF lookup (bedrock_CDF.f90:49-81); cdf[(bin) * stride]
__device__ double cdf_fraction(const double* __restrict__ cdf, size_t stride, int nbins, double Hb_float) {
  if (Hb_float <= cdf[0]) return 1.0;
  if (Hb_float >= cdf[(size_t)(nbins - 1) * stride]) return 0.0;
  int iu = 1;
  while (cdf[(size_t)(iu - 1) * stride] < Hb_float) iu = iu + 1;
  const int il = iu - 1;
  const double cu = cdf[(size_t)(iu - 1) * stride], cl = cdf[(size_t)(il - 1) * stride];
  const double wl = (cu - Hb_float) / (cu - cl), wu = 1.0 - wl;
  const double f = 1.0 - ((double)(il - 1) * wl + (double)(iu - 1) * wu) / (double)(nbins - 1);
  return fmin(1.0, fmax(0.0, f));
}

// triangles: TAF_b = map_a_b TAF (for the vertex pass), the TAF fraction of the triangle itself, the CDF fraction, and their combination
__global__ void k_geo_frac_b(GeoMesh m, int n, int off, int choice, int nbins, double ri_over_rw, const int* __restrict__ ptr,
                             const int* __restrict__ ind, const double* __restrict__ w_map, const double* __restrict__ TAF,
                             const double* __restrict__ Hi, const double* __restrict__ SL, const double* __restrict__ dHb,
                             const int* __restrict__ kind, const double* __restrict__ cdf_b, double* __restrict__ TAF_b,
                             double* __restrict__ fraction_gr_b) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const int ti = m.ti1 + (int)r;
    double tb = 0.0, Hi_b = 0.0, SL_b = 0.0, dHb_b = 0.0;
    for (int k = ptr[r]; k < ptr[r + 1]; ++k) {
      const int j = ind[k];
      const double wk = w_map[k];
      tb += wk * TAF[j];
      if (choice != UFE_GF_BILIN_INTERP_TAF) { Hi_b += wk * Hi[j]; SL_b += wk * SL[j]; dHb_b += wk * dHb[j]; }
    }
    TAF_b[off + r] = tb;
    const int via = m.Tri[ti - 1], vib = m.Tri[(size_t)m.nTri + ti - 1], vic = m.Tri[2 * (size_t)m.nTri + ti - 1];
    const double a = TAF[via - m.a_lo], b = TAF[vib - m.a_lo], c = TAF[vic - m.a_lo];
    double f_taf = 0.0, f_cdf = 0.0;
    if (choice != UFE_GF_BEDROCK_CDF) {
      const double tmax = fmax(a, fmax(b, c)), tmin = fmin(a, fmin(b, c));
      if (tmin >= 0.0) f_taf = 1.0;
      else if (tmax <= 0.0) f_taf = 0.0;
      else {
        double At, Ag;
        grounded_area(vert(m, via), vert(m, vib), vert(m, vic), a, b, c, At, Ag);
        f_taf = fmin(1.0, fmax(0.0, Ag / At));
      }
    }
    if (choice != UFE_GF_BILIN_INTERP_TAF) {
      if (m.TriBI[ti - 1] > 0) f_cdf = (a > 0.0 || b > 0.0 || c > 0.0) ? 1.0 : 0.0;
      else f_cdf = cdf_fraction(cdf_b + r, (size_t)n, nbins, SL_b - Hi_b * ri_over_rw - dHb_b);
    }
    double f = f_taf;
    if (choice == UFE_GF_BEDROCK_CDF) f = f_cdf;
    else if (choice == UFE_GF_BILIN_INTERP_TAF_PLUS_BEDROCK_CDF) {
      const bool gl = kind[via - m.a_lo] == 4 || kind[vib - m.a_lo] == 4 || kind[vic - m.a_lo] == 4;   // mask_floating_ice of a corner
      f = gl ? f_taf : f_cdf;
    }
    fraction_gr_b[off + r] = f;
  }
}

// vertices: TAF fraction of the Voronoi cell (bilin_TAF.f90:49-168), CDF fraction, combination (main.f90:76-88)
__global__ void k_geo_frac_a(GeoMesh m, int n, int off, int choice, int nbins, double ri_over_rw, const double* __restrict__ TAF,
                             const double* __restrict__ TAF_b, const double* __restrict__ Hi, const double* __restrict__ SL,
                             const double* __restrict__ dHb, const double* __restrict__ cdf, double* __restrict__ fraction_gr) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const int vi = m.vi1 + (int)r;
    const size_t i = off + r;
    double f_taf = 0.0, f_cdf = 0.0;
    if (choice != UFE_GF_BEDROCK_CDF) {
      const double TAFa = TAF[i];
      double tmax = fmax(-1e6, TAFa), tmin = fmin(1e6, TAFa);
      const int nc = m.nC[vi - 1];
      for (int ci = 0; ci < nc; ++ci) {
        const double t = TAF[m.C[(size_t)ci * m.nV + vi - 1] - m.a_lo];
        tmax = fmax(tmax, t); tmin = fmin(tmin, t);
      }
      if (tmin >= 0.0) f_taf = 1.0;
      else if (tmax <= 0.0) f_taf = 0.0;
      else {
        double A_vor = 0.0, A_grnd = 0.0, At, Ag;
        const P2 va = vert(m, vi);
        const int nt = m.niTri[vi - 1];
        auto itri = [&](int k) { return m.iTri[(size_t)k * m.nV + vi - 1]; };
        if (m.VBI[vi - 1] == 0) {
          for (int k1 = 0; k1 < nt; ++k1) {
            const int t1 = itri(k1), t2 = itri(k1 + 1 == nt ? 0 : k1 + 1);
            grounded_area(va, tricc(m, t1), tricc(m, t2), TAFa, TAF_b[t1 - m.b_lo], TAF_b[t2 - m.b_lo], At, Ag);
            A_vor = A_vor + At; A_grnd = A_grnd + Ag;
          }
        } else {
          int vj = m.C[vi - 1];
          P2 pj = vert(m, vj);
          int t = itri(0);
          grounded_area(va, P2{0.5 * (va.x + pj.x), 0.5 * (va.y + pj.y)}, tricc(m, t), TAFa, 0.5 * (TAFa + TAF[vj - m.a_lo]), TAF_b[t - m.b_lo], At, Ag);
          A_vor = A_vor + At; A_grnd = A_grnd + Ag;
          for (int k1 = 0; k1 < nt - 1; ++k1) {
            const int t1 = itri(k1), t2 = itri(k1 + 1);
            grounded_area(va, tricc(m, t1), tricc(m, t2), TAFa, TAF_b[t1 - m.b_lo], TAF_b[t2 - m.b_lo], At, Ag);
            A_vor = A_vor + At; A_grnd = A_grnd + Ag;
          }
          t = itri(nt - 1);
          vj = m.C[(size_t)(nc - 1) * m.nV + vi - 1];
          pj = vert(m, vj);
          grounded_area(va, tricc(m, t), P2{0.5 * (va.x + pj.x), 0.5 * (va.y + pj.y)}, TAFa, TAF_b[t - m.b_lo], 0.5 * (TAFa + TAF[vj - m.a_lo]), At, Ag);
          A_vor = A_vor + At; A_grnd = A_grnd + Ag;
        }
        f_taf = fmin(1.0, fmax(0.0, A_grnd / A_vor));
      }
    }
    if (choice != UFE_GF_BILIN_INTERP_TAF) f_cdf = cdf_fraction(cdf + r, (size_t)n, nbins, SL[i] - Hi[i] * ri_over_rw - dHb[i]);
    fraction_gr[i] = choice == UFE_GF_BILIN_INTERP_TAF ? f_taf : choice == UFE_GF_BEDROCK_CDF ? f_cdf : fmin(f_taf, f_cdf);
  }
}

__device__ __forceinline__ void block_min_to(double v, unsigned long long* dtbits) {
  for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
  if ((threadIdx.x & 31) == 0) atomicMin(dtbits, (unsigned long long)__double_as_longlong(v));   // v > 0: bit order = value order
}

// calc_critical_timestep_adv (time_step_criteria.f90:113-131): every edge is seen from the vertex (or two) that owns an end of it
__global__ void k_geo_dt_adv(GeoMesh m, int n, int off, int grounded_only, double dt_max, const double* __restrict__ Hi, const int* __restrict__ fl,
                             const double* __restrict__ u_b, const double* __restrict__ v_b, unsigned long long* dtbits) {
  double best = dt_max;
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const int vi = m.vi1 + (int)r;
    if (Hi[off + r] == 0.0 || (grounded_only && fl[off + r])) continue;
    const P2 pi = vert(m, vi);
    for (int ci = 0; ci < m.nC[vi - 1]; ++ci) {
      const size_t k = (size_t)ci * m.nV + vi - 1;
      const int vj = m.C[k], ei = m.VE[k];
      if (Hi[vj - m.a_lo] == 0.0 || (grounded_only && fl[vj - m.a_lo])) continue;
      const int til = m.ETri[ei - 1], tir = m.ETri[(size_t)m.nE + ei - 1];
      double uc, vc;
      if (til == 0) { uc = u_b[tir - m.b_lo]; vc = v_b[tir - m.b_lo]; }
      else if (tir == 0) { uc = u_b[til - m.b_lo]; vc = v_b[til - m.b_lo]; }
      else { uc = (u_b[til - m.b_lo] + u_b[tir - m.b_lo]) / 2.0; vc = (v_b[til - m.b_lo] + v_b[tir - m.b_lo]) / 2.0; }
      const P2 pj = vert(m, vj);
      const double dx = pi.x - pj.x, dy = pi.y - pj.y;
      best = fmin(best, sqrt(dx * dx + dy * dy) / fmax(0.1, fabs(uc) + fabs(vc)) * 0.9);
    }
  }
  block_min_to(best, dtbits);
}

// calc_critical_timestep_SIA (:46-72)
__global__ void k_geo_dt_sia(GeoMesh m, int n, int off, int nb_ext, int nz, double dt_max, const double* __restrict__ Hi, const double* __restrict__ D_3D_b,
                             unsigned long long* dtbits) {
  double best = dt_max;
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const int ti = m.ti1 + (int)r;
    const int via = m.Tri[ti - 1], vib = m.Tri[(size_t)m.nTri + ti - 1], vic = m.Tri[2 * (size_t)m.nTri + ti - 1];
    const P2 a = vert(m, via), b = vert(m, vib), c = vert(m, vic);
    auto dist = [](P2 p, P2 q) { return sqrt((q.x - p.x) * (q.x - p.x) + (q.y - p.y) * (q.y - p.y)); };
    const double d_min = fmin(dist(a, b), fmin(dist(b, c), dist(c, a)));
    double Dm = 0.0;
    for (int k = 0; k < nz; ++k) Dm = fmax(Dm, fabs(D_3D_b[(size_t)k * nb_ext + off + r]));
    const double H = fmax(fmax(0.1, Hi[via - m.a_lo]), fmax(Hi[vib - m.a_lo], Hi[vic - m.a_lo]));
    best = fmin(best, d_min * d_min / (6.0 * H * fmax(1e2, Dm)) * 0.9);
  }
  block_min_to(best, dtbits);
}

// mask bits packed by pack_mask_bits
enum { MB_LAND = 1, MB_OCEAN = 2, MB_GROUNDED = 4, MB_FLOATING = 8, MB_GL_GR = 16, MB_GL_FL = 32, MB_CF_GR = 64, MB_CF_FL = 128 };

// ---- apply_grounded_fractions_to_bed_roughness (sliding_laws.f90:592-631) -------------------------------------------------------------------
__global__ void k_bed_roughness_weight(int n, int off, const int* __restrict__ bits, const double* __restrict__ Hi, const double* __restrict__ Hs_slope,
                                       const double* __restrict__ fraction_gr, const double* __restrict__ rough, double* __restrict__ applied) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const size_t i = off + r;
    const int mb = bits[i];
    double w = 1.0;
    const double exponent_hi = log10(fmax(1.0, Hi[i]));
    const double exponent_hs = Hs_slope[i] / 0.005;
    const double exponent_gr = fmax(0.0, exponent_hi - exponent_hs);
    if (mb & (MB_GL_GR | MB_CF_GR | MB_GL_FL)) w = pow(fraction_gr[i], exponent_gr);
    else if (mb & MB_GROUNDED) w = 1.0;
    else if (mb & (MB_FLOATING | MB_OCEAN)) w = 0.0;
    w = fmin(1.0, fmax(0.0, w));
    applied[i] = rough[i] * w;
  }
}

// ---- calc_pc_truncation_error + the guilty count of run_ice_dynamics_model_pc (predictor_corrector_scheme.f90:327-361, 254-279) ------------
// counts[0] = n_tot, counts[1] = n_guilty; etabits = bits of the (positive) running maximum, initialised with C%pc_eta_min
__global__ void k_pc_truncation_error(int n, int off, double zeta_t, double dt_n, double eps, const double* __restrict__ Hi_np1,
                                      const double* __restrict__ Hi_star, const int* __restrict__ grounded, const int* __restrict__ gl_gr,
                                      const double* __restrict__ fraction_gr, double* __restrict__ tau, int* __restrict__ guilty_rec,
                                      unsigned long long* etabits, int* counts) {
  double best = 0.0;
  int ntot = 0, ngui = 0;
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const size_t i = off + r;
    const double t = zeta_t * fabs(Hi_np1[i] - Hi_star[i]) / ((3.0 * zeta_t + 3.0) * dt_n);
    tau[i] = t;
    const bool full = fraction_gr[i] == 1.0;
    if (grounded[i] && !gl_gr[i] && full) best = fmax(best, t);
    if (full) {     // fraction_gr < 1: not considered (:262)
      ++ntot;
      if (t > eps) { ++ngui; if (guilty_rec) guilty_rec[i] = guilty_rec[i] + 1; }
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    best = fmax(best, __shfl_xor_sync(0xffffffffu, best, o));
    ntot += __shfl_xor_sync(0xffffffffu, ntot, o);
    ngui += __shfl_xor_sync(0xffffffffu, ngui, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (best > 0.0) atomicMax(etabits, (unsigned long long)__double_as_longlong(best));   // positive doubles: bit order = value order
    if (ntot) atomicAdd(&counts[0], ntot);
    if (ngui) atomicAdd(&counts[1], ngui);
  }
}

// ---- alter_ice_thickness (ice_thickness_safeties.f90:26-288), after the two effective-thickness passes on Hi_new ------------------------------
struct AlterDev {
  ufe_b200_alter_params p;
  double rw_over_ri, fixiness, limitness;
  int protect;
};
__global__ void k_geo_alter(int n, int off, AlterDev a, const int* __restrict__ bits, const double* __restrict__ Hi_old, const double* __restrict__ Hb,
                            const double* __restrict__ SL, const double* __restrict__ Hi_eff, const double* __restrict__ ref_Hi,
                            const double* __restrict__ ref_Hb, const double* __restrict__ Ti_hom, double* __restrict__ Hi_new, int* err) {
  const ufe_b200_alter_params& P = a.p;
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const size_t i = off + r;
    const int mb = bits[i];
    double H = Hi_new[i];
    if (a.protect && (mb & MB_GROUNDED)) H = fmax(H, (SL[i] - Hb[i]) * a.rw_over_ri + 0.1);
    const double he = Hi_eff[i];
    const bool fl = is_floating(a.rw_over_ri, he, Hb[i], SL[i]);
    if (he < P.Hi_min) H = 0.0;
    if (P.calving_law_threshold_thickness && fl && he < P.calving_threshold_thickness_shelf) H = 0.0;
    if (P.remove_ice_absent_at_PD && ref_Hi[i] == 0.0) H = 0.0;
    if (P.do_remove_shelves && fl) H = 0.0;
    if (P.remove_shelves_larger_than_PD && ref_Hi[i] == 0.0 && ref_Hb[i] < 0.0) H = 0.0;
    if (P.continental_shelf_calving && ref_Hi[i] == 0.0 && ref_Hb[i] < P.continental_shelf_min_height) H = 0.0;
    double up = 0.0, down = 0.0;
    const bool inland = (mb & MB_GROUNDED) && !(mb & MB_GL_GR);
    switch (P.modiness_H_style) {
      case UFE_MODINESS_TI_HOM: up = down = 1.0 - exp(Ti_hom[i] / P.modiness_T_hom_ref); break;
      case UFE_MODINESS_TI_HOM_UP: up = 1.0 - exp(Ti_hom[i] / P.modiness_T_hom_ref); break;
      case UFE_MODINESS_TI_HOM_DOWN: down = 1.0 - exp(Ti_hom[i] / P.modiness_T_hom_ref); break;
      case UFE_MODINESS_NO_THICK_INLAND: if (inland) up = 1.0; break;
      case UFE_MODINESS_NO_THIN_INLAND: if (inland) down = 1.0; break;
      default: break;
    }
    up = fmin(1.0, fmax(0.0, up)); down = fmin(1.0, fmax(0.0, down));
    double fix = 0.0, lim = 0.0;
    if (mb & MB_GL_GR) { fix = P.fixiness_H_gl_gr * a.fixiness; lim = P.limitness_H_gl_gr * a.limitness; }
    else if (mb & MB_GL_FL) { fix = P.fixiness_H_gl_fl * a.fixiness; lim = P.limitness_H_gl_fl * a.limitness; }
    else if (mb & MB_GROUNDED) { fix = P.fixiness_H_grounded * a.fixiness; lim = P.limitness_H_grounded * a.limitness; }
    else if (mb & MB_FLOATING) { fix = P.fixiness_H_floating * a.fixiness; lim = P.limitness_H_floating * a.limitness; }
    else if (mb & MB_LAND) { if (P.fixiness_H_freeland && a.fixiness > 0.0) fix = 1.0; lim = P.limitness_H_grounded * a.limitness; }
    else if (mb & MB_OCEAN) { if (P.fixiness_H_freeocean && a.fixiness > 0.0) fix = 1.0; lim = P.limitness_H_floating * a.limitness; }
    else { *err = 1; }   // 'vertex neither grounded, floating, nor ice-free?' (:259)
    H = Hi_old[i] * fix + H * (1.0 - fix);
    H = fmin(H, ref_Hi[i] + (1.0 - up) * lim + (1.0 - a.limitness) * (H - ref_Hi[i]));
    H = fmax(H, ref_Hi[i] - (1.0 - down) * lim - (1.0 - a.limitness) * (ref_Hi[i] - H));
    Hi_new[i] = H;
  }
}

// the six mask bits alter_ice_thickness reads, from the basic mask code of the geometry the masks belong to (masks_mod.f90:63-203)
__global__ void k_geo_alter_bits(GeoMesh m, int n, int off, const int* __restrict__ kind, int* __restrict__ bits) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const int vi = m.vi1 + (int)r;
    const int k = kind[off + r];
    bool nb_fl = false, nb_gr = false;
    for (int ci = 0; ci < m.nC[vi - 1]; ++ci) {
      const int kj = kind[m.C[(size_t)ci * m.nV + vi - 1] - m.a_lo];
      nb_fl |= kj == 4; nb_gr |= kj == 3;
    }
    int b = (k == 1 ? MB_LAND : 0) | (k == 2 ? MB_OCEAN : 0) | (k == 3 ? MB_GROUNDED : 0) | (k == 4 ? MB_FLOATING : 0);
    if (k == 3 && nb_fl) b |= MB_GL_GR;
    if (k == 4 && nb_gr) b |= MB_GL_FL;
    bits[off + r] = b;
  }
}
// alter_ice_thickness with every optional modification switched off: what is left is the removal of very thin ice (:75-80)
__global__ void k_geo_alter_thin(int n, int off, double Hi_min, const double* __restrict__ Hi_eff, double* __restrict__ Hi_new) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x)
    if (Hi_eff[off + r] < Hi_min) Hi_new[off + r] = 0.0;
}

// ---- calc_and_apply_spill_over_flux (ice_thickness_safeties.f90:290-437) -------------------------------------------------------------------
struct SpillMesh { const double *A, *D_x, *D_y, *D; };
__device__ __forceinline__ double u_perp_of(const GeoMesh& m, const SpillMesh& g, int vi, int ci, const double* __restrict__ u_b, const double* __restrict__ v_b) {
  const size_t k = (size_t)ci * m.nV + vi - 1;
  const int ei = m.VE[k];
  const int til = m.ETri[ei - 1], tir = m.ETri[(size_t)m.nE + ei - 1];
  double uc, vc;
  if (til == 0) { uc = u_b[tir - m.b_lo]; vc = v_b[tir - m.b_lo]; }
  else if (tir == 0) { uc = u_b[til - m.b_lo]; vc = v_b[til - m.b_lo]; }
  else { uc = (u_b[til - m.b_lo] + u_b[tir - m.b_lo]) / 2.0; vc = (v_b[til - m.b_lo] + v_b[tir - m.b_lo]) / 2.0; }
  return uc * g.D_x[k] / g.D[k] + vc * g.D_y[k] / g.D[k];
}
// sources: Q_src and the sum of the outflow weights of every calving-front vertex (:323-392); the relative weights are re-formed by the receiver
__global__ void k_geo_spill_src(GeoMesh m, SpillMesh g, int n, int off, double dt, double w_eps, const int* __restrict__ bits, const double* __restrict__ Hi_new,
                                const double* __restrict__ u_b, const double* __restrict__ v_b, double* __restrict__ Q_src, double* __restrict__ sumw) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const int vi = m.vi1 + (int)r;
    const size_t i = off + r;
    double q = 0.0, sw = 0.0;
    if (bits[i] & (MB_CF_GR | MB_CF_FL)) {
      const int nc = m.nC[vi - 1];
      int cm = 0; double umin = 1.7976931348623157e308;
      for (int ci = 0; ci < nc; ++ci) {
        const double up = u_perp_of(m, g, vi, ci, u_b, v_b);
        if (up < umin) { umin = up; cm = ci; }
      }
      const double Hi_ups = Hi_new[m.C[(size_t)cm * m.nV + vi - 1] - m.a_lo];
      if (Hi_ups != 0.0 && Hi_new[i] > Hi_ups) {
        q = -(Hi_new[i] - Hi_ups) * g.A[vi - 1] / dt;
        for (int ci = 0; ci < nc; ++ci)
          if (bits[m.C[(size_t)ci * m.nV + vi - 1] - m.a_lo] & MB_OCEAN) sw += fmax(0.0, u_perp_of(m, g, vi, ci, u_b, v_b)) + w_eps;
        if (sw < w_eps) q = 0.0;
      }
    }
    Q_src[i] = q; sumw[i] = sw;
  }
}
// destinations (:398-424) and the update (:427-435)
__global__ void k_geo_spill_dst(GeoMesh m, SpillMesh g, int n, int off, double dt, double w_eps, const int* __restrict__ bits, const double* __restrict__ u_b,
                                const double* __restrict__ v_b, const double* __restrict__ Q_src, const double* __restrict__ sumw, double* __restrict__ Hi_new,
                                double* __restrict__ Qspill) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const int vi = m.vi1 + (int)r;
    const size_t i = off + r;
    double qd = 0.0;
    if (bits[i] & MB_OCEAN) {
      for (int ci = 0; ci < m.nC[vi - 1]; ++ci) {
        const int vj = m.C[(size_t)ci * m.nV + vi - 1];
        const double qs = Q_src[vj - m.a_lo];
        if (!(qs < -1.e-2)) continue;
        for (int cj = 0; cj < m.nC[vj - 1]; ++cj)
          if (m.C[(size_t)cj * m.nV + vj - 1] == vi) {
            const double rel = (fmax(0.0, u_perp_of(m, g, vj, cj, u_b, v_b)) + w_eps) / sumw[vj - m.a_lo];
            if (rel > 1.e-6) qd = qd - qs * rel;
          }
      }
    }
    const double q = (Q_src[i] + qd) / g.A[vi - 1];
    Qspill[i] = q;
    Hi_new[i] = Hi_new[i] + q * dt;
  }
}

int pack_mask_bits(Ctx* c, Geometry& G, const ufe_b200_masks* mk, int which) {   // which: MB_* bits the caller needs; missing pointers are an error
  Mesh& M = *c->mesh;
  const int na = M.a.n_loc();
  const int32_t* src[8] = {mk->mask_icefree_land, mk->mask_icefree_ocean, mk->mask_grounded_ice, mk->mask_floating_ice, mk->mask_gl_gr, mk->mask_gl_fl,
                           mk->mask_cf_gr, mk->mask_cf_fl};
  std::vector<int> bits(na, 0);
  for (int q = 0; q < 8; ++q) {
    if (!(which & (1 << q))) continue;
    UFE_CHECK(src[q], "a required mask is NULL (bit %d of land, ocean, grounded, floating, gl_gr, gl_fl, cf_gr, cf_fl)", q);
    for (int i = 0; i < na; ++i) bits[i] |= src[q][i] ? (1 << q) : 0;
  }
  UFE_CUDA(cudaMemcpyAsync(G.ia[2].p + M.a.off(), bits.data(), sizeof(int) * na, cudaMemcpyHostToDevice, c->stream));
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

double decay_weight(double time, double t0, double t1, int before_start) {   // ice_thickness_safeties.f90:118-172
  double f = 1.0;
  if (t0 >= t1) f = 0.0;
  else if (time <= t0) f = before_start ? 1.0 : 0.0;
  else if (time >= t1) f = 0.0;
  else f = 1.0 - (time - t0) / (t1 - t0);
  return f < 0.0 ? 0.0 : (f > 1.0 ? 1.0 : f);
}

int ensure_geometry(Ctx* c) {
  UFE_CHECK(c->mesh && c->mesh->have_ops, "the geometry stage needs the mesh and its operators (ufe_b200_calc_matrix_operators)");
  Mesh& M = *c->mesh;
  if (!c->geo) c->geo = new Geometry();
  Geometry& G = *c->geo;
  for (auto& x : G.a) UFE_TRY(x.alloc(M.a.n_ext()));
  for (auto& x : G.b) UFE_TRY(x.alloc(M.b.n_ext()));
  for (auto& x : G.ia) UFE_TRY(x.alloc(M.a.n_ext()));
  UFE_TRY(G.masks.alloc((size_t)11 * M.a.n_loc()));
  UFE_TRY(G.dtbits.alloc(1));
  return 0;
}

int up_a(Ctx* c, double* d, const double* h) {
  Mesh& M = *c->mesh;
  UFE_CUDA(cudaMemcpyAsync(d + M.a.off(), h, sizeof(double) * M.a.n_loc(), cudaMemcpyHostToDevice, c->stream));
  return 0;
}
int down_a(Ctx* c, double* h, const double* d) {
  Mesh& M = *c->mesh;
  if (h) UFE_CUDA(cudaMemcpyAsync(h, d + M.a.off(), sizeof(double) * M.a.n_loc(), cudaMemcpyDeviceToHost, c->stream));
  return 0;
}

int finish_dt(Ctx* c, Geometry& G, double dt_ice_max, double* out) {
  cudaStream_t st = c->stream;
  unsigned long long bits = 0;
  UFE_CUDA(cudaMemcpyAsync(&bits, G.dtbits.p, sizeof(bits), cudaMemcpyDeviceToHost, st));
  UFE_CUDA(cudaStreamSynchronize(st));
  double dt = __builtin_bit_cast(double, bits);
  if (c->nranks > 1) {   // MPI_ALLREDUCE( ..., MPI_MIN, ...) (time_step_criteria.f90:74,133)
    UFE_CUDA(cudaMemcpyAsync(G.a[7].p, &dt, sizeof(double), cudaMemcpyHostToDevice, st));
    UFE_TRY(allreduce_min(c, G.a[7].p, 1));
    UFE_CUDA(cudaMemcpyAsync(&dt, G.a[7].p, sizeof(double), cudaMemcpyDeviceToHost, st));
    UFE_CUDA(cudaStreamSynchronize(st));
  }
  *out = dt < dt_ice_max ? dt : dt_ice_max;
  return 0;
}

// velocity-solve masks from the basic mask code
__global__ void k_geo_vel_masks(int n, int off, const int* __restrict__ kind, int* __restrict__ mask_grounded_ice, int* __restrict__ mask_icefree_land,
                                int* any_grounded) {
  int any = 0;
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const int k = kind[off + r];
    mask_grounded_ice[off + r] = k == 3; mask_icefree_land[off + r] = k == 1;
    any |= k == 3;
  }
  if (__any_sync(0xffffffffu, any) && (threadIdx.x & 31) == 0) atomicOr(any_grounded, 1);
}

}  // namespace

bool geometry_dynamic(const Ctx* c) { return c->geo && c->geo->dynamic; }

int geometry_for_velocity(Ctx* c, const ufe_b200_config* cfg, double* Hi, const double* Hb, double* SL, int* mask_grounded_ice, int* mask_icefree_land,
                          double* fraction_gr_b, int* any_grounded) {
  UFE_TRY(ensure_geometry(c));
  Mesh& M = *c->mesh; Geometry& G = *c->geo;
  cudaStream_t st = c->stream;
  const int na = M.a.n_loc(), nb = M.b.n_loc(), off = M.a.off();
  const GeoMesh m = geo_mesh(M);
  const double rw_ri = cfg->seawater_density / cfg->ice_density, ri_rw = cfg->ice_density / cfg->seawater_density;
  UFE_TRY(G.flag.alloc(1));
  UFE_CUDA(cudaMemsetAsync(G.flag.p, 0, sizeof(int), st));
  { UFE_LAUNCH(c, TC_COLUMN); k_geo_kind<<<grid_for(c, na, 128, 16), 128, 0, st>>>(na, off, rw_ri, Hi, Hb, SL, G.ia[0].p, G.a[4].p); }
  { UFE_LAUNCH(c, TC_COLUMN); k_geo_vel_masks<<<grid_for(c, na, 128, 16), 128, 0, st>>>(na, off, G.ia[0].p, mask_grounded_ice, mask_icefree_land, G.flag.p); }
  UFE_TRY(halo_exchange(c, M.a, G.a[4].p)); UFE_TRY(halo_exchange_int(c, M.a, G.ia[0].p));
  const bool need_cdf = G.choice != UFE_GF_BILIN_INTERP_TAF;
  if (need_cdf) { UFE_TRY(halo_exchange(c, M.a, Hi)); UFE_TRY(halo_exchange(c, M.a, SL)); }
  const Mesh::OpGroup& gab = M.grp[Mesh::G_AB];
  {   // only the b-grid fraction feeds the velocity solve (basal friction); TAF_b is a by-product
    UFE_LAUNCH(c, TC_TRI);
    k_geo_frac_b<<<grid_for(c, nb, 128, 16), 128, 0, st>>>(m, nb, M.b.off(), G.choice, G.nbins, ri_rw, gab.ptr.p, gab.ind.p, gab.val[0].p, G.a[4].p, Hi, SL,
                                                            G.dHb_res.p, G.ia[0].p, G.cdf_b_res.p, G.b[0].p, fraction_gr_b);
  }
  UFE_CUDA(cudaGetLastError());
  UFE_CUDA(cudaMemcpyAsync(any_grounded, G.flag.p, sizeof(int), cudaMemcpyDeviceToHost, st));
  UFE_CUDA(cudaStreamSynchronize(st));
  return 0;
}

int geometry_for_thickness(Ctx* c, const ufe_b200_config* cfg, double* Hi, const double* Hb, const double* SL, double* fraction_margin) {
  UFE_TRY(ensure_geometry(c));
  Mesh& M = *c->mesh; Geometry& G = *c->geo;
  cudaStream_t st = c->stream;
  const int na = M.a.n_loc(), off = M.a.off();
  const GeoMesh m = geo_mesh(M);
  const int grid = grid_for(c, na, 128, 16);
  UFE_TRY(halo_exchange(c, M.a, Hi));
  { UFE_LAUNCH(c, TC_COLUMN); k_geo_margin<<<grid, 128, 0, st>>>(m, na, off, Hi, G.ia[1].p); }
  UFE_TRY(halo_exchange_int(c, M.a, G.ia[1].p));
  { UFE_LAUNCH(c, TC_COLUMN); k_geo_hi_eff<<<grid, 128, 0, st>>>(m, na, off, cfg->seawater_density / cfg->ice_density, Hi, Hb, SL, G.ia[1].p, G.a[5].p, fraction_margin); }
  UFE_CUDA(cudaGetLastError());
  return 0;
}

// alter_ice_thickness( mesh, ice, Hi_old, Hb, SL, Hi_new, refgeo, time) on device fields (extended a-range vectors, owned part valid on entry).
// prm == nullptr: every optional modification off, only `Hi_eff_new < C%Hi_min -> 0` remains (the scaling configuration's settings).
// Hi_of_masks: the thickness ice%mask_* belong to at the time of the call (region%ice%Hi of the last determine_masks).
int geometry_alter(Ctx* c, const ufe_b200_config* cfg, const ufe_b200_alter_params* prm, double time, double* Hi_of_masks, const double* Hi_old,
                   const double* Hb, const double* SL, const double* ref_Hi, const double* ref_Hb, const double* Ti_hom, double* Hi_new) {
  UFE_TRY(ensure_geometry(c));
  Mesh& M = *c->mesh; Geometry& G = *c->geo;
  cudaStream_t st = c->stream;
  const int na = M.a.n_loc(), off = M.a.off();
  const GeoMesh m = geo_mesh(M);
  const int grid = grid_for(c, na, 128, 16);
  const double rw_ri = cfg->seawater_density / cfg->ice_density;
  // the would-be effective thickness of Hi_new (:55)
  UFE_TRY(halo_exchange(c, M.a, Hi_new));
  { UFE_LAUNCH(c, TC_COLUMN); k_geo_margin<<<grid, 128, 0, st>>>(m, na, off, Hi_new, G.ia[1].p); }
  UFE_TRY(halo_exchange_int(c, M.a, G.ia[1].p));
  { UFE_LAUNCH(c, TC_COLUMN); k_geo_hi_eff<<<grid, 128, 0, st>>>(m, na, off, rw_ri, Hi_new, Hb, SL, G.ia[1].p, G.a[5].p, G.a[6].p); }
  if (!prm) {
    UFE_LAUNCH(c, TC_COLUMN);
    k_geo_alter_thin<<<grid, 128, 0, st>>>(na, off, cfg->Hi_min, G.a[5].p, Hi_new);
    UFE_CUDA(cudaGetLastError());
    return 0;
  }
  { UFE_LAUNCH(c, TC_COLUMN); k_geo_kind<<<grid, 128, 0, st>>>(na, off, rw_ri, Hi_of_masks, Hb, SL, G.ia[0].p, nullptr); }
  UFE_TRY(halo_exchange_int(c, M.a, G.ia[0].p));
  { UFE_LAUNCH(c, TC_COLUMN); k_geo_alter_bits<<<grid, 128, 0, st>>>(m, na, off, G.ia[0].p, G.ia[2].p); }
  AlterDev a;
  a.p = *prm; a.rw_over_ri = rw_ri;
  a.fixiness = decay_weight(time, prm->fixiness_t_start, prm->fixiness_t_end, prm->do_fixiness_before_start);
  a.limitness = decay_weight(time, prm->limitness_t_start, prm->limitness_t_end, prm->do_limitness_before_start);
  a.protect = prm->do_protect_grounded_mask && time <= prm->protect_grounded_mask_t_end;
  UFE_TRY(G.flag.alloc(1));
  UFE_CUDA(cudaMemsetAsync(G.flag.p, 0, sizeof(int), st));
  { UFE_LAUNCH(c, TC_COLUMN); k_geo_alter<<<grid, 128, 0, st>>>(na, off, a, G.ia[2].p, Hi_old, Hb, SL, G.a[5].p, ref_Hi, ref_Hb, Ti_hom, Hi_new, G.flag.p); }
  UFE_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace ufe

using namespace ufe;

extern "C" {

int ufe_b200_geometry_resident(ufe_b200_handle h, int choice, int nbins, const double* dHb, const double* bedrock_cdf, const double* bedrock_cdf_b) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c, "NULL handle");
  UFE_CUDA(cudaSetDevice(c->device));
  UFE_TRY(ensure_geometry(c));
  Mesh& M = *c->mesh; Geometry& G = *c->geo;
  if (choice < 0) { G.dynamic = false; return 0; }
  UFE_CHECK(choice <= UFE_GF_BILIN_INTERP_TAF_PLUS_BEDROCK_CDF, "unknown choice_subgrid_grounded_fraction %d", choice);
  const bool need_cdf = choice != UFE_GF_BILIN_INTERP_TAF;
  UFE_CHECK(!need_cdf || (dHb && bedrock_cdf && bedrock_cdf_b && nbins >= 2), "ufe_b200_geometry_resident: the bedrock CDFs and dHb are required");
  UFE_CHECK(choice == UFE_GF_BEDROCK_CDF || M.Tricc.p, "ufe_b200_geometry_resident: bilin_interp_TAF needs mesh%%Tricc (ufe_b200_mesh_desc.Tricc)");
  const int na = M.a.n_loc(), nb = M.b.n_loc();
  cudaStream_t st = c->stream;
  UFE_TRY(G.dHb_res.alloc(M.a.n_ext()));
  UFE_CUDA(cudaMemsetAsync(G.dHb_res.p, 0, sizeof(double) * M.a.n_ext(), st));
  if (need_cdf) {
    UFE_TRY(G.cdf_res.alloc((size_t)na * nbins)); UFE_TRY(G.cdf_b_res.alloc((size_t)nb * nbins));
    UFE_CUDA(cudaMemcpyAsync(G.dHb_res.p + M.a.off(), dHb, sizeof(double) * na, cudaMemcpyHostToDevice, st));
    UFE_CUDA(cudaMemcpyAsync(G.cdf_res.p, bedrock_cdf, sizeof(double) * na * nbins, cudaMemcpyHostToDevice, st));
    UFE_CUDA(cudaMemcpyAsync(G.cdf_b_res.p, bedrock_cdf_b, sizeof(double) * nb * nbins, cudaMemcpyHostToDevice, st));
    UFE_TRY(halo_exchange(c, M.a, G.dHb_res.p));
  }
  UFE_CUDA(cudaStreamSynchronize(st));
  G.choice = choice; G.nbins = nbins; G.dynamic = true;
  return 0;
}

int ufe_b200_determine_masks(ufe_b200_handle h, const ufe_b200_config* cfg, const double* Hi, const double* Hb, const double* SL,
                             const ufe_b200_masks* out) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && cfg && Hi && Hb && SL && out, "ufe_b200_determine_masks: NULL argument");
  UFE_CUDA(cudaSetDevice(c->device));
  UFE_TRY(ensure_geometry(c));
  Mesh& M = *c->mesh; Geometry& G = *c->geo;
  cudaStream_t st = c->stream;
  const int na = M.a.n_loc(), off = M.a.off();
  { UFE_REGION(c, TC_H2D); UFE_TRY(up_a(c, G.a[0].p, Hi)); UFE_TRY(up_a(c, G.a[1].p, Hb)); UFE_TRY(up_a(c, G.a[2].p, SL)); }
  const GeoMesh m = geo_mesh(M);
  const int grid = grid_for(c, na, 128, 16);
  { UFE_LAUNCH(c, TC_COLUMN); k_geo_kind<<<grid, 128, 0, st>>>(na, off, cfg->seawater_density / cfg->ice_density, G.a[0].p, G.a[1].p, G.a[2].p, G.ia[0].p, nullptr); }
  UFE_TRY(halo_exchange_int(c, M.a, G.ia[0].p));
  { UFE_LAUNCH(c, TC_COLUMN); k_geo_masks<<<grid, 128, 0, st>>>(m, na, off, G.ia[0].p, G.masks.p); }
  UFE_CUDA(cudaGetLastError());
  {
    UFE_REGION(c, TC_D2H);
    int32_t* const dst[11] = {out->mask, out->mask_icefree_land, out->mask_icefree_ocean, out->mask_grounded_ice, out->mask_floating_ice, out->mask_margin,
                              out->mask_gl_gr, out->mask_gl_fl, out->mask_cf_gr, out->mask_cf_fl, out->mask_coastline};
    for (int q = 0; q < 11; ++q)
      if (dst[q]) UFE_CUDA(cudaMemcpyAsync(dst[q], G.masks.p + (size_t)q * na, sizeof(int) * na, cudaMemcpyDeviceToHost, st));
  }
  UFE_CUDA(cudaStreamSynchronize(st));
  return 0;
}

int ufe_b200_calc_effective_thickness(ufe_b200_handle h, const ufe_b200_config* cfg, const double* Hi, const double* Hb, const double* SL,
                                      double* Hi_eff, double* fraction_margin) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && cfg && Hi && Hb && SL && Hi_eff && fraction_margin, "ufe_b200_calc_effective_thickness: NULL argument");
  UFE_CUDA(cudaSetDevice(c->device));
  UFE_TRY(ensure_geometry(c));
  Mesh& M = *c->mesh; Geometry& G = *c->geo;
  cudaStream_t st = c->stream;
  const int na = M.a.n_loc(), off = M.a.off();
  { UFE_REGION(c, TC_H2D); UFE_TRY(up_a(c, G.a[0].p, Hi)); UFE_TRY(up_a(c, G.a[1].p, Hb)); UFE_TRY(up_a(c, G.a[2].p, SL)); }
  UFE_TRY(halo_exchange(c, M.a, G.a[0].p));
  const GeoMesh m = geo_mesh(M);
  const int grid = grid_for(c, na, 128, 16);
  { UFE_LAUNCH(c, TC_COLUMN); k_geo_margin<<<grid, 128, 0, st>>>(m, na, off, G.a[0].p, G.ia[1].p); }
  UFE_TRY(halo_exchange_int(c, M.a, G.ia[1].p));
  {
    UFE_LAUNCH(c, TC_COLUMN);
    k_geo_hi_eff<<<grid, 128, 0, st>>>(m, na, off, cfg->seawater_density / cfg->ice_density, G.a[0].p, G.a[1].p, G.a[2].p, G.ia[1].p, G.a[5].p, G.a[6].p);
  }
  UFE_CUDA(cudaGetLastError());
  { UFE_REGION(c, TC_D2H); UFE_TRY(down_a(c, Hi_eff, G.a[5].p)); UFE_TRY(down_a(c, fraction_margin, G.a[6].p)); }
  UFE_CUDA(cudaStreamSynchronize(st));
  return 0;
}

int ufe_b200_calc_grounded_fractions(ufe_b200_handle h, const ufe_b200_config* cfg, int choice, int nbins, const double* Hi, const double* Hb,
                                     const double* SL, const double* dHb, const int32_t* mask_floating_ice, const double* bedrock_cdf,
                                     const double* bedrock_cdf_b, double* fraction_gr, double* fraction_gr_b) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && cfg && Hi && Hb && SL && fraction_gr && fraction_gr_b, "ufe_b200_calc_grounded_fractions: NULL argument");
  UFE_CHECK(choice >= UFE_GF_BILIN_INTERP_TAF && choice <= UFE_GF_BILIN_INTERP_TAF_PLUS_BEDROCK_CDF,
            "unknown choice_subgrid_grounded_fraction %d", choice);                                        // main.f90:58-60
  const bool need_cdf = choice != UFE_GF_BILIN_INTERP_TAF;
  UFE_CHECK(!need_cdf || (dHb && bedrock_cdf && bedrock_cdf_b && nbins >= 2), "ufe_b200_calc_grounded_fractions: the bedrock CDFs and dHb are required");
  UFE_CHECK(choice != UFE_GF_BILIN_INTERP_TAF_PLUS_BEDROCK_CDF || mask_floating_ice, "ufe_b200_calc_grounded_fractions: mask_floating_ice is required");
  UFE_CUDA(cudaSetDevice(c->device));
  UFE_TRY(ensure_geometry(c));
  Mesh& M = *c->mesh; Geometry& G = *c->geo;
  UFE_CHECK(choice == UFE_GF_BEDROCK_CDF || M.Tricc.p, "ufe_b200_calc_grounded_fractions: bilin_interp_TAF needs mesh%%Tricc (ufe_b200_mesh_desc.Tricc)");
  cudaStream_t st = c->stream;
  const int na = M.a.n_loc(), nb = M.b.n_loc(), off = M.a.off();
  {
    UFE_REGION(c, TC_H2D);
    UFE_TRY(up_a(c, G.a[0].p, Hi)); UFE_TRY(up_a(c, G.a[1].p, Hb)); UFE_TRY(up_a(c, G.a[2].p, SL));
    if (need_cdf) {
      UFE_TRY(up_a(c, G.a[3].p, dHb));
      UFE_TRY(G.cdf.alloc((size_t)na * nbins)); UFE_TRY(G.cdf_b.alloc((size_t)nb * nbins));
      UFE_CUDA(cudaMemcpyAsync(G.cdf.p, bedrock_cdf, sizeof(double) * na * nbins, cudaMemcpyHostToDevice, st));
      UFE_CUDA(cudaMemcpyAsync(G.cdf_b.p, bedrock_cdf_b, sizeof(double) * nb * nbins, cudaMemcpyHostToDevice, st));
    }
  }
  const GeoMesh m = geo_mesh(M);
  const double rw_ri = cfg->seawater_density / cfg->ice_density, ri_rw = cfg->ice_density / cfg->seawater_density;
  { UFE_LAUNCH(c, TC_COLUMN); k_geo_kind<<<grid_for(c, na, 128, 16), 128, 0, st>>>(na, off, rw_ri, G.a[0].p, G.a[1].p, G.a[2].p, G.ia[0].p, G.a[4].p); }
  if (mask_floating_ice && choice == UFE_GF_BILIN_INTERP_TAF_PLUS_BEDROCK_CDF) {
    // the caller's mask decides which triangles touch the grounding line (main.f90:91-104); 4 = floating ice, 0 = anything else
    std::vector<int> k(na);
    for (int i = 0; i < na; ++i) k[i] = mask_floating_ice[i] ? 4 : 0;
    UFE_CUDA(cudaMemcpyAsync(G.ia[0].p + off, k.data(), sizeof(int) * na, cudaMemcpyHostToDevice, st));
    UFE_CUDA(cudaStreamSynchronize(st));
  }
  UFE_TRY(halo_exchange(c, M.a, G.a[4].p)); UFE_TRY(halo_exchange_int(c, M.a, G.ia[0].p));
  if (need_cdf) { UFE_TRY(halo_exchange(c, M.a, G.a[0].p)); UFE_TRY(halo_exchange(c, M.a, G.a[2].p)); UFE_TRY(halo_exchange(c, M.a, G.a[3].p)); }
  const Mesh::OpGroup& gab = M.grp[Mesh::G_AB];
  {
    UFE_LAUNCH(c, TC_TRI);
    k_geo_frac_b<<<grid_for(c, nb, 128, 16), 128, 0, st>>>(m, nb, M.b.off(), choice, nbins, ri_rw, gab.ptr.p, gab.ind.p, gab.val[0].p, G.a[4].p, G.a[0].p,
                                                            G.a[2].p, G.a[3].p, G.ia[0].p, G.cdf_b.p, G.b[0].p, G.b[1].p);
  }
  UFE_TRY(halo_exchange(c, M.b, G.b[0].p));
  {
    UFE_LAUNCH(c, TC_COLUMN);
    k_geo_frac_a<<<grid_for(c, na, 128, 16), 128, 0, st>>>(m, na, off, choice, nbins, ri_rw, G.a[4].p, G.b[0].p, G.a[0].p, G.a[2].p, G.a[3].p, G.cdf.p,
                                                            G.a[5].p);
  }
  UFE_CUDA(cudaGetLastError());
  {
    UFE_REGION(c, TC_D2H);
    UFE_TRY(down_a(c, fraction_gr, G.a[5].p));
    UFE_CUDA(cudaMemcpyAsync(fraction_gr_b, G.b[1].p + M.b.off(), sizeof(double) * nb, cudaMemcpyDeviceToHost, st));
  }
  UFE_CUDA(cudaStreamSynchronize(st));
  return 0;
}

void ufe_b200_alter_params_defaults(ufe_b200_alter_params* p) {   // model_configuration_type_and_namelist.f90:424-469
  if (!p) return;
  *p = ufe_b200_alter_params{};
  p->protect_grounded_mask_t_end = -9.9e9;
  p->calving_law_threshold_thickness = 1; p->calving_threshold_thickness_shelf = 100.0;
  p->continental_shelf_min_height = -2000.0;
  p->fixiness_t_start = p->fixiness_t_end = 9.9e9; p->limitness_t_start = p->limitness_t_end = 9.9e9;
  p->modiness_H_style = UFE_MODINESS_NONE; p->modiness_T_hom_ref = 20.0;
  p->limitness_H_gl_gr = p->limitness_H_gl_fl = p->limitness_H_grounded = p->limitness_H_floating = 9.9e9;
}

int ufe_b200_alter_ice_thickness(ufe_b200_handle h, const ufe_b200_config* cfg, const ufe_b200_alter_params* prm, double time, const ufe_b200_masks* masks,
                                 const double* Hi_old, const double* Hb, const double* SL, const double* refgeo_Hi, const double* refgeo_Hb,
                                 const double* Ti_hom, double* Hi_new) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && cfg && prm && masks && Hi_old && Hb && SL && refgeo_Hi && refgeo_Hb && Hi_new, "ufe_b200_alter_ice_thickness: NULL argument");
  UFE_CHECK(prm->modiness_H_style >= UFE_MODINESS_NONE && prm->modiness_H_style <= UFE_MODINESS_NO_THIN_INLAND, "unknown modiness_H_choice %d",
            prm->modiness_H_style);                                                                   // :180-181
  const bool need_T = prm->modiness_H_style >= UFE_MODINESS_TI_HOM && prm->modiness_H_style <= UFE_MODINESS_TI_HOM_DOWN;
  UFE_CHECK(!need_T || Ti_hom, "ufe_b200_alter_ice_thickness: this modiness style needs Ti_hom");
  UFE_CUDA(cudaSetDevice(c->device));
  UFE_TRY(ensure_geometry(c));
  Mesh& M = *c->mesh; Geometry& G = *c->geo;
  cudaStream_t st = c->stream;
  const int na = M.a.n_loc(), off = M.a.off();
  {
    UFE_REGION(c, TC_H2D);
    UFE_TRY(up_a(c, G.a[0].p, Hi_new)); UFE_TRY(up_a(c, G.a[1].p, Hb)); UFE_TRY(up_a(c, G.a[2].p, SL)); UFE_TRY(up_a(c, G.a[3].p, Hi_old));
    UFE_TRY(up_a(c, G.a[4].p, refgeo_Hi)); UFE_TRY(up_a(c, G.a[7].p, refgeo_Hb));
    if (need_T) UFE_TRY(up_a(c, G.a[8].p, Ti_hom));
    UFE_TRY(pack_mask_bits(c, G, masks, MB_LAND | MB_OCEAN | MB_GROUNDED | MB_FLOATING | MB_GL_GR | MB_GL_FL));
  }
  // the would-be effective thickness of Hi_new (:55)
  UFE_TRY(halo_exchange(c, M.a, G.a[0].p));
  const GeoMesh m = geo_mesh(M);
  const int grid = grid_for(c, na, 128, 16);
  const double rw_ri = cfg->seawater_density / cfg->ice_density;
  { UFE_LAUNCH(c, TC_COLUMN); k_geo_margin<<<grid, 128, 0, st>>>(m, na, off, G.a[0].p, G.ia[1].p); }
  UFE_TRY(halo_exchange_int(c, M.a, G.ia[1].p));
  { UFE_LAUNCH(c, TC_COLUMN); k_geo_hi_eff<<<grid, 128, 0, st>>>(m, na, off, rw_ri, G.a[0].p, G.a[1].p, G.a[2].p, G.ia[1].p, G.a[5].p, G.a[6].p); }
  AlterDev a;
  a.p = *prm; a.rw_over_ri = rw_ri;
  a.fixiness = decay_weight(time, prm->fixiness_t_start, prm->fixiness_t_end, prm->do_fixiness_before_start);
  a.limitness = decay_weight(time, prm->limitness_t_start, prm->limitness_t_end, prm->do_limitness_before_start);
  a.protect = prm->do_protect_grounded_mask && time <= prm->protect_grounded_mask_t_end;
  UFE_CUDA(cudaMemsetAsync(G.ia[0].p, 0, sizeof(int), st));
  {
    UFE_LAUNCH(c, TC_COLUMN);
    k_geo_alter<<<grid, 128, 0, st>>>(na, off, a, G.ia[2].p, G.a[3].p, G.a[1].p, G.a[2].p, G.a[5].p, G.a[4].p, G.a[7].p, G.a[8].p, G.a[0].p, G.ia[0].p);
  }
  UFE_CUDA(cudaGetLastError());
  int err = 0;
  { UFE_REGION(c, TC_D2H); UFE_TRY(down_a(c, Hi_new, G.a[0].p)); UFE_CUDA(cudaMemcpyAsync(&err, G.ia[0].p, sizeof(int), cudaMemcpyDeviceToHost, st)); }
  UFE_CUDA(cudaStreamSynchronize(st));
  UFE_CHECK(!err, "vertex neither grounded, floating, nor ice-free?");
  return 0;
}

int ufe_b200_calc_and_apply_spill_over_flux(ufe_b200_handle h, const ufe_b200_masks* masks, const double* Hi_eff, const double* u_vav_b,
                                            const double* v_vav_b, double dt, double* Hi_new, double* Qspill) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && masks && Hi_eff && Hi_new && dt > 0.0, "ufe_b200_calc_and_apply_spill_over_flux: NULL argument or dt <= 0");
  UFE_CHECK((u_vav_b == nullptr) == (v_vav_b == nullptr), "ufe_b200_calc_and_apply_spill_over_flux: give both velocities or neither");
  UFE_CUDA(cudaSetDevice(c->device));
  UFE_TRY(ensure_geometry(c));
  Mesh& M = *c->mesh; Geometry& G = *c->geo;
  cudaStream_t st = c->stream;
  const int na = M.a.n_loc(), nb = M.b.n_loc(), off = M.a.off();
  const double* d_u; const double* d_v;
  {
    UFE_REGION(c, TC_H2D);
    UFE_TRY(up_a(c, G.a[0].p, Hi_new));
    UFE_TRY(pack_mask_bits(c, G, masks, MB_OCEAN | MB_CF_GR | MB_CF_FL));
    if (u_vav_b) {
      UFE_TRY(G.u.alloc(M.b.n_ext())); UFE_TRY(G.v.alloc(M.b.n_ext()));
      UFE_CUDA(cudaMemcpyAsync(G.u.p + M.b.off(), u_vav_b, sizeof(double) * nb, cudaMemcpyHostToDevice, st));
      UFE_CUDA(cudaMemcpyAsync(G.v.p + M.b.off(), v_vav_b, sizeof(double) * nb, cudaMemcpyHostToDevice, st));
      d_u = G.u.p; d_v = G.v.p;
    } else {
      UFE_CHECK(c->vel && c->vel->uploaded, "ufe_b200_calc_and_apply_spill_over_flux: no velocities given and none resident on the device");
      d_u = c->vel->u.p; d_v = c->vel->v.p;
    }
  }
  (void)Hi_eff;   // ice%Hi_eff only feeds the non-calving-front branch of Hi_ups, which never produces a source (:349-356)
  UFE_TRY(halo_exchange(c, M.a, G.a[0].p)); UFE_TRY(halo_exchange_int(c, M.a, G.ia[2].p));
  UFE_TRY(halo_exchange(c, M.b, const_cast<double*>(d_u))); UFE_TRY(halo_exchange(c, M.b, const_cast<double*>(d_v)));
  const GeoMesh m = geo_mesh(M);
  const SpillMesh g{M.A.p, M.D_x.p, M.D_y.p, M.D.p};
  const double w_eps = (double)1.e-2f;   // `real(dp) :: w_eps = 1.e-2`, a default-real literal (:309)
  const int grid = grid_for(c, na, 128, 16);
  { UFE_LAUNCH(c, TC_THICKNESS); k_geo_spill_src<<<grid, 128, 0, st>>>(m, g, na, off, dt, w_eps, G.ia[2].p, G.a[0].p, d_u, d_v, G.a[5].p, G.a[6].p); }
  UFE_TRY(halo_exchange(c, M.a, G.a[5].p)); UFE_TRY(halo_exchange(c, M.a, G.a[6].p));
  { UFE_LAUNCH(c, TC_THICKNESS); k_geo_spill_dst<<<grid, 128, 0, st>>>(m, g, na, off, dt, w_eps, G.ia[2].p, d_u, d_v, G.a[5].p, G.a[6].p, G.a[0].p, G.a[7].p); }
  UFE_CUDA(cudaGetLastError());
  { UFE_REGION(c, TC_D2H); UFE_TRY(down_a(c, Hi_new, G.a[0].p)); UFE_TRY(down_a(c, Qspill, G.a[7].p)); }
  UFE_CUDA(cudaStreamSynchronize(st));
  return 0;
}

int ufe_b200_calc_critical_timestep_adv(ufe_b200_handle h, const ufe_b200_config* cfg, int do_grounded_only_adv_dt, const double* Hi,
                                        const int32_t* mask_floating_ice, const double* u_vav_b, const double* v_vav_b, double* dt_crit_adv) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && cfg && Hi && dt_crit_adv, "ufe_b200_calc_critical_timestep_adv: NULL argument");
  UFE_CHECK(!do_grounded_only_adv_dt || mask_floating_ice, "ufe_b200_calc_critical_timestep_adv: mask_floating_ice is required");
  UFE_CHECK((u_vav_b == nullptr) == (v_vav_b == nullptr), "ufe_b200_calc_critical_timestep_adv: give both velocities or neither");
  UFE_CUDA(cudaSetDevice(c->device));
  UFE_TRY(ensure_geometry(c));
  Mesh& M = *c->mesh; Geometry& G = *c->geo;
  cudaStream_t st = c->stream;
  const int na = M.a.n_loc(), nb = M.b.n_loc();
  const double* d_u; const double* d_v;
  {
    UFE_REGION(c, TC_H2D);
    UFE_TRY(up_a(c, G.a[0].p, Hi));
    if (mask_floating_ice) UFE_CUDA(cudaMemcpyAsync(G.ia[2].p + M.a.off(), mask_floating_ice, sizeof(int) * na, cudaMemcpyHostToDevice, st));
    else UFE_CUDA(cudaMemsetAsync(G.ia[2].p, 0, sizeof(int) * M.a.n_ext(), st));
    if (u_vav_b) {
      UFE_TRY(G.u.alloc(M.b.n_ext())); UFE_TRY(G.v.alloc(M.b.n_ext()));
      UFE_CUDA(cudaMemcpyAsync(G.u.p + M.b.off(), u_vav_b, sizeof(double) * nb, cudaMemcpyHostToDevice, st));
      UFE_CUDA(cudaMemcpyAsync(G.v.p + M.b.off(), v_vav_b, sizeof(double) * nb, cudaMemcpyHostToDevice, st));
      d_u = G.u.p; d_v = G.v.p;
    } else {
      UFE_CHECK(c->vel && c->vel->uploaded, "ufe_b200_calc_critical_timestep_adv: no velocities given and none resident on the device");
      d_u = c->vel->u.p; d_v = c->vel->v.p;
    }
  }
  UFE_TRY(halo_exchange(c, M.a, G.a[0].p)); UFE_TRY(halo_exchange_int(c, M.a, G.ia[2].p));
  UFE_TRY(halo_exchange(c, M.b, const_cast<double*>(d_u))); UFE_TRY(halo_exchange(c, M.b, const_cast<double*>(d_v)));
  const unsigned long long init = (unsigned long long)__builtin_bit_cast(long long, cfg->dt_ice_max);
  UFE_CUDA(cudaMemcpyAsync(G.dtbits.p, &init, sizeof(init), cudaMemcpyHostToDevice, st));
  {
    UFE_LAUNCH(c, TC_COLUMN);
    k_geo_dt_adv<<<grid_for(c, na, 128, 16), 128, 0, st>>>(geo_mesh(M), na, M.a.off(), do_grounded_only_adv_dt, cfg->dt_ice_max, G.a[0].p, G.ia[2].p, d_u, d_v,
                                                            G.dtbits.p);
  }
  UFE_CUDA(cudaGetLastError());
  return finish_dt(c, G, cfg->dt_ice_max, dt_crit_adv);
}

int ufe_b200_apply_grounded_fractions_to_bed_roughness(ufe_b200_handle h, const double* Hi, const double* Hs_slope, const double* fraction_gr,
                                                       const ufe_b200_masks* masks, const double* bed_roughness, double* bed_roughness_applied) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && Hi && Hs_slope && fraction_gr && masks && bed_roughness && bed_roughness_applied,
            "ufe_b200_apply_grounded_fractions_to_bed_roughness: NULL argument");
  UFE_CUDA(cudaSetDevice(c->device));
  UFE_TRY(ensure_geometry(c));
  Mesh& M = *c->mesh; Geometry& G = *c->geo;
  {
    UFE_REGION(c, TC_H2D);
    UFE_TRY(up_a(c, G.a[0].p, Hi)); UFE_TRY(up_a(c, G.a[1].p, Hs_slope)); UFE_TRY(up_a(c, G.a[2].p, fraction_gr)); UFE_TRY(up_a(c, G.a[3].p, bed_roughness));
    UFE_TRY(pack_mask_bits(c, G, masks, MB_OCEAN | MB_GROUNDED | MB_FLOATING | MB_GL_GR | MB_GL_FL | MB_CF_GR));
  }
  {
    UFE_LAUNCH(c, TC_COLUMN);
    k_bed_roughness_weight<<<grid_for(c, M.a.n_loc(), 256), 256, 0, c->stream>>>(M.a.n_loc(), M.a.off(), G.ia[2].p, G.a[0].p, G.a[1].p, G.a[2].p, G.a[3].p, G.a[5].p);
  }
  UFE_CUDA(cudaGetLastError());
  { UFE_REGION(c, TC_D2H); UFE_TRY(down_a(c, bed_roughness_applied, G.a[5].p)); }
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

int ufe_b200_calc_pc_truncation_error(ufe_b200_handle h, double zeta_t, double dt_n, double pc_eta_min, double pc_epsilon, const double* Hi_np1,
                                      const double* Hi_star_np1, const int32_t* mask_grounded_ice, const int32_t* mask_gl_gr,
                                      const double* fraction_gr, double* tau_np1, int32_t* tau_n_guilty, double* eta_np1, int32_t* n_tot,
                                      int32_t* n_guilty) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && Hi_np1 && Hi_star_np1 && mask_grounded_ice && mask_gl_gr && fraction_gr && eta_np1 && n_tot && n_guilty,
            "ufe_b200_calc_pc_truncation_error: NULL argument");
  UFE_CHECK(dt_n > 0.0 && pc_eta_min > 0.0, "ufe_b200_calc_pc_truncation_error: dt_n and pc_eta_min must be positive");
  UFE_CUDA(cudaSetDevice(c->device));
  UFE_TRY(ensure_geometry(c));
  Mesh& M = *c->mesh; Geometry& G = *c->geo;
  cudaStream_t st = c->stream;
  const int na = M.a.n_loc();
  auto up_i = [&](int* d, const int32_t* s) { return cudaMemcpyAsync(d + M.a.off(), s, sizeof(int) * na, cudaMemcpyHostToDevice, st); };
  UFE_TRY(G.flag.alloc(2));
  {
    UFE_REGION(c, TC_H2D);
    UFE_TRY(up_a(c, G.a[0].p, Hi_np1)); UFE_TRY(up_a(c, G.a[1].p, Hi_star_np1)); UFE_TRY(up_a(c, G.a[2].p, fraction_gr));
    UFE_CUDA(up_i(G.ia[0].p, mask_grounded_ice)); UFE_CUDA(up_i(G.ia[1].p, mask_gl_gr));
    if (tau_n_guilty) UFE_CUDA(up_i(G.ia[2].p, tau_n_guilty));
    const unsigned long long init = (unsigned long long)__builtin_bit_cast(long long, pc_eta_min);
    UFE_CUDA(cudaMemcpyAsync(G.dtbits.p, &init, sizeof(init), cudaMemcpyHostToDevice, st));
    UFE_CUDA(cudaMemsetAsync(G.flag.p, 0, 2 * sizeof(int), st));
  }
  {
    UFE_LAUNCH(c, TC_COLUMN);
    k_pc_truncation_error<<<grid_for(c, na, 256), 256, 0, st>>>(na, M.a.off(), zeta_t, dt_n, pc_epsilon, G.a[0].p, G.a[1].p, G.ia[0].p, G.ia[1].p, G.a[2].p,
                                                               G.a[5].p, tau_n_guilty ? G.ia[2].p : nullptr, G.dtbits.p, G.flag.p);
  }
  UFE_CUDA(cudaGetLastError());
  unsigned long long bits = 0; int counts[2] = {0, 0};
  {
    UFE_REGION(c, TC_D2H);
    UFE_TRY(down_a(c, tau_np1, G.a[5].p));
    if (tau_n_guilty) UFE_CUDA(cudaMemcpyAsync(tau_n_guilty, G.ia[2].p + M.a.off(), sizeof(int) * na, cudaMemcpyDeviceToHost, st));
    UFE_CUDA(cudaMemcpyAsync(&bits, G.dtbits.p, sizeof(bits), cudaMemcpyDeviceToHost, st));
    UFE_CUDA(cudaMemcpyAsync(counts, G.flag.p, sizeof(counts), cudaMemcpyDeviceToHost, st));
  }
  UFE_CUDA(cudaStreamSynchronize(st));
  double red[3] = {-__builtin_bit_cast(double, bits), (double)counts[0], (double)counts[1]};
  if (c->nranks > 1) {   // MPI_ALLREDUCE MAX of eta (:353), SUM of the two counts (:276-277); the counts are exact in doubles
    UFE_CUDA(cudaMemcpyAsync(G.a[7].p, red, sizeof(red), cudaMemcpyHostToDevice, st));
    UFE_TRY(allreduce_min(c, G.a[7].p, 1));
    UFE_TRY(allreduce_sum(c, G.a[7].p + 1, 2));
    UFE_CUDA(cudaMemcpyAsync(red, G.a[7].p, sizeof(red), cudaMemcpyDeviceToHost, st));
    UFE_CUDA(cudaStreamSynchronize(st));
  }
  *eta_np1 = -red[0]; *n_tot = (int32_t)red[1]; *n_guilty = (int32_t)red[2];
  return 0;
}

int ufe_b200_calc_critical_timestep_SIA(ufe_b200_handle h, const ufe_b200_config* cfg, const double* Hi, const double* D_3D_b, double* dt_crit_SIA) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && cfg && Hi && D_3D_b && dt_crit_SIA, "ufe_b200_calc_critical_timestep_SIA: NULL argument");
  UFE_CUDA(cudaSetDevice(c->device));
  UFE_TRY(ensure_geometry(c));
  Mesh& M = *c->mesh; Geometry& G = *c->geo;
  cudaStream_t st = c->stream;
  const int nb = M.b.n_loc(), nz = M.nz;
  const size_t nb_ext = M.b.n_ext();
  {
    UFE_REGION(c, TC_H2D);
    UFE_TRY(up_a(c, G.a[0].p, Hi));
    UFE_TRY(G.D3.alloc(nb_ext * nz));
    UFE_CUDA(cudaMemcpy2DAsync(G.D3.p + M.b.off(), sizeof(double) * nb_ext, D_3D_b, sizeof(double) * nb, sizeof(double) * nb, nz, cudaMemcpyHostToDevice, st));
  }
  UFE_TRY(halo_exchange(c, M.a, G.a[0].p));
  const unsigned long long init = (unsigned long long)__builtin_bit_cast(long long, cfg->dt_ice_max);
  UFE_CUDA(cudaMemcpyAsync(G.dtbits.p, &init, sizeof(init), cudaMemcpyHostToDevice, st));
  {
    UFE_LAUNCH(c, TC_TRI);
    k_geo_dt_sia<<<grid_for(c, nb, 128, 16), 128, 0, st>>>(geo_mesh(M), nb, M.b.off(), (int)nb_ext, nz, cfg->dt_ice_max, G.a[0].p, G.D3.p, G.dtbits.p);
  }
  UFE_CUDA(cudaGetLastError());
  return finish_dt(c, G, cfg->dt_ice_max, dt_crit_SIA);
}

}  // extern "C"
