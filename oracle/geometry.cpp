// TEST INFRASTRUCTURE ONLY -- CPU restatement of the geometry / mask / sub-grid stage between the velocity and thickness solves.
// Single rank, global arrays (the reference's gather_to_all copies are the arrays themselves here).
//   determine_masks                           src/UFEMISM/ice_dynamics/utilities/masks_mod.f90:25-217
//   calc_effective_thickness                  src/UFEMISM/ice_dynamics/utilities/subgrid_ice_margin.f90:19-158
//   calc_grounded_fractions                   src/UFEMISM/ice_dynamics/utilities/subgrid_grounded_fractions_main.f90:22-120
//   ..._bilin_interp_TAF_{a,b}, calc_grounded_area_triangle*   .../subgrid_grounded_fractions_bilin_TAF.f90:20-369
//   ..._bedrock_CDF_{a,b}                     .../subgrid_grounded_fractions_bedrock_CDF.f90:22-181
//   calc_critical_timestep_{SIA,adv}          src/UFEMISM/ice_dynamics/utilities/time_step_criteria.f90:20-141
//   is_floating, thickness_above_floatation   src/UPSY/basic/math_utilities/ice_geometry_basics.f90:20-64
//   triangle_area                             src/UPSY/basic/math_utilities/plane_geometry.f90:322-330
#include <algorithm>
#include <cmath>
#include <limits>
#include <stdexcept>

#include "oracle.hpp"

namespace orc {

namespace {

inline bool is_floating(const Config& C, double Hi, double Hb, double SL) { return Hi < (SL - Hb) * C.seawater_density / C.ice_density; }
inline double thickness_above_floatation(const Config& C, double Hi, double Hb, double SL) {
  return Hi - std::max(0.0, (SL - Hb) * (C.seawater_density / C.ice_density));
}
struct P2 { double x, y; };
inline double triangle_area(P2 p, P2 q, P2 r) { return std::fabs((q.x - p.x) * (r.y - p.y) - (q.y - p.y) * (r.x - p.x)) / 2.0; }

// calc_grounded_area_triangle_1grnd_2flt / _1flt_2grnd (bilin_TAF.f90:323-369): the same formula for the corner with the odd sign
inline double corner_area(P2 va, P2 vb, P2 vc, double TAFa, double TAFb, double TAFc) {
  const double lab = TAFa / (TAFa - TAFb), lac = TAFa / (TAFa - TAFc);
  const P2 pab{(va.x * (1.0 - lab)) + (vb.x * lab), (va.y * (1.0 - lab)) + (vb.y * lab)};
  const P2 pac{(va.x * (1.0 - lac)) + (vc.x * lac), (va.y * (1.0 - lac)) + (vc.y * lac)};
  return triangle_area(va, pab, pac);
}

// calc_grounded_area_triangle (bilin_TAF.f90:244-321)
void grounded_area_triangle(P2 va, P2 vb, P2 vc, double TAFa, double TAFb, double TAFc, double& A_tot, double& A_grnd) {
  const double tol = 1e-9;
  A_tot = triangle_area(va, vb, vc);
  const double a = TAFa >= 0.0 ? std::max(tol, TAFa) : std::min(-tol, TAFa);
  const double b = TAFb >= 0.0 ? std::max(tol, TAFb) : std::min(-tol, TAFb);
  const double c = TAFc >= 0.0 ? std::max(tol, TAFc) : std::min(-tol, TAFc);
  if (a >= 0 && b >= 0 && c >= 0) A_grnd = A_tot;
  else if (a <= 0 && b <= 0 && c <= 0) A_grnd = 0.0;
  else if (a >= 0 && b <= 0 && c <= 0) A_grnd = corner_area(va, vb, vc, a, b, c);
  else if (a <= 0 && b >= 0 && c <= 0) A_grnd = corner_area(vb, vc, va, b, c, a);
  else if (a <= 0 && b <= 0 && c >= 0) A_grnd = corner_area(vc, va, vb, c, a, b);
  else if (a <= 0 && b >= 0 && c >= 0) A_grnd = A_tot - corner_area(va, vb, vc, a, b, c);
  else if (a >= 0 && b <= 0 && c >= 0) A_grnd = A_tot - corner_area(vb, vc, va, b, c, a);
  else A_grnd = A_tot - corner_area(vc, va, vb, c, a, b);
}

inline P2 vert(const MeshView& m, int vi) { return P2{m.V[vi], m.V[vi + m.nV]}; }
inline P2 tricc(const MeshView& m, int ti) { return P2{m.Tricc[ti], m.Tricc[ti + m.nTri]}; }

// the linear interpolation in the bedrock CDF shared by the a- and b-grid versions (bedrock_CDF.f90:49-81, 147-173)
double cdf_fraction(const double* cdf, size_t stride, int nbins, double Hb_float) {
  if (Hb_float <= cdf[0]) return 1.0;
  if (Hb_float >= cdf[(size_t)(nbins - 1) * stride]) return 0.0;
  int iu = 1;                                     // 1-based as in the reference
  while (cdf[(size_t)(iu - 1) * stride] < Hb_float) iu = iu + 1;
  const int il = iu - 1;
  const double cu = cdf[(size_t)(iu - 1) * stride], cl = cdf[(size_t)(il - 1) * stride];
  const double wl = (cu - Hb_float) / (cu - cl), wu = 1.0 - wl;
  const double f = 1.0 - ((double)(il - 1) * wl + (double)(float)(iu - 1) * wu) / (double)(nbins - 1);
  return std::min(1.0, std::max(0.0, f));
}

}  // namespace

void determine_masks(const MeshView& mesh, const Config& C, const double* Hi, const double* Hb, const double* SL, const MaskSet& o) {
  const int nV = mesh.nV;
  for (int vi = 0; vi < nV; ++vi) {   // basic masks :63-103
    o.icefree_land[vi] = o.icefree_ocean[vi] = o.grounded_ice[vi] = o.floating_ice[vi] = 0;
    if (is_floating(C, Hi[vi], Hb[vi], SL[vi])) {
      if (Hi[vi] > 0.0) { o.floating_ice[vi] = 1; o.mask[vi] = 4; } else { o.icefree_ocean[vi] = 1; o.mask[vi] = 2; }
    } else {
      if (Hi[vi] > 0.0) { o.grounded_ice[vi] = 1; o.mask[vi] = 3; } else { o.icefree_land[vi] = 1; o.mask[vi] = 1; }
    }
  }
  for (int vi = 0; vi < nV; ++vi) {   // transitional masks :121-203; the order of the tests decides the final value of `mask`
    o.margin[vi] = o.gl_gr[vi] = o.gl_fl[vi] = o.cf_gr[vi] = o.cf_fl[vi] = o.coastline[vi] = 0;
    const bool ice_i = o.grounded_ice[vi] || o.floating_ice[vi];
    for (int pass = 0; pass < 6; ++pass)
      for (int ci = 0; ci < mesh.nC[vi]; ++ci) {
        const int vj = mesh.C[vi + (size_t)ci * nV] - 1;
        switch (pass) {
          case 0: if (ice_i && !(o.grounded_ice[vj] || o.floating_ice[vj])) { o.margin[vi] = 1; o.mask[vi] = 9; } break;
          case 1: if (o.grounded_ice[vi] && o.floating_ice[vj]) { o.gl_gr[vi] = 1; o.mask[vi] = 5; } break;
          case 2: if (o.floating_ice[vi] && o.grounded_ice[vj]) { o.gl_fl[vi] = 1; o.mask[vi] = 6; } break;
          case 3: if (o.grounded_ice[vi] && o.icefree_ocean[vj]) { o.cf_gr[vi] = 1; o.mask[vi] = 7; } break;
          case 4: if (o.floating_ice[vi] && o.icefree_ocean[vj]) { o.cf_fl[vi] = 1; o.mask[vi] = 8; } break;
          case 5: if (o.icefree_land[vi] && o.icefree_ocean[vj]) { o.coastline[vi] = 1; o.mask[vi] = 10; } break;
        }
      }
  }
}

void calc_effective_thickness(const MeshView& mesh, const Config& C, const double* Hi, const double* Hb, const double* SL, double* Hi_eff,
                              double* fraction_margin) {
  const int nV = mesh.nV;
  std::vector<char> margin(nV, 0), floating(nV, 0);
  for (int vi = 0; vi < nV; ++vi)
    for (int ci = 0; ci < mesh.nC[vi]; ++ci) {
      const int vc = mesh.C[vi + (size_t)ci * nV] - 1;
      if (Hi[vi] > 0.0 && Hi[vc] == 0.0) margin[vi] = 1;
    }
  for (int vi = 0; vi < nV; ++vi) floating[vi] = is_floating(C, Hi[vi], Hb[vi], SL[vi]);
  for (int vi = 0; vi < nV; ++vi) {   // :87-104
    if (!floating[vi]) { fraction_margin[vi] = 1.0; Hi_eff[vi] = Hi[vi]; }
    else if (Hi[vi] > 0.0) { fraction_margin[vi] = 1.0; Hi_eff[vi] = Hi[vi]; }
    else { fraction_margin[vi] = 0.0; Hi_eff[vi] = 0.0; }
  }
  for (int vi = 0; vi < nV; ++vi) {   // :109-146
    if (!margin[vi]) continue;
    double Hmax = 0.0;
    for (int ci = 0; ci < mesh.nC[vi]; ++ci) {
      const int vc = mesh.C[vi + (size_t)ci * nV] - 1;
      if (margin[vc]) continue;
      if (floating[vi]) Hmax = std::max(Hmax, Hi[vc]);
    }
    if (Hmax > Hi[vi]) { Hi_eff[vi] = Hmax; fraction_margin[vi] = Hi[vi] / Hmax; }
  }
}

static void fractions_TAF_a(const MeshView& mesh, const Operators& ops, const double* TAF, double* fr) {
  const int nV = mesh.nV;
  if (!mesh.Tricc) throw std::runtime_error("bilin_interp_TAF needs mesh%Tricc");
  std::vector<double> TAF_b(mesh.nTri);
  spmv(ops.M[OP_map_a_b], TAF, TAF_b.data());
  for (int vi = 0; vi < nV; ++vi) {
    double tmax = -1e6, tmin = 1e6;
    tmax = std::max(tmax, TAF[vi]); tmin = std::min(tmin, TAF[vi]);
    for (int ci = 0; ci < mesh.nC[vi]; ++ci) {
      const int vj = mesh.C[vi + (size_t)ci * nV] - 1;
      tmax = std::max(tmax, TAF[vj]); tmin = std::min(tmin, TAF[vj]);
    }
    if (tmin >= 0.0) { fr[vi] = 1.0; continue; }
    if (tmax <= 0.0) { fr[vi] = 0.0; continue; }
    double A_vor = 0.0, A_grnd = 0.0, At, Ag;
    const P2 va = vert(mesh, vi);
    const double TAFa = TAF[vi];
    const int nt = mesh.niTri[vi];
    auto itri = [&](int k) { return mesh.iTri[vi + (size_t)k * nV] - 1; };
    if (mesh.VBI[vi] == 0) {
      for (int k1 = 0; k1 < nt; ++k1) {
        const int k2 = (k1 + 1 == nt) ? 0 : k1 + 1;
        grounded_area_triangle(va, tricc(mesh, itri(k1)), tricc(mesh, itri(k2)), TAFa, TAF_b[itri(k1)], TAF_b[itri(k2)], At, Ag);
        A_vor = A_vor + At; A_grnd = A_grnd + Ag;
      }
    } else {
      int vj = mesh.C[vi] - 1;
      P2 vb{0.5 * (va.x + mesh.V[vj]), 0.5 * (va.y + mesh.V[vj + nV])};
      grounded_area_triangle(va, vb, tricc(mesh, itri(0)), TAFa, 0.5 * (TAF[vi] + TAF[vj]), TAF_b[itri(0)], At, Ag);
      A_vor = A_vor + At; A_grnd = A_grnd + Ag;
      for (int k1 = 0; k1 < nt - 1; ++k1) {
        grounded_area_triangle(va, tricc(mesh, itri(k1)), tricc(mesh, itri(k1 + 1)), TAFa, TAF_b[itri(k1)], TAF_b[itri(k1 + 1)], At, Ag);
        A_vor = A_vor + At; A_grnd = A_grnd + Ag;
      }
      vj = mesh.C[vi + (size_t)(mesh.nC[vi] - 1) * nV] - 1;
      P2 vc{0.5 * (va.x + mesh.V[vj]), 0.5 * (va.y + mesh.V[vj + nV])};
      grounded_area_triangle(va, tricc(mesh, itri(nt - 1)), vc, TAFa, TAF_b[itri(nt - 1)], 0.5 * (TAF[vi] + TAF[vj]), At, Ag);
      A_vor = A_vor + At; A_grnd = A_grnd + Ag;
    }
    fr[vi] = std::min(1.0, std::max(0.0, A_grnd / A_vor));
  }
}

static void fractions_TAF_b(const MeshView& mesh, const double* TAF, double* fr_b) {
  for (int ti = 0; ti < mesh.nTri; ++ti) {
    const int via = mesh.Tri[ti] - 1, vib = mesh.Tri[ti + mesh.nTri] - 1, vic = mesh.Tri[ti + 2 * (size_t)mesh.nTri] - 1;
    const double a = TAF[via], b = TAF[vib], c = TAF[vic];
    const double tmax = std::max(a, std::max(b, c)), tmin = std::min(a, std::min(b, c));
    if (tmin >= 0.0) { fr_b[ti] = 1.0; continue; }
    if (tmax <= 0.0) { fr_b[ti] = 0.0; continue; }
    double At, Ag;
    grounded_area_triangle(vert(mesh, via), vert(mesh, vib), vert(mesh, vic), a, b, c, At, Ag);
    fr_b[ti] = std::min(1.0, std::max(0.0, Ag / At));
  }
}

void calc_grounded_fractions(const MeshView& mesh, const Operators& ops, const Config& C, int choice, int nbins, const double* Hi, const double* Hb,
                             const double* SL, const double* dHb, const int* mask_floating_ice, const double* bedrock_cdf,
                             const double* bedrock_cdf_b, double* fraction_gr, double* fraction_gr_b) {
  const int nV = mesh.nV, nTri = mesh.nTri;
  std::vector<double> TAF(nV);
  for (int vi = 0; vi < nV; ++vi) TAF[vi] = thickness_above_floatation(C, Hi[vi], Hb[vi], SL[vi]);
  std::vector<double> ta, tb, ca, cb;
  if (choice == GF_bilin_interp_TAF || choice == GF_bilin_interp_TAF_plus_bedrock_CDF) {
    ta.resize(nV); tb.resize(nTri);
    fractions_TAF_a(mesh, ops, TAF.data(), ta.data());
    fractions_TAF_b(mesh, TAF.data(), tb.data());
  }
  if (choice == GF_bedrock_CDF || choice == GF_bilin_interp_TAF_plus_bedrock_CDF) {
    ca.resize(nV); cb.resize(nTri);
    for (int vi = 0; vi < nV; ++vi)
      ca[vi] = cdf_fraction(bedrock_cdf + vi, nV, nbins, SL[vi] - Hi[vi] * C.ice_density / C.seawater_density - dHb[vi]);
    std::vector<double> Hi_b(nTri), SL_b(nTri), dHb_b(nTri);
    spmv(ops.M[OP_map_a_b], Hi, Hi_b.data()); spmv(ops.M[OP_map_a_b], SL, SL_b.data()); spmv(ops.M[OP_map_a_b], dHb, dHb_b.data());
    for (int ti = 0; ti < nTri; ++ti) {
      if (mesh.TriBI[ti] > 0) {
        const int via = mesh.Tri[ti] - 1, vib = mesh.Tri[ti + nTri] - 1, vic = mesh.Tri[ti + 2 * (size_t)nTri] - 1;
        cb[ti] = (TAF[via] > 0.0 || TAF[vib] > 0.0 || TAF[vic] > 0.0) ? 1.0 : 0.0;
        continue;
      }
      cb[ti] = cdf_fraction(bedrock_cdf_b + ti, nTri, nbins, SL_b[ti] - Hi_b[ti] * C.ice_density / C.seawater_density - dHb_b[ti]);
    }
  }
  if (choice == GF_bilin_interp_TAF) {
    std::copy(ta.begin(), ta.end(), fraction_gr); std::copy(tb.begin(), tb.end(), fraction_gr_b);
  } else if (choice == GF_bedrock_CDF) {
    std::copy(ca.begin(), ca.end(), fraction_gr); std::copy(cb.begin(), cb.end(), fraction_gr_b);
  } else if (choice == GF_bilin_interp_TAF_plus_bedrock_CDF) {   // main.f90:76-107
    for (int vi = 0; vi < nV; ++vi) fraction_gr[vi] = std::min(ta[vi], ca[vi]);
    for (int ti = 0; ti < nTri; ++ti) {
      const int via = mesh.Tri[ti] - 1, vib = mesh.Tri[ti + nTri] - 1, vic = mesh.Tri[ti + 2 * (size_t)nTri] - 1;
      fraction_gr_b[ti] = (mask_floating_ice[via] || mask_floating_ice[vib] || mask_floating_ice[vic]) ? tb[ti] : cb[ti];
    }
  } else {
    throw std::runtime_error("unknown choice_subgrid_grounded_fraction");
  }
}

double calc_critical_timestep_SIA(const MeshView& mesh, const Config& C, const double* Hi, const double* D_3D_b) {
  double dt_crit = C.dt_ice_max;
  const int nV = mesh.nV, nTri = mesh.nTri;
  for (int ti = 0; ti < nTri; ++ti) {
    const int via = mesh.Tri[ti] - 1, vib = mesh.Tri[ti + nTri] - 1, vic = mesh.Tri[ti + 2 * (size_t)nTri] - 1;
    auto dist = [&](int p, int q) { return std::sqrt((mesh.V[q] - mesh.V[p]) * (mesh.V[q] - mesh.V[p]) + (mesh.V[q + nV] - mesh.V[p + nV]) * (mesh.V[q + nV] - mesh.V[p + nV])); };
    const double d_min = std::min(dist(via, vib), std::min(dist(vib, vic), dist(vic, via)));
    double Dm = 0.0;
    for (int k = 0; k < mesh.nz; ++k) Dm = std::max(Dm, std::fabs(D_3D_b[ti + (size_t)k * nTri]));
    const double D_SIA = std::max(1e2, Dm);
    const double H = std::max(std::max(0.1, Hi[via]), std::max(Hi[vib], Hi[vic]));
    const double dt = d_min * d_min / (6.0 * H * D_SIA) * 0.9;
    dt_crit = std::min(dt_crit, dt);
  }
  return std::min(C.dt_ice_max, dt_crit);
}

double calc_critical_timestep_adv(const MeshView& mesh, const Config& C, int do_grounded_only_adv_dt, const double* Hi, const int* mask_floating_ice,
                                  const double* u_vav_b, const double* v_vav_b) {
  double dt_crit = C.dt_ice_max;
  const int nV = mesh.nV, nE = mesh.nE;
  // every edge ei = VE( vi,ci) joins vi and C( vi,ci) (= EV( ei,1:2)); visiting it from both ends gives the same minimum
  for (int vi = 0; vi < nV; ++vi)
    for (int ci = 0; ci < mesh.nC[vi]; ++ci) {
      const int vj = mesh.C[vi + (size_t)ci * nV] - 1, ei = mesh.VE[vi + (size_t)ci * nV] - 1;
      if (Hi[vi] == 0.0 || Hi[vj] == 0.0) continue;
      if (do_grounded_only_adv_dt && (mask_floating_ice[vi] || mask_floating_ice[vj])) continue;
      const int til = mesh.ETri[ei], tir = mesh.ETri[ei + nE];
      double uc, vc;   // map_velocities_from_b_to_c_2D (map_velocities_to_c_grid.f90:51-70)
      if (til == 0) { uc = u_vav_b[tir - 1]; vc = v_vav_b[tir - 1]; }
      else if (tir == 0) { uc = u_vav_b[til - 1]; vc = v_vav_b[til - 1]; }
      else { uc = (u_vav_b[til - 1] + u_vav_b[tir - 1]) / 2.0; vc = (v_vav_b[til - 1] + v_vav_b[tir - 1]) / 2.0; }
      const double dx = mesh.V[vi] - mesh.V[vj], dy = mesh.V[vi + nV] - mesh.V[vj + nV];
      const double dist = std::sqrt(dx * dx + dy * dy);
      const double dt = dist / std::max(0.1, std::fabs(uc) + std::fabs(vc)) * 0.9;
      dt_crit = std::min(dt_crit, dt);
    }
  return std::min(C.dt_ice_max, dt_crit);
}


// decaying weight shared by fixiness and limitness (ice_thickness_safeties.f90:118-172)
static double decay_weight(double time, double t_start, double t_end, int before_start) {
  double f = 1.0;
  if (t_start >= t_end) f = 0.0;
  else if (time <= t_start) f = before_start ? 1.0 : 0.0;
  else if (time >= t_end) f = 0.0;
  else f = 1.0 - (time - t_start) / (t_end - t_start);
  return std::min(1.0, std::max(0.0, f));
}

void alter_ice_thickness(const MeshView& mesh, const Config& C, const AlterParams& P, double time, const MaskSet& M, const double* Hi_old,
                         const double* Hb, const double* SL, const double* refgeo_Hi, const double* refgeo_Hb, const double* Ti_hom, double* Hi_new) {
  const int nV = mesh.nV;
  std::vector<double> Hi_eff(nV), fm(nV);
  calc_effective_thickness(mesh, C, Hi_new, Hb, SL, Hi_eff.data(), fm.data());   // :55
  if (P.do_protect_grounded_mask && time <= P.protect_grounded_mask_t_end)
    for (int vi = 0; vi < nV; ++vi)
      if (M.grounded_ice[vi]) Hi_new[vi] = std::max(Hi_new[vi], (SL[vi] - Hb[vi]) * C.seawater_density / C.ice_density + 0.1);
  for (int vi = 0; vi < nV; ++vi) {   // :72-119, in the reference's order
    if (Hi_eff[vi] < P.Hi_min) Hi_new[vi] = 0.0;
    const bool fl = is_floating(C, Hi_eff[vi], Hb[vi], SL[vi]);
    if (P.calving_law_threshold_thickness && fl && Hi_eff[vi] < P.calving_threshold_thickness_shelf) Hi_new[vi] = 0.0;
    if (P.remove_ice_absent_at_PD && refgeo_Hi[vi] == 0.0) Hi_new[vi] = 0.0;
    if (P.do_remove_shelves && fl) Hi_new[vi] = 0.0;
    if (P.remove_shelves_larger_than_PD && refgeo_Hi[vi] == 0.0 && refgeo_Hb[vi] < 0.0) Hi_new[vi] = 0.0;
    if (P.continental_shelf_calving && refgeo_Hi[vi] == 0.0 && refgeo_Hb[vi] < P.continental_shelf_min_height) Hi_new[vi] = 0.0;
  }
  const double fixiness = decay_weight(time, P.fixiness_t_start, P.fixiness_t_end, P.do_fixiness_before_start);
  const double limitness = decay_weight(time, P.limitness_t_start, P.limitness_t_end, P.do_limitness_before_start);
  for (int vi = 0; vi < nV; ++vi) {
    double up = 0.0, down = 0.0;   // modiness :177-219
    const bool inland = M.grounded_ice[vi] && !M.gl_gr[vi];
    switch (P.modiness_H_style) {
      case 0: break;
      case 1: up = down = 1.0 - std::exp(Ti_hom[vi] / P.modiness_T_hom_ref); break;
      case 2: up = 1.0 - std::exp(Ti_hom[vi] / P.modiness_T_hom_ref); break;
      case 3: down = 1.0 - std::exp(Ti_hom[vi] / P.modiness_T_hom_ref); break;
      case 4: if (inland) up = 1.0; break;
      case 5: if (inland) down = 1.0; break;
      default: throw std::runtime_error("unknown modiness_H_choice");
    }
    up = std::min(1.0, std::max(0.0, up)); down = std::min(1.0, std::max(0.0, down));
    double fix = 0.0, lim = 0.0;   // :224-262
    if (M.gl_gr[vi]) { fix = P.fixiness_H_gl_gr * fixiness; lim = P.limitness_H_gl_gr * limitness; }
    else if (M.gl_fl[vi]) { fix = P.fixiness_H_gl_fl * fixiness; lim = P.limitness_H_gl_fl * limitness; }
    else if (M.grounded_ice[vi]) { fix = P.fixiness_H_grounded * fixiness; lim = P.limitness_H_grounded * limitness; }
    else if (M.floating_ice[vi]) { fix = P.fixiness_H_floating * fixiness; lim = P.limitness_H_floating * limitness; }
    else if (M.icefree_land[vi]) { if (P.fixiness_H_freeland && fixiness > 0.0) fix = 1.0; lim = P.limitness_H_grounded * limitness; }
    else if (M.icefree_ocean[vi]) { if (P.fixiness_H_freeocean && fixiness > 0.0) fix = 1.0; lim = P.limitness_H_floating * limitness; }
    else throw std::runtime_error("vertex neither grounded, floating, nor ice-free?");
    double H = Hi_old[vi] * fix + Hi_new[vi] * (1.0 - fix);
    H = std::min(H, refgeo_Hi[vi] + (1.0 - up) * lim + (1.0 - limitness) * (H - refgeo_Hi[vi]));
    H = std::max(H, refgeo_Hi[vi] - (1.0 - down) * lim - (1.0 - limitness) * (refgeo_Hi[vi] - H));
    Hi_new[vi] = H;
  }
}

void calc_and_apply_spill_over_flux(const MeshView& mesh, const MaskSet& M, const double* Hi_eff, const double* u_vav_b, const double* v_vav_b,
                                    double dt, double* Hi_new, double* Qspill) {
  const int nV = mesh.nV, nE = mesh.nE, ncm = mesh.nC_mem;
  const double w_eps = (double)1.e-2f;   // `real(dp) :: w_eps = 1.e-2`: a default-real literal (:309)
  // ice%u_perp as calc_ice_flux_divergence_matrix_upwind leaves it (conservation_of_mass_utilities.f90:78-99)
  std::vector<double> u_perp((size_t)nV * ncm, 0.0), relweight((size_t)nV * ncm, 0.0), Q_src(nV, 0.0), Q_dst(nV, 0.0);
  for (int vi = 0; vi < nV; ++vi)
    for (int ci = 0; ci < mesh.nC[vi]; ++ci) {
      const size_t k = vi + (size_t)ci * nV;
      const int ei = mesh.VE[k] - 1, til = mesh.ETri[ei], tir = mesh.ETri[ei + nE];
      double uc, vc;
      if (til == 0) { uc = u_vav_b[tir - 1]; vc = v_vav_b[tir - 1]; }
      else if (tir == 0) { uc = u_vav_b[til - 1]; vc = v_vav_b[til - 1]; }
      else { uc = (u_vav_b[til - 1] + u_vav_b[tir - 1]) / 2.0; vc = (v_vav_b[til - 1] + v_vav_b[tir - 1]) / 2.0; }
      u_perp[k] = uc * mesh.D_x[k] / mesh.D[k] + vc * mesh.D_y[k] / mesh.D[k];
    }
  for (int vi = 0; vi < nV; ++vi) {   // sources :323-392
    const bool cf = M.cf_fl[vi] || M.cf_gr[vi];
    double Hi_ups;
    if (cf) {
      int cm = -1; double umin = std::numeric_limits<double>::max();
      for (int ci = 0; ci < mesh.nC[vi]; ++ci)
        if (u_perp[vi + (size_t)ci * nV] < umin) { umin = u_perp[vi + (size_t)ci * nV]; cm = ci; }
      const int vj = mesh.C[vi + (size_t)cm * nV] - 1;
      if (Hi_new[vj] == 0.0) continue;
      Hi_ups = Hi_new[vj];
    } else {
      Hi_ups = Hi_eff[vi];
    }
    if (cf && Hi_new[vi] > Hi_ups) {
      Q_src[vi] = -(Hi_new[vi] - Hi_ups) * mesh.A[vi] / dt;
      std::vector<double> weight(ncm, 0.0);
      for (int ci = 0; ci < mesh.nC[vi]; ++ci) {
        const int vj = mesh.C[vi + (size_t)ci * nV] - 1;
        if (M.icefree_ocean[vj]) weight[ci] = std::max(0.0, u_perp[vi + (size_t)ci * nV]) + w_eps;
      }
      double sw = 0.0;
      for (int ci = 0; ci < ncm; ++ci) sw += weight[ci];   // sum( weight) over all nC_mem entries
      if (sw < w_eps) Q_src[vi] = 0.0;
      else for (int ci = 0; ci < mesh.nC[vi]; ++ci) relweight[vi + (size_t)ci * nV] = weight[ci] / sw;
    }
  }
  for (int vi = 0; vi < nV; ++vi) {   // destinations :398-424
    if (!M.icefree_ocean[vi]) continue;
    for (int ci = 0; ci < mesh.nC[vi]; ++ci) {
      const int vj = mesh.C[vi + (size_t)ci * nV] - 1;
      for (int cj = 0; cj < mesh.nC[vj]; ++cj)
        if (mesh.C[vj + (size_t)cj * nV] - 1 == vi)
          if (Q_src[vj] < -1.e-2 && relweight[vj + (size_t)cj * nV] > 1.e-6) Q_dst[vi] = Q_dst[vi] - Q_src[vj] * relweight[vj + (size_t)cj * nV];
    }
  }
  for (int vi = 0; vi < nV; ++vi) {
    const double q = (Q_src[vi] + Q_dst[vi]) / mesh.A[vi];
    if (Qspill) Qspill[vi] = q;
    Hi_new[vi] = Hi_new[vi] + q * dt;
  }
}

}  // namespace orc
