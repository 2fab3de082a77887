// Secondary mesh data the hot path reads: edges (c-grid), TriBI, Voronoi cell areas, connection
// widths/lengths, triangle centres/areas, zeta.  Host-side mirror of calc_all_secondary_mesh_data
// (reference: src/UPSY/mesh/mesh_secondary.f90:31-70), needed as *inputs* of the device path.
#include <algorithm>
#include <cmath>

#include "ufe_mesh.hpp"

namespace ufe {

namespace {
struct Acc {  // Fortran-style accessors on the column-major arrays (1-based)
  const HostMesh& m;
  int C(int vi, int ci) const { return m.C[(size_t)(ci - 1) * m.nV + vi - 1]; }
  int iTri(int vi, int k) const { return m.iTri[(size_t)(k - 1) * m.nV + vi - 1]; }
  int Tri(int ti, int n) const { return m.Tri[(size_t)(n - 1) * m.nTri + ti - 1]; }
  int TriC(int ti, int n) const { return m.TriC[(size_t)(n - 1) * m.nTri + ti - 1]; }
  double Vx(int vi) const { return m.V[vi - 1]; }
  double Vy(int vi) const { return m.V[(size_t)m.nV + vi - 1]; }
  double ccx(int ti) const { return m.Tricc[ti - 1]; }
  double ccy(int ti) const { return m.Tricc[(size_t)m.nTri + ti - 1]; }
};

// edge_border_index, mesh_edges.f90:191-224
int edge_border_index(const HostMesh& m, int vi, int vj) {
  int a = m.VBI[vi - 1], b = m.VBI[vj - 1];
  auto in = [](int v, int p, int q, int r) { return v == p || v == q || v == r; };
  if (a == 0 || b == 0) return 0;
  if (in(a, 8, 1, 2) && in(b, 8, 1, 2)) return 1;
  if (in(a, 2, 3, 4) && in(b, 2, 3, 4)) return 3;
  if (in(a, 4, 5, 6) && in(b, 4, 5, 6)) return 5;
  if (in(a, 6, 7, 8) && in(b, 6, 7, 8)) return 7;
  return 0;
}
}  // namespace

// construct_mesh_edges, mesh_edges.f90:19-189
static bool construct_edges(HostMesh& m) {
  Acc a{m};
  const int nV = m.nV, nTri = m.nTri, ncm = m.nC_mem;
  long sumC = 0;
  for (int v = 0; v < nV; ++v) sumC += m.nC[v];
  m.nE = (int)(sumC / 2);
  const int nE = m.nE;
  m.E.assign((size_t)nE * 2, 0.0);
  m.VE.assign((size_t)nV * ncm, 0);
  m.EV.assign((size_t)nE * 4, 0);
  m.ETri.assign((size_t)nE * 2, 0);
  m.TriE.assign((size_t)nTri * 3, 0);
  m.EBI.assign(nE, 0);
  int ei = 0;
  for (int vi = 1; vi <= nV; ++vi)
    for (int ci = 1; ci <= m.nC[vi - 1]; ++ci) {
      int vj = a.C(vi, ci);
      if (m.VE[(size_t)(ci - 1) * nV + vi - 1] > 0) continue;  // already seen from the other side
      ++ei;
      if (ei > nE) { m.last_error = "edge count mismatch"; return false; }
      m.E[ei - 1] = (a.Vx(vi) + a.Vx(vj)) / 2.0;
      m.E[(size_t)nE + ei - 1] = (a.Vy(vi) + a.Vy(vj)) / 2.0;
      m.VE[(size_t)(ci - 1) * nV + vi - 1] = ei;
      for (int cj = 1; cj <= m.nC[vj - 1]; ++cj)
        if (a.C(vj, cj) == vi) { m.VE[(size_t)(cj - 1) * nV + vj - 1] = ei; break; }
      m.EBI[ei - 1] = edge_border_index(m, vi, vj);
      int til = 0, vil = 0, tir = 0, vir = 0;
      for (int iti = 1; iti <= m.niTri[vi - 1]; ++iti) {
        int ti = a.iTri(vi, iti);
        for (int n1 = 1; n1 <= 3; ++n1) {
          int n2 = n1 % 3 + 1, n3 = n2 % 3 + 1;
          if (a.Tri(ti, n1) == vi && a.Tri(ti, n2) == vj) { til = ti; vil = a.Tri(ti, n3); }
          if (a.Tri(ti, n1) == vj && a.Tri(ti, n2) == vi) { tir = ti; vir = a.Tri(ti, n3); }
        }
      }
      m.EV[ei - 1] = vi; m.EV[(size_t)nE + ei - 1] = vj;
      m.EV[(size_t)2 * nE + ei - 1] = vil; m.EV[(size_t)3 * nE + ei - 1] = vir;
      m.ETri[ei - 1] = til; m.ETri[(size_t)nE + ei - 1] = tir;
    }
  if (ei != nE) { m.last_error = "edge count mismatch"; return false; }
  // TriE: edge across from vertex ci
  for (int ti = 1; ti <= nTri; ++ti)
    for (int ci = 1; ci <= 3; ++ci) {
      int v1 = a.Tri(ti, ci % 3 + 1), v2 = a.Tri(ti, (ci + 1) % 3 + 1);
      for (int cj = 1; cj <= m.nC[v1 - 1]; ++cj)
        if (a.C(v1, cj) == v2) { m.TriE[(size_t)(ci - 1) * nTri + ti - 1] = m.VE[(size_t)(cj - 1) * nV + v1 - 1]; break; }
    }
  return true;
}

// calc_TriBI, mesh_secondary.f90:72-135: trace the border from the SW corner
static bool calc_TriBI(HostMesh& m) {
  Acc a{m};
  const int nV = m.nV;
  m.TriBI.assign(m.nTri, 0);
  int vi_SW = 0;
  for (int vi = 1; vi <= nV; ++vi) if (m.VBI[vi - 1] == 6) { vi_SW = vi; break; }
  if (!vi_SW) { m.last_error = "no south-west corner vertex"; return false; }
  int vi = vi_SW, n_cycles = 0, vi_SE = 0, vi_NE = 0, vi_NW = 0;
  while (true) {
    if (++n_cycles > nV) { m.last_error = "got stuck tracing the domain border"; return false; }
    for (int iti = 1; iti <= m.niTri[vi - 1]; ++iti) m.TriBI[a.iTri(vi, iti) - 1] = m.VBI[vi - 1];
    vi = a.C(vi, m.nC[vi - 1]);
    if (m.VBI[vi - 1] == 4) vi_SE = vi;
    if (m.VBI[vi - 1] == 2) vi_NE = vi;
    if (m.VBI[vi - 1] == 8) vi_NW = vi;
    if (vi == vi_SW) break;
  }
  if (!vi_SE || !vi_NE || !vi_NW) { m.last_error = "missing corner vertex"; return false; }
  if (m.niTri[vi_SW - 1] == 1) m.TriBI[a.iTri(vi_SW, 1) - 1] = 6;
  if (m.niTri[vi_SE - 1] == 1) m.TriBI[a.iTri(vi_SE, 1) - 1] = 4;
  if (m.niTri[vi_NE - 1] == 1) m.TriBI[a.iTri(vi_NE, 1) - 1] = 2;
  if (m.niTri[vi_NW - 1] == 1) m.TriBI[a.iTri(vi_NW, 1) - 1] = 8;
  return true;
}

// calc_Voronoi_cell (dx = 0) + shoelace, mesh_utilities.f90:72-303, mesh_secondary.f90:137-186
static void voronoi_areas(HostMesh& m) {
  Acc a{m};
  const int nV = m.nV;
  m.A.assign(nV, 0.0);
  std::vector<double> vx(m.nC_mem + 4), vy(m.nC_mem + 4);
  for (int vi = 1; vi <= nV; ++vi) {
    int nVor = 0;
    const int bi = m.VBI[vi - 1];
    if (bi == 0) {
      for (int iti = 1; iti <= m.niTri[vi - 1]; ++iti) {
        int ti = a.iTri(vi, iti);
        vx[nVor] = a.ccx(ti); vy[nVor] = a.ccy(ti); ++nVor;
      }
    } else {
      // slot 0 reserved for the projection of the first circumcentre on the border
      nVor = 1;
      for (int ci = 2; ci <= m.nC[vi - 1]; ++ci) {
        int vj = a.C(vi, ci), ti = 0;
        for (int iti = 1; iti <= m.niTri[vi - 1] && !ti; ++iti) {
          int tj = a.iTri(vi, iti);
          for (int n = 1; n <= 3; ++n)
            if (a.Tri(tj, n) == vj && a.Tri(tj, n % 3 + 1) == vi) { ti = tj; break; }
        }
        vx[nVor] = a.ccx(ti); vy[nVor] = a.ccy(ti); ++nVor;
      }
      if (bi == 1 || bi == 2) { vx[0] = vx[1]; vy[0] = m.ymax; }
      else if (bi == 3 || bi == 4) { vx[0] = m.xmax; vy[0] = vy[1]; }
      else if (bi == 5 || bi == 6) { vx[0] = vx[1]; vy[0] = m.ymin; }
      else { vx[0] = m.xmin; vy[0] = vy[1]; }
      if (bi == 2 || bi == 3) { vx[nVor] = m.xmax; vy[nVor] = vy[nVor - 1]; }
      else if (bi == 4 || bi == 5) { vx[nVor] = vx[nVor - 1]; vy[nVor] = m.ymin; }
      else if (bi == 6 || bi == 7) { vx[nVor] = m.xmin; vy[nVor] = vy[nVor - 1]; }
      else { vx[nVor] = vx[nVor - 1]; vy[nVor] = m.ymax; }
      ++nVor;
      if (bi == 2) { vx[nVor] = m.xmax; vy[nVor] = m.ymax; ++nVor; }
      else if (bi == 4) { vx[nVor] = m.xmax; vy[nVor] = m.ymin; ++nVor; }
      else if (bi == 6) { vx[nVor] = m.xmin; vy[nVor] = m.ymin; ++nVor; }
      else if (bi == 8) { vx[nVor] = m.xmin; vy[nVor] = m.ymax; ++nVor; }
    }
    double area = 0.0;
    const double x0 = a.Vx(vi), y0 = a.Vy(vi);
    for (int k = 0; k < nVor; ++k) {
      int kn = (k + 1 == nVor) ? 0 : k + 1;
      double ax = vx[kn] - x0, ay = vy[kn] - y0, bx = vx[k] - x0, by = vy[k] - y0;
      area += std::fabs(ax * by - ay * bx) / 2.0;
    }
    m.A[vi - 1] = area;
  }
}

// calc_connection_widths / find_shared_Voronoi_boundary, mesh_secondary.f90:247-298, mesh_utilities.f90:305-372
static void connection_widths_lengths(HostMesh& m) {
  Acc a{m};
  const int nV = m.nV, nE = m.nE, ncm = m.nC_mem;
  const double dx = ((m.xmax - m.xmin) + (m.ymax - m.ymin)) / 100.0;
  m.Cw.assign((size_t)nV * ncm, 0.0);
  m.D_x.assign((size_t)nV * ncm, 0.0);
  m.D_y.assign((size_t)nV * ncm, 0.0);
  m.D.assign((size_t)nV * ncm, 0.0);
  for (int vi = 1; vi <= nV; ++vi)
    for (int ci = 1; ci <= m.nC[vi - 1]; ++ci) {
      const size_t k = (size_t)(ci - 1) * nV + vi - 1;
      int ei = m.VE[k];
      int til = m.ETri[ei - 1], tir = m.ETri[(size_t)nE + ei - 1];
      double p[2], q[2];
      if (m.EBI[ei - 1] == 0) {
        p[0] = a.ccx(til); p[1] = a.ccy(til); q[0] = a.ccx(tir); q[1] = a.ccy(tir);
      } else {
        int ti = til > 0 ? til : tir;
        p[0] = a.ccx(ti); p[1] = a.ccy(ti);
        switch (m.EBI[ei - 1]) {
          case 1: q[0] = p[0]; q[1] = m.ymax + dx; break;
          case 3: q[0] = m.xmax + dx; q[1] = p[1]; break;
          case 5: q[0] = p[0]; q[1] = m.ymin - dx; break;
          default: q[0] = m.xmin - dx; q[1] = p[1]; break;
        }
        p[0] = std::min(m.xmax, std::max(m.xmin, p[0]));
        p[1] = std::min(m.ymax, std::max(m.ymin, p[1]));
      }
      m.Cw[k] = std::hypot(p[0] - q[0], p[1] - q[1]);
      int vj = a.C(vi, ci);
      double ddx = a.Vx(vj) - a.Vx(vi), ddy = a.Vy(vj) - a.Vy(vi);
      m.D_x[k] = ddx; m.D_y[k] = ddy; m.D[k] = std::sqrt(ddx * ddx + ddy * ddy);
    }
}

static void triangle_centres_areas(HostMesh& m) {
  Acc a{m};
  const int nTri = m.nTri;
  m.TriGC.assign((size_t)nTri * 2, 0.0);
  m.TriA.assign(nTri, 0.0);
  for (int ti = 1; ti <= nTri; ++ti) {
    int p = a.Tri(ti, 1), q = a.Tri(ti, 2), r = a.Tri(ti, 3);
    m.TriGC[ti - 1] = (a.Vx(p) + a.Vx(q) + a.Vx(r)) / 3.0;
    m.TriGC[(size_t)nTri + ti - 1] = (a.Vy(p) + a.Vy(q) + a.Vy(r)) / 3.0;
    double cr = (a.Vx(q) - a.Vx(p)) * (a.Vy(r) - a.Vy(p)) - (a.Vy(q) - a.Vy(p)) * (a.Vx(r) - a.Vx(p));
    m.TriA[ti - 1] = std::fabs(cr) / 2.0;
  }
}

// mesh_zeta.f90:63-124
static void make_zeta(HostMesh& m, int kind, double R) {
  const int nz = m.nz;
  m.zeta.assign(nz, 0.0);
  if (kind == 0 || R == 1.0) {
    for (int k = 1; k <= nz; ++k) m.zeta[k - 1] = (double)(k - 1) / (double)(nz - 1);
  } else {
    for (int k = 1; k <= nz; ++k) {
      double sigma = (double)(k - 1) / (double)(nz - 1);
      m.zeta[nz - k] = 1.0 - (std::pow(R, sigma) - 1.0) / (R - 1.0);
    }
  }
}

bool calc_secondary(HostMesh& m, int zeta_kind, double zeta_R) {
  if (!construct_edges(m)) return false;
  if (!calc_TriBI(m)) return false;
  voronoi_areas(m);
  connection_widths_lengths(m);
  triangle_centres_areas(m);
  make_zeta(m, zeta_kind, zeta_R);
  return true;
}

// partition_list, mpi_distributed_memory.f90:47-73
void partition_list(int ntot, int i, int n, int& i1, int& i2) {
  if (ntot > n * 2) {
    int remainder = ntot % n, slice = ntot / n;
    i1 = slice * i + std::min(i, remainder) + 1;
    i2 = slice * (i + 1) + std::min(i + 1, remainder);
  } else if (i == 0) {
    i1 = 1; i2 = ntot;
  } else {
    i1 = 1; i2 = 0;
  }
}

namespace {
int owner(int idx, int ntot, int n) {  // which rank owns 1-based idx
  for (int r = 0; r < n; ++r) { int i1, i2; partition_list(ntot, r, n, i1, i2); if (idx >= i1 && idx <= i2) return r; }
  return -1;
}
struct Own {
  std::vector<int> b;  // start index of every rank (1-based), size n+1
  int of(int idx) const { return (int)(std::upper_bound(b.begin(), b.end(), idx) - b.begin()) - 1; }
};
Own make_own(int ntot, int n) {
  Own o; o.b.resize(n + 1);
  for (int r = 0; r < n; ++r) { int i1, i2; partition_list(ntot, r, n, i1, i2); o.b[r] = (i2 >= i1) ? i1 : ntot + 1; }
  o.b[n] = ntot + 1;
  // ranks with empty ranges must not capture indices
  for (int r = n - 1; r >= 0; --r) if (o.b[r] > o.b[r + 1]) o.b[r] = o.b[r + 1];
  (void)owner;
  return o;
}
}  // namespace

// determine_halo_range_{a,b,c}, mesh_parallelisation.f90:410-625
void determine_halos(const HostMesh& m, int rank, int nranks, HaloRanges& ha, HaloRanges& hb, HaloRanges& hc) {
  Acc a{m};
  const int nV = m.nV, nTri = m.nTri, nE = m.nE;
  Own oV = make_own(nV, nranks), oT = make_own(nTri, nranks), oE = make_own(nE, nranks);
  auto VEat = [&](int vi, int ci) { return m.VE[(size_t)(ci - 1) * nV + vi - 1]; };
  auto TriEat = [&](int ti, int n) { return m.TriE[(size_t)(n - 1) * nTri + ti - 1]; };

  auto range_a = [&](int node_in, int node_next, int& r1, int& r2) {
    r1 = nV; r2 = 1;
    int i1, i2; partition_list(nV, node_in, nranks, i1, i2);
    for (int vi = i1; vi <= i2; ++vi) {
      bool hit = false;
      for (int ci = 1; ci <= m.nC[vi - 1] && !hit; ++ci) hit = oV.of(a.C(vi, ci)) == node_next;
      for (int k = 1; k <= m.niTri[vi - 1] && !hit; ++k) hit = oT.of(a.iTri(vi, k)) == node_next;
      for (int ci = 1; ci <= m.nC[vi - 1] && !hit; ++ci) hit = oE.of(VEat(vi, ci)) == node_next;
      if (hit) { r1 = std::min(r1, vi); r2 = std::max(r2, vi); }
    }
  };
  auto range_b = [&](int node_in, int node_next, int& r1, int& r2) {
    r1 = nTri; r2 = 1;
    int i1, i2; partition_list(nTri, node_in, nranks, i1, i2);
    for (int ti = i1; ti <= i2; ++ti) {
      bool hit = false;
      for (int n = 1; n <= 3 && !hit; ++n) hit = oV.of(a.Tri(ti, n)) == node_next;
      for (int n = 1; n <= 3 && !hit; ++n) { int tj = a.TriC(ti, n); if (tj > 0) hit = oT.of(tj) == node_next; }
      for (int n = 1; n <= 3 && !hit; ++n) hit = oE.of(TriEat(ti, n)) == node_next;
      if (hit) { r1 = std::min(r1, ti); r2 = std::max(r2, ti); }
    }
  };
  auto range_c = [&](int node_in, int node_next, int& r1, int& r2) {
    r1 = nE; r2 = 1;
    int i1, i2; partition_list(nE, node_in, nranks, i1, i2);
    for (int ei = i1; ei <= i2; ++ei) {
      bool hit = false;
      for (int k = 0; k < 4 && !hit; ++k) { int v = m.EV[(size_t)k * nE + ei - 1]; if (v > 0) hit = oV.of(v) == node_next; }
      for (int k = 0; k < 2 && !hit; ++k) { int t = m.ETri[(size_t)k * nE + ei - 1]; if (t > 0) hit = oT.of(t) == node_next; }
      if (hit) { r1 = std::min(r1, ei); r2 = std::max(r2, ei); }
    }
  };

  auto fill = [&](HaloRanges& h, int ntot, auto range_fn) {
    partition_list(ntot, rank, nranks, h.i1, h.i2);
    if (rank == 0) { h.i1_nih = h.i1; h.i1_hle = 0; h.i2_hle = -1; h.i1_hli = 0; h.i2_hli = -1; }
    else {
      range_fn(rank - 1, rank, h.i1_hle, h.i2_hle);
      range_fn(rank, rank - 1, h.i1_hli, h.i2_hli);
      h.i1_nih = h.i1_hle;
    }
    if (rank == nranks - 1) { h.i2_nih = h.i2; h.i1_hre = 0; h.i2_hre = -1; h.i1_hri = 0; h.i2_hri = -1; }
    else {
      range_fn(rank + 1, rank, h.i1_hre, h.i2_hre);
      range_fn(rank, rank + 1, h.i1_hri, h.i2_hri);
      h.i2_nih = h.i2_hre;
    }
  };
  fill(ha, nV, range_a);
  fill(hb, nTri, range_b);
  fill(hc, nE, range_c);
}

}  // namespace ufe
