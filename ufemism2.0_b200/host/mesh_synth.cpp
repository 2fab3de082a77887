// Synthetic primary meshes for the UFEMISM hot path (input generator; not a reference restatement).
//
// UFEMISM builds its meshes by serial Ruppert refinement + Lloyd smoothing (src/UPSY/mesh/Delaunay,
// mesh_refinement_*.f90 -- out of scope).  The hot path only needs *a* valid
// type_mesh: ccw triangles, Delaunay (so Voronoi cells are well defined), circumcentres inside
// the domain (mesh_utilities.f90:139-142 crashes otherwise), vertices ordered in x
// (mesh_parallelisation.f90:14-22).  This file produces such meshes at any size:
// a near-equilateral lattice, jittered, repaired to Delaunay by Lawson flips, then x-sorted.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <numeric>

#include "ufe_mesh.hpp"

namespace ufe {

namespace {

inline uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
inline double u01(uint64_t x) { return (double)(splitmix64(x) >> 11) * (1.0 / 9007199254740992.0); }

// > 0 if d lies strictly inside the circumcircle of ccw triangle (a,b,c); scaled to be dimensionless
inline double incircle(const double* a, const double* b, const double* c, const double* d) {
  const double ax = a[0] - d[0], ay = a[1] - d[1];
  const double bx = b[0] - d[0], by = b[1] - d[1];
  const double cx = c[0] - d[0], cy = c[1] - d[1];
  const double a2 = ax * ax + ay * ay, b2 = bx * bx + by * by, c2 = cx * cx + cy * cy;
  const double det = ax * (by * c2 - b2 * cy) - ay * (bx * c2 - b2 * cx) + a2 * (bx * cy - by * cx);
  const double scale = (a2 + b2 + c2);
  return det / (scale * scale + 1e-300);
}

struct Work {
  std::vector<double> x, y;   // 0-based
  std::vector<int> tri;       // [t*3+n]
  std::vector<int> nb;        // [t*3+n] neighbour across from vertex n, -1 none
};

// neighbour table from shared edges (sort-based, O(n log n), no hashing)
void build_neighbours(int nTri, const std::vector<int>& tri, std::vector<int>& nb) {
  struct HE { uint64_t key; int slot; };
  std::vector<HE> he((size_t)nTri * 3);
  for (int t = 0; t < nTri; ++t)
    for (int n = 0; n < 3; ++n) {
      int p = tri[t * 3 + (n + 1) % 3], q = tri[t * 3 + (n + 2) % 3];
      uint64_t lo = (uint64_t)std::min(p, q), hi = (uint64_t)std::max(p, q);
      he[(size_t)t * 3 + n] = {(lo << 32) | hi, t * 3 + n};
    }
  std::sort(he.begin(), he.end(), [](const HE& a, const HE& b) { return a.key < b.key; });
  nb.assign((size_t)nTri * 3, -1);
  for (size_t i = 0; i + 1 < he.size(); ++i)
    if (he[i].key == he[i + 1].key) {
      nb[he[i].slot] = he[i + 1].slot / 3;
      nb[he[i + 1].slot] = he[i].slot / 3;
      ++i;
    }
}

// Lawson edge flipping until locally Delaunay everywhere
long lawson_flips(Work& w) {
  const int nTri = (int)w.tri.size() / 3;
  std::vector<int> stack;
  stack.reserve((size_t)nTri * 3);
  for (int s = 0; s < nTri * 3; ++s) stack.push_back(s);
  long nflip = 0;
  const double eps = 1e-12;
  while (!stack.empty()) {
    int s = stack.back();
    stack.pop_back();
    int t = s / 3, n = s % 3;
    int t2 = w.nb[s];
    if (t2 < 0) continue;
    int a = w.tri[t * 3 + n], b = w.tri[t * 3 + (n + 1) % 3], c = w.tri[t * 3 + (n + 2) % 3];
    int m = -1;
    for (int k = 0; k < 3; ++k)
      if (w.nb[t2 * 3 + k] == t) {
        // make sure this is the shared edge (b,c), not another edge shared with the same triangle
        int p = w.tri[t2 * 3 + (k + 1) % 3], q = w.tri[t2 * 3 + (k + 2) % 3];
        if ((p == c && q == b) || (p == b && q == c)) { m = k; break; }
      }
    if (m < 0) continue;  // stale entry
    int d = w.tri[t2 * 3 + m];
    double pa[2] = {w.x[a], w.y[a]}, pb[2] = {w.x[b], w.y[b]}, pc[2] = {w.x[c], w.y[c]}, pd[2] = {w.x[d], w.y[d]};
    if (incircle(pa, pb, pc, pd) <= eps) continue;
    // t = (a,b,c): nb across a = t2, across b = Nb (edge c-a), across c = Nc (edge a-b)
    // t2 = (d,c,b): across d = t, across c = Mc (edge b-d), across b = Mb (edge d-c)
    int Nb = w.nb[t * 3 + (n + 1) % 3], Nc = w.nb[t * 3 + (n + 2) % 3];
    int Mc = w.nb[t2 * 3 + (m + 1) % 3], Mb = w.nb[t2 * 3 + (m + 2) % 3];
    // sanity: t2's vertex after d must be c
    if (w.tri[t2 * 3 + (m + 1) % 3] != c) std::swap(Mc, Mb);
    // new t = (a,b,d): across a = Mc (edge b-d), across b = t2 (edge d-a), across d = Nc (edge a-b)
    w.tri[t * 3 + 0] = a; w.tri[t * 3 + 1] = b; w.tri[t * 3 + 2] = d;
    w.nb[t * 3 + 0] = Mc; w.nb[t * 3 + 1] = t2; w.nb[t * 3 + 2] = Nc;
    // new t2 = (a,d,c): across a = Mb (edge d-c), across d = Nb (edge c-a), across c = t (edge a-d)
    w.tri[t2 * 3 + 0] = a; w.tri[t2 * 3 + 1] = d; w.tri[t2 * 3 + 2] = c;
    w.nb[t2 * 3 + 0] = Mb; w.nb[t2 * 3 + 1] = Nb; w.nb[t2 * 3 + 2] = t;
    if (Mc >= 0)
      for (int k = 0; k < 3; ++k)
        if (w.nb[Mc * 3 + k] == t2) {
          int p = w.tri[Mc * 3 + (k + 1) % 3], q = w.tri[Mc * 3 + (k + 2) % 3];
          if ((p == b && q == d) || (p == d && q == b)) w.nb[Mc * 3 + k] = t;
        }
    if (Nb >= 0)
      for (int k = 0; k < 3; ++k)
        if (w.nb[Nb * 3 + k] == t) {
          int p = w.tri[Nb * 3 + (k + 1) % 3], q = w.tri[Nb * 3 + (k + 2) % 3];
          if ((p == c && q == a) || (p == a && q == c)) w.nb[Nb * 3 + k] = t2;
        }
    ++nflip;
    stack.push_back(t * 3 + 0); stack.push_back(t * 3 + 2);
    stack.push_back(t2 * 3 + 0); stack.push_back(t2 * 3 + 1);
  }
  return nflip;
}

}  // namespace

bool make_synthetic_mesh(const SynthParams& p, HostMesh& m) {
  if (!(p.xmax > p.xmin) || !(p.ymax > p.ymin) || !(p.h > 0)) { m.last_error = "bad domain/h"; return false; }
  m = HostMesh();
  m.xmin = p.xmin; m.xmax = p.xmax; m.ymin = p.ymin; m.ymax = p.ymax;
  m.nC_mem = p.nC_mem; m.nz = p.nz;
  m.tol_dist = ((p.xmax - p.xmin) + (p.ymax - p.ymin)) * 1e-9 / 2.0;  // mesh_dummy_meshes.f90:58

  // column positions (optionally graded in x)
  std::vector<double> xc;
  {
    double x = p.xmin;
    xc.push_back(x);
    const double L = p.xmax - p.xmin;
    while (x < p.xmax) {
      double s = p.h;
      if (p.grade_depth > 0 && p.grade_width > 0) {
        double r = (x - p.grade_x0) / p.grade_width;
        s = p.h * (1.0 - p.grade_depth * std::exp(-r * r));
      }
      x += s;
      xc.push_back(x);
      if (xc.size() > 50000000) { m.last_error = "too many columns"; return false; }
    }
    // rescale so the last column lands exactly on xmax
    double xl = xc.back();
    for (auto& v : xc) v = p.xmin + (v - p.xmin) * (L / (xl - p.xmin));
    xc.back() = p.xmax;
  }
  const int nx = (int)xc.size() - 1;
  const double dy0 = p.h * std::sqrt(3.0) / 2.0;
  const int ny = std::max(2, (int)std::lround((p.ymax - p.ymin) / dy0));
  if (nx < 2) { m.last_error = "mesh too coarse"; return false; }
  const double dy = (p.ymax - p.ymin) / ny;

  Work w;
  const size_t nV = (size_t)(nx + 1) * (ny + 1);
  w.x.resize(nV); w.y.resize(nV);
  auto vid = [&](int i, int j) { return j * (nx + 1) + i; };
  for (int j = 0; j <= ny; ++j)
    for (int i = 0; i <= nx; ++i) {
      double x = xc[i], y = p.ymin + j * dy;
      if (j == ny) y = p.ymax;
      if (i > 0 && i < nx) {
        double hl = 0.5 * (xc[i + 1] - xc[i - 1]);
        x += ((j % 2 == 0) ? 0.25 : -0.25) * hl;
        if (j > 0 && j < ny && p.jitter > 0) {
          double amp = p.jitter * std::min(hl, dy0);
          if (i == 1 || i == nx - 1 || j == 1 || j == ny - 1) amp *= 0.5;
          uint64_t key = p.seed * 0x100000001B3ull + (uint64_t)vid(i, j) * 2ull;
          x += amp * (2.0 * u01(key) - 1.0);
          y += amp * (2.0 * u01(key + 1) - 1.0);
        }
      }
      w.x[vid(i, j)] = x; w.y[vid(i, j)] = y;
    }

  const size_t nTri = (size_t)nx * ny * 2;
  w.tri.resize(nTri * 3);
  {
    size_t t = 0;
    for (int j = 0; j < ny; ++j)
      for (int i = 0; i < nx; ++i) {
        int v00 = vid(i, j), v10 = vid(i + 1, j), v01 = vid(i, j + 1), v11 = vid(i + 1, j + 1);
        if (j % 2 == 0) {
          w.tri[t * 3 + 0] = v00; w.tri[t * 3 + 1] = v10; w.tri[t * 3 + 2] = v11; ++t;
          w.tri[t * 3 + 0] = v00; w.tri[t * 3 + 1] = v11; w.tri[t * 3 + 2] = v01; ++t;
        } else {
          w.tri[t * 3 + 0] = v00; w.tri[t * 3 + 1] = v10; w.tri[t * 3 + 2] = v01; ++t;
          w.tri[t * 3 + 0] = v10; w.tri[t * 3 + 1] = v11; w.tri[t * 3 + 2] = v01; ++t;
        }
      }
  }
  build_neighbours((int)nTri, w.tri, w.nb);
  lawson_flips(w);

  // x-sort vertices (ties by y), then triangles by geometric centre
  std::vector<int> vorder(nV);
  std::iota(vorder.begin(), vorder.end(), 0);
  std::sort(vorder.begin(), vorder.end(), [&](int a, int b) {
    if (w.x[a] != w.x[b]) return w.x[a] < w.x[b];
    return w.y[a] < w.y[b];
  });
  std::vector<int> vnew(nV);
  for (size_t k = 0; k < nV; ++k) vnew[vorder[k]] = (int)k;

  std::vector<double> gx(nTri), gy(nTri);
  for (size_t t = 0; t < nTri; ++t) {
    int a = w.tri[t * 3], b = w.tri[t * 3 + 1], c = w.tri[t * 3 + 2];
    gx[t] = (w.x[a] + w.x[b] + w.x[c]) / 3.0;
    gy[t] = (w.y[a] + w.y[b] + w.y[c]) / 3.0;
  }
  std::vector<int> torder(nTri);
  std::iota(torder.begin(), torder.end(), 0);
  std::sort(torder.begin(), torder.end(), [&](int a, int b) {
    if (gx[a] != gx[b]) return gx[a] < gx[b];
    return gy[a] < gy[b];
  });

  m.nV = (int)nV; m.nTri = (int)nTri;
  m.V.resize(nV * 2);
  for (size_t k = 0; k < nV; ++k) { m.V[k] = w.x[vorder[k]]; m.V[nV + k] = w.y[vorder[k]]; }
  m.Tri.resize(nTri * 3);
  for (size_t k = 0; k < nTri; ++k) {
    int t = torder[k];
    for (int n = 0; n < 3; ++n) m.Tri[(size_t)n * nTri + k] = vnew[w.tri[t * 3 + n]] + 1;
  }
  w = Work();
  if (!build_connectivity(m)) return false;
  return calc_secondary(m, p.zeta_kind, p.zeta_R);
}

bool make_dummy_mesh_5(double xmin, double xmax, double ymin, double ymax, HostMesh& m) {
  m = HostMesh();
  m.xmin = xmin; m.xmax = xmax; m.ymin = ymin; m.ymax = ymax;
  m.tol_dist = ((xmax - xmin) + (ymax - ymin)) * 1e-9 / 2.0;
  m.nV = 5; m.nTri = 4;
  const double xs[5] = {xmin, xmax, xmax, xmin, 0.5 * (xmin + xmax)};
  const double ys[5] = {ymin, ymin, ymax, ymax, 0.5 * (ymin + ymax)};
  m.V.resize(10);
  for (int i = 0; i < 5; ++i) { m.V[i] = xs[i]; m.V[5 + i] = ys[i]; }
  const int tri[4][3] = {{1, 2, 5}, {2, 3, 5}, {3, 4, 5}, {4, 1, 5}};
  m.Tri.resize(12);
  for (int t = 0; t < 4; ++t)
    for (int n = 0; n < 3; ++n) m.Tri[n * 4 + t] = tri[t][n];
  if (!build_connectivity(m)) return false;
  return calc_secondary(m, 0, 10.0);
}

bool build_connectivity(HostMesh& m) {
  const int nV = m.nV, nTri = m.nTri, ncm = m.nC_mem;
  const double* X = m.V.data();
  const double* Y = m.V.data() + nV;
  auto T = [&](int t, int n) { return m.Tri[(size_t)n * nTri + t] - 1; };

  // orientation check + neighbours
  std::vector<int> tri0((size_t)nTri * 3), nb;
  for (int t = 0; t < nTri; ++t) {
    for (int n = 0; n < 3; ++n) tri0[(size_t)t * 3 + n] = T(t, n);
    int a = T(t, 0), b = T(t, 1), c = T(t, 2);
    double cr = (X[b] - X[a]) * (Y[c] - Y[a]) - (Y[b] - Y[a]) * (X[c] - X[a]);
    if (!(cr > 0)) { m.last_error = "triangle not counter-clockwise / degenerate"; return false; }
  }
  build_neighbours(nTri, tri0, nb);
  m.TriC.assign((size_t)nTri * 3, 0);
  for (int t = 0; t < nTri; ++t)
    for (int n = 0; n < 3; ++n) m.TriC[(size_t)n * nTri + t] = nb[(size_t)t * 3 + n] + 1;

  // border index from coordinates
  m.VBI.assign(nV, 0);
  for (int v = 0; v < nV; ++v) {
    bool N = Y[v] == m.ymax, S = Y[v] == m.ymin, Ea = X[v] == m.xmax, W = X[v] == m.xmin;
    int bi = 0;
    if (N && Ea) bi = 2; else if (S && Ea) bi = 4; else if (S && W) bi = 6; else if (N && W) bi = 8;
    else if (N) bi = 1; else if (Ea) bi = 3; else if (S) bi = 5; else if (W) bi = 7;
    m.VBI[v] = bi;
  }

  // one incident (triangle, local index) per vertex
  std::vector<int> v2s(nV, -1);
  for (int t = 0; t < nTri; ++t)
    for (int n = 0; n < 3; ++n) v2s[tri0[(size_t)t * 3 + n]] = t * 3 + n;

  m.nC.assign(nV, 0); m.niTri.assign(nV, 0);
  m.C.assign((size_t)nV * ncm, 0); m.iTri.assign((size_t)nV * ncm, 0);
  std::vector<int> fanC(ncm + 2), fanT(ncm + 2);
  for (int v = 0; v < nV; ++v) {
    int s = v2s[v];
    if (s < 0) { m.last_error = "isolated vertex"; return false; }
    // rotate clockwise until the border or back at the start
    int t = s / 3, n = s % 3;
    int t_start = t;
    bool border = false;
    int guard = 0;
    while (true) {
      int tprev = nb[(size_t)t * 3 + (n + 2) % 3];  // shares edge v - Tri(t,n+1)
      if (tprev < 0) { border = true; break; }
      t = tprev;
      n = 0; while (tri0[(size_t)t * 3 + n] != v) ++n;
      if (t == t_start) break;
      if (++guard > ncm + 2) { m.last_error = "vertex fan exceeds nC_mem"; return false; }
    }
    // walk counter-clockwise collecting the fan
    int nc = 0, nt = 0;
    int t0 = t;
    while (true) {
      int pnext = tri0[(size_t)t * 3 + (n + 1) % 3], q = tri0[(size_t)t * 3 + (n + 2) % 3];
      if (nc + 2 > ncm) { m.last_error = "vertex fan exceeds nC_mem"; return false; }
      if (nt == 0) fanC[nc++] = pnext;
      fanT[nt++] = t;
      int tnext = nb[(size_t)t * 3 + (n + 1) % 3];    // shares edge v - q
      if (tnext < 0) { fanC[nc++] = q; break; }
      if (tnext == t0) break;                          // closed fan: q == first neighbour
      fanC[nc++] = q;
      t = tnext;
      n = 0; while (tri0[(size_t)t * 3 + n] != v) ++n;
    }
    if (border != (m.VBI[v] != 0)) { m.last_error = "border vertex does not lie on the domain border"; return false; }
    int rot = 0;
    if (!border) {  // start interior fans at the lowest-numbered neighbour (as in mesh_dummy_meshes.f90:81)
      for (int k = 1; k < nc; ++k) if (fanC[k] < fanC[rot]) rot = k;
    }
    m.nC[v] = nc; m.niTri[v] = nt;
    for (int k = 0; k < nc; ++k) m.C[(size_t)k * nV + v] = fanC[(k + rot) % nc] + 1;
    for (int k = 0; k < nt; ++k) m.iTri[(size_t)k * nV + v] = fanT[border ? k : (k + rot) % nt] + 1;
  }

  // circumcentres (plane_geometry.f90:282-309: intersection of two perpendicular bisectors)
  m.Tricc.assign((size_t)nTri * 2, 0.0);
  for (int t = 0; t < nTri; ++t) {
    int a = T(t, 0), b = T(t, 1), c = T(t, 2);
    double bx = X[b] - X[a], by = Y[b] - Y[a], cx = X[c] - X[a], cy = Y[c] - Y[a];
    double d = 2.0 * (bx * cy - by * cx);
    double b2 = bx * bx + by * by, c2 = cx * cx + cy * cy;
    double ux = (cy * b2 - by * c2) / d, uy = (bx * c2 - cx * b2) / d;
    double ccx = X[a] + ux, ccy = Y[a] + uy;
    // tolerate round-off for right-angled border triangles
    const double tol = m.tol_dist;
    if (ccx < m.xmin && ccx > m.xmin - tol) ccx = m.xmin;
    if (ccx > m.xmax && ccx < m.xmax + tol) ccx = m.xmax;
    if (ccy < m.ymin && ccy > m.ymin - tol) ccy = m.ymin;
    if (ccy > m.ymax && ccy < m.ymax + tol) ccy = m.ymax;
    if (ccx < m.xmin || ccx > m.xmax || ccy < m.ymin || ccy > m.ymax) {
      m.last_error = "triangle circumcentre outside the mesh domain";
      return false;
    }
    m.Tricc[t] = ccx; m.Tricc[(size_t)nTri + t] = ccy;
  }
  return true;
}

}  // namespace ufe
