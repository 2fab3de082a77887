// TEST INFRASTRUCTURE ONLY (see oracle.hpp).
// CPU restatement of the one-off mesh operator construction:
//   calc_all_matrix_operators_mesh          src/UPSY/mesh/discretisation/mesh_disc_calc_matrix_operators_2D.f90:26-764
//   extend_group_single_iteration_{a,b}     src/UPSY/mesh/mesh_utilities.f90:1856-1935
//   calc_shape_functions_2D_*               src/UPSY/basic/math_utilities/shape_functions.f90:140-451
//   2x2 / 3x3 / 5x5 determinant + inverse   src/UPSY/basic/math_utilities/matrix_algebra.f90:11-457
#include <cfloat>
#include <cmath>
#include <cstring>

#include "oracle.hpp"

namespace orc {

void spmv(const Csr& A, const double* x, double* y) {
#pragma omp parallel for schedule(static)
  for (int i = 1; i <= A.m; ++i) {
    double s = 0.0;
    for (int k = A.ptr[i - 1]; k <= A.ptr[i] - 1; ++k) s = s + A.val[k - 1] * x[A.ind[k - 1] - 1];
    y[i - 1] = s;
  }
}

namespace {

constexpr double q_weight = 1.5;  // shape_functions.f90:13 (Syrakos et al., 2017)
constexpr double tiny_dp = DBL_MIN;

// generic determinant by Laplace expansion along the first row (n <= 5)
double det_n(const double* a, int n, int lda) {
  if (n == 1) return a[0];
  if (n == 2) return a[0] * a[lda + 1] - a[1] * a[lda];
  double sub[16];
  double d = 0.0;
  for (int j = 0; j < n; ++j) {
    for (int r = 1; r < n; ++r) {
      int cc = 0;
      for (int c = 0; c < n; ++c) if (c != j) sub[(r - 1) * (n - 1) + cc++] = a[r * lda + c];
    }
    double t = a[j] * det_n(sub, n - 1, n - 1);
    d += (j % 2 == 0) ? t : -t;
  }
  return d;
}

// inverse by the adjugate (the reference's 3x3 and 5x5 routines are explicit cofactor formulas)
bool inverse_adj(const double* a, int n, double* inv, double* det_out) {
  double det = det_n(a, n, n);
  *det_out = det;
  if (std::fabs(det) < tiny_dp) return false;
  double sub[16];
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) {
      int rr = 0;
      for (int r = 0; r < n; ++r) {
        if (r == i) continue;
        int cc = 0;
        for (int c = 0; c < n; ++c) if (c != j) sub[rr * (n - 1) + cc++] = a[r * n + c];
        ++rr;
      }
      double cof = det_n(sub, n - 1, n - 1);
      if ((i + j) % 2) cof = -cof;
      inv[j * n + i] = cof / det;  // transpose of the cofactor matrix
    }
  return true;
}

struct Nbr { int n; const int* id; const double* x; const double* y; };

// calc_shape_functions_2D_reg_1st_order (shape_functions.f90:140-216)
bool shape_reg_1st(double x, double y, int n_c, const double* xc, const double* yc,
                   double& Nfx_i, double& Nfy_i, double* Nfx_c, double* Nfy_c) {
  double M[4] = {0, 0, 0, 0}, Minv[4], det;
  for (int c = 0; c < n_c; ++c) {
    double dx = xc[c] - x, dy = yc[c] - y;
    double w = 1.0 / std::pow(std::sqrt(dx * dx + dy * dy), q_weight), w2 = w * w;
    M[0] += w2 * dx * dx; M[1] += w2 * dx * dy; M[2] += w2 * dy * dx; M[3] += w2 * dy * dy;
  }
  if (!inverse_adj(M, 2, Minv, &det)) return false;
  Nfx_i = 0; Nfy_i = 0;
  for (int c = 0; c < n_c; ++c) {
    double dx = xc[c] - x, dy = yc[c] - y;
    double w = 1.0 / std::pow(std::sqrt(dx * dx + dy * dy), q_weight), w2 = w * w;
    Nfx_c[c] = w2 * ((Minv[0] * dx) + (Minv[1] * dy));
    Nfy_c[c] = w2 * ((Minv[2] * dx) + (Minv[3] * dy));
  }
  double sx = 0, sy = 0;
  for (int c = 0; c < n_c; ++c) { sx += Nfx_c[c]; sy += Nfy_c[c]; }
  Nfx_i = -sx; Nfy_i = -sy;
  return true;
}

// calc_shape_functions_2D_stag_1st_order (shape_functions.f90:366-451); note "<= tiny" here
bool shape_stag_1st(double x, double y, int n_c, const double* xc, const double* yc,
                    double* Nf_c, double* Nfx_c, double* Nfy_c) {
  double M[9] = {0}, Minv[9];
  for (int c = 0; c < n_c; ++c) {
    double dx = xc[c] - x, dy = yc[c] - y;
    double w = 1.0 / std::pow(std::sqrt(dx * dx + dy * dy), q_weight), w2 = w * w;
    const double b[3] = {1.0, dx, dy};
    for (int r = 0; r < 3; ++r) for (int s = 0; s < 3; ++s) M[r * 3 + s] += (w2 * b[r] * b[s]);
  }
  double det = det_n(M, 3, 3);
  if (std::fabs(det) <= tiny_dp) return false;
  if (!inverse_adj(M, 3, Minv, &det)) return false;
  for (int c = 0; c < n_c; ++c) {
    double dx = xc[c] - x, dy = yc[c] - y;
    double w = 1.0 / std::pow(std::sqrt(dx * dx + dy * dy), q_weight), w2 = w * w;
    Nf_c[c]  = w2 * ((Minv[0] * 1.0) + (Minv[1] * dx) + (Minv[2] * dy));
    Nfx_c[c] = w2 * ((Minv[3] * 1.0) + (Minv[4] * dx) + (Minv[5] * dy));
    Nfy_c[c] = w2 * ((Minv[6] * 1.0) + (Minv[7] * dx) + (Minv[8] * dy));
  }
  return true;
}

// calc_shape_functions_2D_reg_2nd_order (shape_functions.f90:218-364)
bool shape_reg_2nd(double x, double y, int n_c, const double* xc, const double* yc,
                   double Ni[5], double* Nc /* [5][n_c] */) {
  double M[25] = {0}, Minv[25], det;
  for (int c = 0; c < n_c; ++c) {
    double dx = xc[c] - x, dy = yc[c] - y;
    double w = 1.0 / std::pow(std::sqrt(dx * dx + dy * dy), q_weight), w2 = w * w;
    // Taylor basis [dx, dy, 1/2 dx^2, dx dy, 1/2 dy^2]
    const double b[5] = {dx, dy, dx * dx / 2.0, dx * dy, dy * dy / 2.0};
    for (int r = 0; r < 5; ++r) for (int s = 0; s < 5; ++s) M[r * 5 + s] += w2 * b[r] * b[s];
  }
  if (!inverse_adj(M, 5, Minv, &det)) return false;
  for (int r = 0; r < 5; ++r) Ni[r] = 0.0;
  for (int c = 0; c < n_c; ++c) {
    double dx = xc[c] - x, dy = yc[c] - y;
    double w = 1.0 / std::pow(std::sqrt(dx * dx + dy * dy), q_weight), w2 = w * w;
    const double b[5] = {dx, dy, dx * dx / 2.0, dx * dy, dy * dy / 2.0};
    for (int r = 0; r < 5; ++r) {
      double s = 0.0;
      for (int k = 0; k < 5; ++k) s += Minv[r * 5 + k] * b[k];
      Nc[(size_t)r * n_c + c] = w2 * s;
    }
  }
  for (int r = 0; r < 5; ++r) { double s = 0; for (int c = 0; c < n_c; ++c) s += Nc[(size_t)r * n_c + c]; Ni[r] = -s; }
  return true;
}

struct Flood {  // map + stack of extend_group_single_iteration_{a,b}
  std::vector<int> map, stack; int stackN = 0;
  explicit Flood(int n) : map(n + 1, 0), stack(n + 1, 0) {}
  void reset() { for (int i = 1; i <= stackN; ++i) map[stack[i]] = 0; stackN = 0; }
  void push(int v) { map[v] = 1; stack[++stackN] = v; }
};

void extend_a(const MeshView& m, Flood& f) {
  int n = f.stackN;
  for (int i = 1; i <= n; ++i) {
    int vi = f.stack[i];
    for (int ci = 1; ci <= m.nC[vi - 1]; ++ci) {
      int vj = m.C[(size_t)(ci - 1) * m.nV + vi - 1];
      if (f.map[vj] == 0) f.push(vj);
    }
  }
}
void extend_b(const MeshView& m, Flood& f) {
  int n = f.stackN;
  for (int i = 1; i <= n; ++i) {
    int ti = f.stack[i];
    for (int n2 = 1; n2 <= 3; ++n2) {
      int tj = m.TriC[(size_t)(n2 - 1) * m.nTri + ti - 1];
      if (tj == 0) continue;
      if (f.map[tj] == 0) f.push(tj);
    }
  }
}

}  // namespace

void calc_all_matrix_operators(const MeshView& m, Operators& ops) {
  const int nV = m.nV, nTri = m.nTri;
  const int nmax = m.nC_mem * m.nC_mem;
  auto Vx = [&](int vi) { return m.V[vi - 1]; };
  auto Vy = [&](int vi) { return m.V[(size_t)nV + vi - 1]; };
  auto Gx = [&](int ti) { return m.TriGC[ti - 1]; };
  auto Gy = [&](int ti) { return m.TriGC[(size_t)nTri + ti - 1]; };
  std::vector<int> i_c(nmax);
  std::vector<double> x_c(nmax), y_c(nmax), N1(nmax), N2(nmax), N3(nmax), N5((size_t)5 * nmax);

  // ---- a -> a (mesh_disc_calc_matrix_operators_2D.f90:61-196), n_neighbours_min = 2
  {
    Csr &Mx = ops.M[OP_ddx_a_a], &My = ops.M[OP_ddy_a_a];
    Mx.init(nV, nV, (size_t)nV * 8); My.init(nV, nV, (size_t)nV * 8);
    Flood f(nV);
    for (int row = 1; row <= nV; ++row) {
      int vi = row;
      f.reset(); f.push(vi);
      while (f.stackN - 1 < 2) extend_a(m, f);
      double Nfx_i = 0, Nfy_i = 0; int n_c = 0;
      for (bool ok = false; !ok;) {
        n_c = 0;
        for (int i = 1; i <= f.stackN && n_c < nmax; ++i) {
          int vj = f.stack[i]; if (vj == vi) continue;
          i_c[n_c] = vj; x_c[n_c] = Vx(vj); y_c[n_c] = Vy(vj); ++n_c;
        }
        ok = shape_reg_1st(Vx(vi), Vy(vi), n_c, x_c.data(), y_c.data(), Nfx_i, Nfy_i, N1.data(), N2.data());
        if (!ok) extend_a(m, f);
      }
      Mx.add(row, row, Nfx_i); My.add(row, row, Nfy_i);
      for (int i = 0; i < n_c; ++i) { Mx.add(row, i_c[i], N1[i]); My.add(row, i_c[i], N2[i]); }
    }
    Mx.finalise(); My.finalise();
  }
  // ---- a -> b (:198-335), start = the three vertices of ti, n_neighbours_min = 3
  {
    Csr &M0 = ops.M[OP_map_a_b], &Mx = ops.M[OP_ddx_a_b], &My = ops.M[OP_ddy_a_b];
    M0.init(nTri, nV, (size_t)nTri * 3); Mx.init(nTri, nV, (size_t)nTri * 3); My.init(nTri, nV, (size_t)nTri * 3);
    Flood f(nV);
    for (int row = 1; row <= nTri; ++row) {
      int ti = row;
      f.reset();
      for (int n = 1; n <= 3; ++n) f.push(m.Tri[(size_t)(n - 1) * nTri + ti - 1]);
      while (f.stackN < 3) extend_a(m, f);
      int n_c = 0;
      for (bool ok = false; !ok;) {
        n_c = 0;
        for (int i = 1; i <= f.stackN && n_c < nmax; ++i) {
          int vi = f.stack[i]; i_c[n_c] = vi; x_c[n_c] = Vx(vi); y_c[n_c] = Vy(vi); ++n_c;
        }
        ok = shape_stag_1st(Gx(ti), Gy(ti), n_c, x_c.data(), y_c.data(), N1.data(), N2.data(), N3.data());
        if (!ok) extend_a(m, f);
      }
      for (int i = 0; i < n_c; ++i) { M0.add(row, i_c[i], N1[i]); Mx.add(row, i_c[i], N2[i]); My.add(row, i_c[i], N3[i]); }
    }
    M0.finalise(); Mx.finalise(); My.finalise();
  }
  // ---- b -> a (:337-474), start = iTri(vi,:), n_neighbours_min = 3
  {
    Csr &M0 = ops.M[OP_map_b_a], &Mx = ops.M[OP_ddx_b_a], &My = ops.M[OP_ddy_b_a];
    M0.init(nV, nTri, (size_t)nV * 7); Mx.init(nV, nTri, (size_t)nV * 7); My.init(nV, nTri, (size_t)nV * 7);
    Flood f(nTri);
    for (int row = 1; row <= nV; ++row) {
      int vi = row;
      f.reset();
      for (int iti = 1; iti <= m.niTri[vi - 1]; ++iti) f.push(m.iTri[(size_t)(iti - 1) * nV + vi - 1]);
      while (f.stackN < 3) extend_b(m, f);
      int n_c = 0;
      for (bool ok = false; !ok;) {
        n_c = 0;
        for (int i = 1; i <= f.stackN && n_c < nmax; ++i) {
          int ti = f.stack[i]; i_c[n_c] = ti; x_c[n_c] = Gx(ti); y_c[n_c] = Gy(ti); ++n_c;
        }
        ok = shape_stag_1st(Vx(vi), Vy(vi), n_c, x_c.data(), y_c.data(), N1.data(), N2.data(), N3.data());
        if (!ok) extend_b(m, f);
      }
      for (int i = 0; i < n_c; ++i) { M0.add(row, i_c[i], N1[i]); Mx.add(row, i_c[i], N2[i]); My.add(row, i_c[i], N3[i]); }
    }
    M0.finalise(); Mx.finalise(); My.finalise();
  }
  // ---- b -> b 1st order (:476-610), n_neighbours_min = 2
  {
    Csr &Mx = ops.M[OP_ddx_b_b], &My = ops.M[OP_ddy_b_b];
    Mx.init(nTri, nTri, (size_t)nTri * 4); My.init(nTri, nTri, (size_t)nTri * 4);
    Flood f(nTri);
    for (int row = 1; row <= nTri; ++row) {
      int ti = row;
      f.reset(); f.push(ti);
      while (f.stackN - 1 < 2) extend_b(m, f);
      double Nfx_i = 0, Nfy_i = 0; int n_c = 0;
      for (bool ok = false; !ok;) {
        n_c = 0;
        for (int i = 1; i <= f.stackN && n_c < nmax; ++i) {
          int tj = f.stack[i]; if (tj == ti) continue;
          i_c[n_c] = tj; x_c[n_c] = Gx(tj); y_c[n_c] = Gy(tj); ++n_c;
        }
        ok = shape_reg_1st(Gx(ti), Gy(ti), n_c, x_c.data(), y_c.data(), Nfx_i, Nfy_i, N1.data(), N2.data());
        if (!ok) extend_b(m, f);
      }
      Mx.add(row, row, Nfx_i); My.add(row, row, Nfy_i);
      for (int i = 0; i < n_c; ++i) { Mx.add(row, i_c[i], N1[i]); My.add(row, i_c[i], N2[i]); }
    }
    Mx.finalise(); My.finalise();
  }
  // ---- b -> b 2nd order (:612-764), n_neighbours_min = 5
  {
    Csr* M5[5] = {&ops.M[OP2_ddx_b_b], &ops.M[OP2_ddy_b_b], &ops.M[OP2_d2dx2_b_b], &ops.M[OP2_d2dxdy_b_b], &ops.M[OP2_d2dy2_b_b]};
    for (auto* M : M5) M->init(nTri, nTri, (size_t)nTri * 10);
    Flood f(nTri);
    for (int row = 1; row <= nTri; ++row) {
      int ti = row;
      f.reset(); f.push(ti);
      while (f.stackN - 1 < 5) extend_b(m, f);
      double Ni[5]; int n_c = 0;
      for (bool ok = false; !ok;) {
        n_c = 0;
        for (int i = 1; i <= f.stackN && n_c < nmax; ++i) {
          int tj = f.stack[i]; if (tj == ti) continue;
          i_c[n_c] = tj; x_c[n_c] = Gx(tj); y_c[n_c] = Gy(tj); ++n_c;
        }
        ok = shape_reg_2nd(Gx(ti), Gy(ti), n_c, x_c.data(), y_c.data(), Ni, N5.data());
        if (!ok) extend_b(m, f);
      }
      for (int r = 0; r < 5; ++r) {
        M5[r]->add(row, row, Ni[r]);
        for (int i = 0; i < n_c; ++i) M5[r]->add(row, i_c[i], N5[(size_t)r * n_c + i]);
      }
    }
    for (auto* M : M5) M->finalise();
  }
}

}  // namespace orc
