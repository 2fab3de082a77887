// Least-squares shape functions (Syrakos et al., 2017) on the device.
// Same mathematics as calc_shape_functions_2D_{reg_1st,reg_2nd,stag_1st}_order
// (reference: src/UPSY/basic/math_utilities/shape_functions.f90:140-451, weight exponent q = 1.5 at :13) with the small
// normal matrices inverted through the adjugate, like the reference's explicit 2x2/3x3/5x5 formulas
// (src/UPSY/basic/math_utilities/matrix_algebra.f90:11-457).
#pragma once
#include <cfloat>

namespace ufe {

template <int N>
struct SmallMat {
  // determinant by Laplace expansion along the first row; a is row-major with leading dimension N
  __device__ static double det(const double* a) {
    double sub[(N - 1) * (N - 1)];
    double d = 0.0;
#pragma unroll
    for (int j = 0; j < N; ++j) {
      for (int r = 1; r < N; ++r) {
        int cc = 0;
        for (int c = 0; c < N; ++c)
          if (c != j) sub[(r - 1) * (N - 1) + cc++] = a[r * N + c];
      }
      const double t = a[j] * SmallMat<N - 1>::det(sub);
      d += (j % 2 == 0) ? t : -t;
    }
    return d;
  }
  // inverse = adj(A) / det(A); returns false when |det| < tiny (the reference's "singular -> add neighbours" test)
  __device__ static bool inverse(const double* a, double* inv, bool le_tiny) {
    const double dt = det(a);
    if (le_tiny ? (fabs(dt) <= DBL_MIN) : (fabs(dt) < DBL_MIN)) return false;
    double sub[(N - 1) * (N - 1)];
    for (int i = 0; i < N; ++i)
      for (int j = 0; j < N; ++j) {
        int rr = 0;
        for (int r = 0; r < N; ++r) {
          if (r == i) continue;
          int cc = 0;
          for (int c = 0; c < N; ++c)
            if (c != j) sub[rr * (N - 1) + cc++] = a[r * N + c];
          ++rr;
        }
        double cof = SmallMat<N - 1>::det(sub);
        if ((i + j) % 2) cof = -cof;
        inv[j * N + i] = cof / dt;
      }
    return true;
  }
};
template <>
struct SmallMat<1> {
  __device__ static double det(const double* a) { return a[0]; }
};
template <>
struct SmallMat<2> {
  __device__ static double det(const double* a) { return a[0] * a[3] - a[1] * a[2]; }
  __device__ static bool inverse(const double* a, double* inv, bool le_tiny) {
    const double dt = det(a);
    if (le_tiny ? (fabs(dt) <= DBL_MIN) : (fabs(dt) < DBL_MIN)) return false;
    inv[0] = a[3] / dt; inv[1] = -a[1] / dt; inv[2] = -a[2] / dt; inv[3] = a[0] / dt;
    return true;
  }
};

__device__ __forceinline__ double sf_weight_sq(double dx, double dy) {
  const double w = 1.0 / pow(sqrt(dx * dx + dy * dy), 1.5);
  return w * w;
}

// regular grid, 1st order: out[0] = d/dx, out[1] = d/dy; Ni = home coefficients, Nc[k*n_c + c] = neighbour coefficients
__device__ inline bool sf_reg_1st(double x, double y, int n_c, const double* xc, const double* yc, double* Ni, double* Nc) {
  double M[4] = {0, 0, 0, 0}, Minv[4];
  for (int c = 0; c < n_c; ++c) {
    const double dx = xc[c] - x, dy = yc[c] - y, w2 = sf_weight_sq(dx, dy);
    M[0] += w2 * dx * dx; M[1] += w2 * dx * dy; M[2] += w2 * dy * dx; M[3] += w2 * dy * dy;
  }
  if (!SmallMat<2>::inverse(M, Minv, false)) return false;
  double sx = 0, sy = 0;
  for (int c = 0; c < n_c; ++c) {
    const double dx = xc[c] - x, dy = yc[c] - y, w2 = sf_weight_sq(dx, dy);
    const double nx = w2 * ((Minv[0] * dx) + (Minv[1] * dy)), ny = w2 * ((Minv[2] * dx) + (Minv[3] * dy));
    Nc[c] = nx; Nc[n_c + c] = ny;
    sx += nx; sy += ny;
  }
  Ni[0] = -sx; Ni[1] = -sy;
  return true;
}

// staggered grid, 1st order: rows map, d/dx, d/dy
__device__ inline bool sf_stag_1st(double x, double y, int n_c, const double* xc, const double* yc, double* Nc) {
  double M[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, Minv[9];
  for (int c = 0; c < n_c; ++c) {
    const double dx = xc[c] - x, dy = yc[c] - y, w2 = sf_weight_sq(dx, dy);
    const double b[3] = {1.0, dx, dy};
    for (int r = 0; r < 3; ++r)
      for (int s = 0; s < 3; ++s) M[r * 3 + s] += (w2 * b[r] * b[s]);
  }
  if (!SmallMat<3>::inverse(M, Minv, true)) return false;
  for (int c = 0; c < n_c; ++c) {
    const double dx = xc[c] - x, dy = yc[c] - y, w2 = sf_weight_sq(dx, dy);
    Nc[c] = w2 * ((Minv[0] * 1.0) + (Minv[1] * dx) + (Minv[2] * dy));
    Nc[n_c + c] = w2 * ((Minv[3] * 1.0) + (Minv[4] * dx) + (Minv[5] * dy));
    Nc[2 * n_c + c] = w2 * ((Minv[6] * 1.0) + (Minv[7] * dx) + (Minv[8] * dy));
  }
  return true;
}

// regular grid, 2nd order: rows d/dx, d/dy, d2/dx2, d2/dxdy, d2/dy2
__device__ inline bool sf_reg_2nd(double x, double y, int n_c, const double* xc, const double* yc, double* Ni, double* Nc) {
  double M[25], Minv[25];
  for (int k = 0; k < 25; ++k) M[k] = 0.0;
  for (int c = 0; c < n_c; ++c) {
    const double dx = xc[c] - x, dy = yc[c] - y, w2 = sf_weight_sq(dx, dy);
    const double b[5] = {dx, dy, dx * dx / 2.0, dx * dy, dy * dy / 2.0};
    for (int r = 0; r < 5; ++r)
      for (int s = 0; s < 5; ++s) M[r * 5 + s] += w2 * b[r] * b[s];
  }
  if (!SmallMat<5>::inverse(M, Minv, false)) return false;
  for (int r = 0; r < 5; ++r) Ni[r] = 0.0;
  for (int c = 0; c < n_c; ++c) {
    const double dx = xc[c] - x, dy = yc[c] - y, w2 = sf_weight_sq(dx, dy);
    const double b[5] = {dx, dy, dx * dx / 2.0, dx * dy, dy * dy / 2.0};
    for (int r = 0; r < 5; ++r) {
      double s = 0.0;
      for (int k = 0; k < 5; ++k) s += Minv[r * 5 + k] * b[k];
      const double v = w2 * s;
      Nc[r * n_c + c] = v;
      Ni[r] -= v;
    }
  }
  return true;
}

}  // namespace ufe
