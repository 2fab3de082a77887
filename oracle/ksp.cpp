// TEST INFRASTRUCTURE ONLY (see oracle.hpp).
// Krylov solvers behind solve_matrix_equation_CSR_PETSc (src/UPSY/basic/petsc_basic.f90:32-242).
//
// The arithmetic lives in PETSc (not vendored; the golden logs record PETSc 3.24.6,
// automated_testing/UFEMISM/integrated_test_MISMIP_mod_small/reference/checksum_logfile.txt:4).
// Restated from PETSc's published algorithms at the reference's call site:
//   * KSPSetTolerances(rtol, abstol, default dtol, maxit=2000)          petsc_basic.f90:166
//   * zero initial guess (KSPSetInitialGuessNonzero is commented out)   petsc_basic.f90:163
//   * default KSP/PC = GMRES(30) + ILU(0) on one rank (block Jacobi of ILU(0) on several)
//   * default (left-preconditioned) residual test  ||B r||_2 <= max(rtol ||B b||_2, abstol)
//   * KSPSetFromOptions lets "-ksp_type bcgs -pc_type jacobi" replace the method  petsc_basic.f90:193
// ksp_bicgstab_jacobi follows KSPBCGS (van der Vorst BiCGStab, left preconditioning) and is the
// algorithm the CUDA path implements; ksp_gmres_ilu0 is the reference's default configuration.
#include <algorithm>
#include <cmath>
#include <cstring>

#include <omp.h>

#include "oracle.hpp"

namespace orc {

namespace {
double dot(const double* a, const double* b, size_t n) {
  double s = 0.0;
#pragma omp parallel for reduction(+ : s) schedule(static)
  for (size_t i = 0; i < n; ++i) s += a[i] * b[i];
  return s;
}

// diagonal of A; duplicate diagonal entries are summed (MatGetDiagonal on an assembled AIJ matrix)
void get_diagonal(const Csr& A, std::vector<double>& d) {
  d.assign(A.m, 0.0);
#pragma omp parallel for schedule(static)
  for (int i = 1; i <= A.m; ++i) {
    double s = 0.0;
    for (int k = A.ptr[i - 1]; k <= A.ptr[i] - 1; ++k) if (A.ind[k - 1] == i) s += A.val[k - 1];
    d[i - 1] = s;
  }
}
}  // namespace

int ksp_bicgstab_jacobi(const Csr& A, const double* b, double* x, double rtol, double abstol, int maxit, int* its, bool use_guess) {
  const size_t n = A.m;
  std::vector<double> dinv, R(n), RP(n), P(n, 0.0), V(n, 0.0), S(n), T(n), tmp(n);
  get_diagonal(A, dinv);
  for (size_t i = 0; i < n; ++i) dinv[i] = (dinv[i] != 0.0) ? 1.0 / dinv[i] : 1.0;  // PCJacobi: zero diagonal -> 1
  for (size_t i = 0; i < n; ++i) R[i] = dinv[i] * b[i];
  const double bnorm = std::sqrt(dot(R.data(), R.data(), n));
  if (use_guess) {  // KSPSetInitialGuessNonzero: r0 = B (b - A x0); the default test still scales with ||B b||
    spmv(A, x, tmp.data());
    for (size_t i = 0; i < n; ++i) R[i] = dinv[i] * (b[i] - tmp[i]);
  } else {
    for (size_t i = 0; i < n; ++i) x[i] = 0.0;
  }
  double dp = std::sqrt(dot(R.data(), R.data(), n));
  const double ttol = std::max(rtol * bnorm, abstol);  // KSPConvergedDefault: rnorm0 = ||B b||
  *its = 0;
  if (dp <= ttol) return 0;
  RP = R;
  double rhoold = 1.0, alpha = 1.0, omegaold = 1.0;
  for (int i = 0; i < maxit; ++i) {
    double rho = dot(R.data(), RP.data(), n);
    if (rho == 0.0) return 1;  // breakdown
    double beta = (rho / rhoold) * (alpha / omegaold);
#pragma omp parallel for schedule(static)
    for (size_t k = 0; k < n; ++k) P[k] = R[k] + beta * (P[k] - omegaold * V[k]);
    spmv(A, P.data(), tmp.data());
#pragma omp parallel for schedule(static)
    for (size_t k = 0; k < n; ++k) V[k] = dinv[k] * tmp[k];
    double d1 = dot(V.data(), RP.data(), n);
    if (d1 == 0.0) return 1;
    alpha = rho / d1;
#pragma omp parallel for schedule(static)
    for (size_t k = 0; k < n; ++k) S[k] = R[k] - alpha * V[k];
    spmv(A, S.data(), tmp.data());
#pragma omp parallel for schedule(static)
    for (size_t k = 0; k < n; ++k) T[k] = dinv[k] * tmp[k];
    double st = dot(S.data(), T.data(), n), tt = dot(T.data(), T.data(), n);
    if (tt == 0.0) {
      // t == 0: s is (numerically) zero too, x + alpha p is the solution
#pragma omp parallel for schedule(static)
      for (size_t k = 0; k < n; ++k) x[k] += alpha * P[k];
      *its = i + 1;
      return 0;
    }
    double omega = st / tt;
#pragma omp parallel for schedule(static)
    for (size_t k = 0; k < n; ++k) { x[k] += alpha * P[k] + omega * S[k]; R[k] = S[k] - omega * T[k]; }
    dp = std::sqrt(dot(R.data(), R.data(), n));
    rhoold = rho; omegaold = omega;
    *its = i + 1;
    if (dp <= ttol) return 0;
    if (!(dp == dp)) return 2;  // NaN
  }
  return 0;  // not converged within maxit: silent, exactly like the reference (iteration count only)
}

namespace {
// canonical CSR (0-based, sorted columns, duplicates summed) as an assembled AIJ matrix holds it
struct Aij { int n; std::vector<int> ptr, ind, diag; std::vector<double> val; };

// ILU(0), natural ordering, in place (row-wise IKJ); zero pivots are an error, as in PETSc
bool ilu0(Aij& F) {
  const int n = F.n;
  F.diag.assign(n, -1);
  std::vector<int> pos(n, -1);
  for (int i = 0; i < n; ++i) {
    for (int k = F.ptr[i]; k < F.ptr[i + 1]; ++k) pos[F.ind[k]] = k;
    for (int k = F.ptr[i]; k < F.ptr[i + 1]; ++k) {
      int j = F.ind[k];
      if (j >= i) break;
      if (F.diag[j] < 0) return false;
      double lij = F.val[k] / F.val[F.diag[j]];
      F.val[k] = lij;
      for (int kk = F.diag[j] + 1; kk < F.ptr[j + 1]; ++kk) {
        int p = pos[F.ind[kk]];
        if (p >= 0) F.val[p] -= lij * F.val[kk];
      }
    }
    for (int k = F.ptr[i]; k < F.ptr[i + 1]; ++k) { if (F.ind[k] == i) F.diag[i] = k; pos[F.ind[k]] = -1; }
    if (F.diag[i] < 0 || F.val[F.diag[i]] == 0.0) return false;
  }
  return true;
}

void ilu_solve(const Aij& F, const double* r, double* z) {
  const int n = F.n;
  for (int i = 0; i < n; ++i) {
    double s = r[i];
    for (int k = F.ptr[i]; k < F.diag[i]; ++k) s -= F.val[k] * z[F.ind[k]];
    z[i] = s;
  }
  for (int i = n - 1; i >= 0; --i) {
    double s = z[i];
    for (int k = F.diag[i] + 1; k < F.ptr[i + 1]; ++k) s -= F.val[k] * z[F.ind[k]];
    z[i] = s / F.val[F.diag[i]];
  }
}

// PCBJACOBI with one ILU(0) block per rank -- PETSc's default preconditioner on more than one MPI rank: rank r owns the rows partition_list gives it
// (mpi_distributed_memory.f90:47-73; petsc_basic.f90:494 hands PETSc exactly those row blocks) and factors the diagonal block of its rows, i.e. the
// entries whose column lies in the same range; everything outside the block is dropped from the preconditioner.  nblocks = 1 is plain ILU(0).
struct BlockIlu {
  std::vector<Aij> F;
  std::vector<int> r0;   // first row (0-based) of every block, size nblocks + 1
};

bool block_ilu0(const Csr& A, int nblocks, BlockIlu& B) {
  const int n = A.m;
  if (n <= 2 * nblocks) nblocks = 1;            // partition_list: rank 0 owns everything
  B.F.assign(nblocks, Aij()); B.r0.assign(nblocks + 1, 0);
  for (int r = 0; r < nblocks; ++r) {
    const int rem = n % nblocks, slice = n / nblocks;
    B.r0[r] = slice * r + std::min(r, rem);
  }
  B.r0[nblocks] = n;
  int bad = 0;
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : bad)
  for (int r = 0; r < nblocks; ++r) {
    const int lo = B.r0[r], hi = B.r0[r + 1];
    Aij& F = B.F[r];
    F.n = hi - lo; F.ptr.assign(F.n + 1, 0);
    std::vector<std::pair<int, double>> row;
    for (int i = lo; i < hi; ++i) {
      row.clear();
      for (int k = A.ptr[i]; k <= A.ptr[i + 1] - 1; ++k) {
        const int j = A.ind[k - 1] - 1;
        if (j >= lo && j < hi) row.push_back({j - lo, A.val[k - 1]});
      }
      std::stable_sort(row.begin(), row.end(), [](auto& a, auto& b) { return a.first < b.first; });
      for (size_t k = 0; k < row.size(); ++k) {
        if (k > 0 && row[k].first == row[k - 1].first) F.val.back() += row[k].second;
        else { F.ind.push_back(row[k].first); F.val.push_back(row[k].second); }
      }
      F.ptr[i - lo + 1] = (int)F.ind.size();
    }
    if (!ilu0(F)) bad += 1;
  }
  return bad == 0;
}

void block_ilu_solve(const BlockIlu& B, const double* r, double* z) {
  const int nb = (int)B.F.size();
#pragma omp parallel for schedule(dynamic, 1)
  for (int b = 0; b < nb; ++b) ilu_solve(B.F[b], r + B.r0[b], z + B.r0[b]);
}
}  // namespace

int ksp_gmres_ilu0(const Csr& A, const double* b, double* x, double rtol, double abstol, int maxit, int restart, int* its, bool use_guess) {
  return ksp_gmres_bjacobi_ilu0(A, 1, b, x, rtol, abstol, maxit, restart, its, use_guess);
}

static int g_bjacobi_blocks = 0;   // 0: one block per OpenMP thread (= one per MPI rank of `mpiexec -n <cores>`)
void set_bjacobi_blocks(int n) { g_bjacobi_blocks = n; }
int bjacobi_blocks() { return g_bjacobi_blocks > 0 ? g_bjacobi_blocks : omp_get_max_threads(); }

int ksp_solve(int ksp_type, const Csr& A, const double* b, double* x, double rtol, double abstol, int maxit, int* its, bool use_guess) {
  if (ksp_type == KSP_gmres_ilu0) return ksp_gmres_ilu0(A, b, x, rtol, abstol, maxit, 30, its, use_guess);
  if (ksp_type == KSP_gmres_bjacobi_ilu0) return ksp_gmres_bjacobi_ilu0(A, bjacobi_blocks(), b, x, rtol, abstol, maxit, 30, its, use_guess);
  return ksp_bicgstab_jacobi(A, b, x, rtol, abstol, maxit, its, use_guess);
}

int ksp_gmres_bjacobi_ilu0(const Csr& A, int nblocks, const double* b, double* x, double rtol, double abstol, int maxit, int restart, int* its,
                           bool use_guess) {
  const size_t n = A.m;
  BlockIlu F;
  if (!block_ilu0(A, std::max(1, nblocks), F)) return 3;
  const int m = restart;
  std::vector<std::vector<double>> Vk(m + 1, std::vector<double>(n));
  std::vector<double> H((size_t)(m + 1) * m, 0.0), cs(m), sn(m), g(m + 1), w(n), tmp(n), y(m);
  if (!use_guess) for (size_t i = 0; i < n; ++i) x[i] = 0.0;
  *its = 0;
  block_ilu_solve(F, b, w.data());
  double ttol = std::max(rtol * std::sqrt(dot(w.data(), w.data(), n)), abstol);
  int total = 0;
  while (true) {
    // preconditioned residual z = B (b - A x)
    spmv(A, x, tmp.data());
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < n; ++i) tmp[i] = b[i] - tmp[i];
    block_ilu_solve(F, tmp.data(), Vk[0].data());
    double beta = std::sqrt(dot(Vk[0].data(), Vk[0].data(), n));
    if (beta <= ttol || total >= maxit) { *its = total; return 0; }
    { double* v0 = Vk[0].data();
#pragma omp parallel for schedule(static)
      for (size_t i = 0; i < n; ++i) v0[i] /= beta; }
    std::fill(g.begin(), g.end(), 0.0);
    g[0] = beta;
    int k = 0;
    bool done = false;
    for (; k < m && total < maxit; ++k) {
      spmv(A, Vk[k].data(), tmp.data());
      block_ilu_solve(F, tmp.data(), w.data());
      for (int j = 0; j <= k; ++j) {  // modified Gram-Schmidt
        const double* vj = Vk[j].data();
        double h = dot(w.data(), vj, n);
        H[(size_t)j * m + k] = h;
        double* ww = w.data();
#pragma omp parallel for schedule(static)
        for (size_t i = 0; i < n; ++i) ww[i] -= h * vj[i];
      }
      double hn = std::sqrt(dot(w.data(), w.data(), n));
      H[(size_t)(k + 1) * m + k] = hn;
      if (hn != 0.0) {
        double* vn = Vk[k + 1].data(); const double* ww = w.data();
#pragma omp parallel for schedule(static)
        for (size_t i = 0; i < n; ++i) vn[i] = ww[i] / hn;
      }
      for (int j = 0; j < k; ++j) {  // apply previous Givens rotations
        double t = cs[j] * H[(size_t)j * m + k] + sn[j] * H[(size_t)(j + 1) * m + k];
        H[(size_t)(j + 1) * m + k] = -sn[j] * H[(size_t)j * m + k] + cs[j] * H[(size_t)(j + 1) * m + k];
        H[(size_t)j * m + k] = t;
      }
      double a = H[(size_t)k * m + k], bb = H[(size_t)(k + 1) * m + k], r = std::hypot(a, bb);
      cs[k] = (r == 0.0) ? 1.0 : a / r; sn[k] = (r == 0.0) ? 0.0 : bb / r;
      H[(size_t)k * m + k] = r; H[(size_t)(k + 1) * m + k] = 0.0;
      g[k + 1] = -sn[k] * g[k]; g[k] = cs[k] * g[k];
      ++total;
      if (std::fabs(g[k + 1]) <= ttol || hn == 0.0) { ++k; done = true; break; }
    }
    // back substitution and update
    for (int i = k - 1; i >= 0; --i) {
      double s = g[i];
      for (int j = i + 1; j < k; ++j) s -= H[(size_t)i * m + j] * y[j];
      y[i] = s / H[(size_t)i * m + i];
    }
    for (int j = 0; j < k; ++j) {
      const double yj = y[j]; const double* vj = Vk[j].data();
#pragma omp parallel for schedule(static)
      for (size_t i = 0; i < n; ++i) x[i] += yj * vj[i];
    }
    if (done || total >= maxit) { *its = total; return 0; }
  }
}

}  // namespace orc
