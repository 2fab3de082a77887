// On-device Krylov solvers (north_star (b); K8/K9 of SURVEY 2e).
// Replaces the PETSc KSP solve behind solve_matrix_equation_CSR_PETSc (src/UPSY/basic/petsc_basic.f90:32-242):
//   * zero initial guess (KSPSetInitialGuessNonzero is commented out, :163)
//   * left preconditioning, convergence on ||B r||_2 <= max(rtol ||B b||_2, abstol) (KSPConvergedDefault)
//   * at most maxit (=2000, :166) iterations; non-convergence is silent (iteration count only)
// Algorithm = PETSc's KSPBCGS recurrence (BiCGStab, "-ksp_type bcgs -pc_type jacobi" through
// KSPSetFromOptions, :193) with every dot product fused into the kernel that produces its operand.
//
// The matrix enters through an "Op" functor so the same solver drives the generic CSR kernels and the
// blocked stiffness-matrix format:
//   struct Op { static constexpr int W;            // lanes cooperating on one row (power of two <= 32)
//               __device__ double row(int r, int lane, const double* x_ext) const; }  // full sum in every lane
#pragma once
#include "common.cuh"
#include "reduce.cuh"

namespace ufe {

// device scalars; two parity banks of 8 doubles + persistent slots
enum { KS_RHO = 0, KS_RR = 1, KS_D1 = 2, KS_ST = 3, KS_TT = 4, KS_SS = 5, KS_SRP = 6, KS_TRP = 7 /* 5-7: merged reduction of mg.cu */, KS_BANK = 8, KS_RR0 = 16, KS_RINIT = 17, KS_RTOL = 18, KS_ABSTOL = 19, KS_COUNT = 24 };
enum { KF_DONE = 0, KF_ITS = 1, KF_REASON = 2, KF_COUNT = 4 };

struct Krylov {
  DBuf<double> R, RP, V, T, dinv;   // local (m_loc)
  DBuf<double> P, S;                // extended (inputs of the SpMV)
  DBuf<double> W1, W2;              // local work vectors of the multigrid-preconditioned variant (mg.cu)
  DBuf<double> binv;                // 2x2 block-Jacobi inverse blocks [4][m_loc/2]
  DBuf<double> scal, partials;
  DBuf<unsigned> counter;
  DBuf<int> flags;
  int grid = 0;
  int ensure(Ctx* c, int m_loc, int n_ext);
};

struct KrylovParams {
  double rtol, abstol;
  int maxit, ksp_type, pc_type;
  int use_guess = 0;   // 1: x_ext holds the initial guess (KSPSetInitialGuessNonzero); the test still uses ||B b||
};

// ---- kernels shared by all matrix formats ------------------------------------------------------------

// R = RP = B (b - A x0) (Ax = nullptr: x0 = 0 and x is zeroed);  RR0 = ||B b||^2, RINIT = RR[0] = RHO[0] = (R,R)
__global__ void k_ks_init(int m, const double* __restrict__ b, const double* __restrict__ Ax, const double* __restrict__ dinv, double* __restrict__ R,
                          double* __restrict__ RP, double* __restrict__ x_own, double* __restrict__ P_own,
                          double* __restrict__ V, double* scal, double* partials, unsigned* counter, int* flags,
                          double rtol, double abstol);
__global__ void k_ks_init_finish(double* scal);

// iteration i, step 1: convergence test on ||R_i||, then P = R + beta (P - omega V)
__global__ void k_ks_update_p(int m, int iter, const double* __restrict__ R, const double* __restrict__ V,
                              double* __restrict__ P_own, const double* __restrict__ scal, int* flags);
// step 3: S = R - alpha V
__global__ void k_ks_update_s(int m, int iter, const double* __restrict__ R, const double* __restrict__ V,
                              double* __restrict__ S_own, const double* __restrict__ scal, const int* flags);
// step 5: x += alpha P + omega S;  R = S - omega T;  RR, RHO of the next iteration
__global__ void k_ks_update_xr(int m, int iter, double* __restrict__ x_own, const double* __restrict__ P_own,
                               const double* __restrict__ S_own, const double* __restrict__ T, double* __restrict__ R,
                               const double* __restrict__ RP, double* scal, double* partials, unsigned* counter,
                               const int* flags);
// final convergence test after the last launched iteration
__global__ void k_ks_final_check(int iter, const double* __restrict__ scal, int* flags);

// ---- SpMV with fused preconditioner and dot products ---------------------------------------------------
// MODE 0: y = A x                      (no preconditioner, no reduction)
// MODE 1: y = B A x, D1  = (y, RP)      (x = P)
// MODE 2: y = B A x, ST  = (x_own, y), TT = (y, y)   (x = S)
template <class Op, int MODE>
__global__ void __launch_bounds__(kBlock) k_spmv_fused(Op op, int m, int off, int iter, const double* __restrict__ x_ext,
                                                       const double* __restrict__ dinv, double* __restrict__ y,
                                                       const double* __restrict__ rp, double* scal, double* partials,
                                                       unsigned* counter, const int* flags) {
  if (MODE != 0 && flags[KF_DONE]) return;
  constexpr int W = Op::W;
  constexpr int RPB = kBlock / W;  // rows per block pass
  const int lane = threadIdx.x % W, rib = threadIdx.x / W;
  double acc[2] = {0.0, 0.0};
  for (long long base = (long long)blockIdx.x * RPB; base < m; base += (long long)gridDim.x * RPB) {
    const int r = (int)base + rib;
    // every lane of the warp takes part in the row's shuffle reduction: clamp the row, guard the store
    double s = op.row(r < m ? r : m - 1, lane, x_ext);
    if (r < m && lane == 0) {
      if (MODE != 0) s *= dinv[r];
      y[r] = s;
      if (MODE == 1) acc[0] += s * rp[r];
      if (MODE == 2) { acc[0] += x_ext[off + r] * s; acc[1] += s * s; }
    }
  }
  if (MODE == 1) { double v[1] = {acc[0]}; grid_sum<1>(v, partials, counter, scal + (iter & 1) * KS_BANK + KS_D1); }
  if (MODE == 2) grid_sum<2>(acc, partials, counter, scal + (iter & 1) * KS_BANK + KS_ST);
}

// generic CSR, W lanes per row, coalesced loads of val/ind along the row (north_star (b): warp-per-row W = 32,
// vector-per-row W < 32)
template <int W_>
struct CsrOp {
  static constexpr int W = W_;
  const int* __restrict__ ptr;
  const int* __restrict__ ind;
  const double* __restrict__ val;
  __device__ __forceinline__ double row(int r, int lane, const double* __restrict__ x) const {
    const int k1 = ptr[r], k2 = ptr[r + 1];
    double s = 0.0;
    for (int k = k1 + lane; k < k2; k += W) s += val[k] * x[ind[k]];
#pragma unroll
    for (int o = W / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o, W);
    return s;
  }
};

// the same matrix as a sliced-ELL copy (DevCsr::sell_*): one row per thread, a warp = one slice of 32 rows whose entries are stored column-major,
// so every val / ind load of the warp is one full 256 B / 128 B segment and each thread keeps several independent loads in flight
// (north_star (b): "coalesced, vectorised loads"; the vector-per-row kernel above has one entry per thread in flight, which is what holds it at
// ~0.4 of the copy bandwidth).  Padding slots hold val = 0 and a column of the same row.
struct SellOp {
  static constexpr int W = 1;
  const int* __restrict__ slice;
  const int* __restrict__ ind;
  const double* __restrict__ val;
  __device__ __forceinline__ double row(int r, int, const double* __restrict__ x) const {
    const int s = r >> 5, b = slice[s], w = (slice[s + 1] - b) >> 5;
    const int* __restrict__ ip = ind + b + (r & 31);
    const double* __restrict__ vp = val + b + (r & 31);
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    int k = 0;
    for (; k + 3 < w; k += 4) {
      const int j0 = ip[k * 32], j1 = ip[(k + 1) * 32], j2 = ip[(k + 2) * 32], j3 = ip[(k + 3) * 32];
      const double v0 = vp[k * 32], v1 = vp[(k + 1) * 32], v2 = vp[(k + 2) * 32], v3 = vp[(k + 3) * 32];
      s0 += v0 * x[j0]; s1 += v1 * x[j1]; s2 += v2 * x[j2]; s3 += v3 * x[j3];
    }
    for (; k < w; ++k) s0 += vp[k * 32] * x[ip[k * 32]];
    return (s0 + s1) + (s2 + s3);
  }
};

// host driver, templated on the Op (defined in krylov_impl.cuh)
template <class Op>
int bicgstab_solve(Ctx* c, const Op& op, const Range& rr, const double* d_b, const double* d_dinv, double* d_x_ext,
                   const KrylovParams& kp, int timing_class_spmv, int* its);

}  // namespace ufe
