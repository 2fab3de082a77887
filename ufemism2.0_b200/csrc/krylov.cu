// Non-templated kernels of the on-device BiCGStab (see krylov.cuh for the algorithm and its reference).
#include "krylov.cuh"
#include "krylov_impl.cuh"

namespace ufe {

int Krylov::ensure(Ctx* c, int m_loc, int n_ext) {
  UFE_TRY(R.alloc(m_loc)); UFE_TRY(RP.alloc(m_loc)); UFE_TRY(V.alloc(m_loc)); UFE_TRY(T.alloc(m_loc));
  UFE_TRY(P.alloc(n_ext)); UFE_TRY(S.alloc(n_ext));
  grid = c->sm_count * 8;
  UFE_TRY(partials.alloc((size_t)8 * grid));
  if (!scal.p) {
    UFE_TRY(scal.alloc(KS_COUNT)); UFE_TRY(counter.alloc(4)); UFE_TRY(flags.alloc(KF_COUNT));
    UFE_CUDA(cudaMemsetAsync(counter.p, 0, 4 * sizeof(unsigned), c->stream));
  }
  return 0;
}

__global__ void k_ks_init(int m, const double* __restrict__ b, const double* __restrict__ Ax, const double* __restrict__ dinv, double* __restrict__ R,
                          double* __restrict__ RP, double* __restrict__ x_own, double* __restrict__ P_own,
                          double* __restrict__ V, double* scal, double* partials, unsigned* counter, int* flags,
                          double rtol, double abstol) {
  double acc[2] = {0.0, 0.0};
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (long long)gridDim.x * blockDim.x) {
    const double pb = dinv[i] * b[i];
    double r = pb;
    if (Ax) r = dinv[i] * (b[i] - Ax[i]); else x_own[i] = 0.0;
    R[i] = r; RP[i] = r; P_own[i] = 0.0; V[i] = 0.0;
    acc[0] += pb * pb;
    acc[1] += r * r;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    flags[KF_DONE] = 0; flags[KF_ITS] = 0; flags[KF_REASON] = 0;
    scal[KS_RTOL] = rtol; scal[KS_ABSTOL] = abstol;
  }
  grid_sum<2>(acc, partials, counter, scal + KS_RR0);
}

// after the (optional) all-reduce of RR0: seed bank 0
__global__ void k_ks_init_finish(double* scal) {
  scal[KS_RHO] = scal[KS_RINIT];
  scal[KS_RR] = scal[KS_RINIT];
}

__device__ __forceinline__ double ks_ttol(const double* scal) {
  return fmax(scal[KS_RTOL] * sqrt(scal[KS_RR0]), scal[KS_ABSTOL]);
}

__global__ void k_ks_update_p(int m, int iter, const double* __restrict__ R, const double* __restrict__ V,
                              double* __restrict__ P_own, const double* __restrict__ scal, int* flags) {
  if (flags[KF_DONE]) return;
  const int p = iter & 1;
  const double* cur = scal + p * KS_BANK;
  const double* prv = scal + (1 - p) * KS_BANK;
  const double dp = sqrt(cur[KS_RR]);
  const bool stop = (dp <= ks_ttol(scal)) || !(dp == dp);
  __syncthreads();
  if (stop) {
    // every block takes the same decision from the same (all-reduced) scalars; one thread records it.
    // The flag is only read by *later* kernels, so there is no intra-kernel race.
    if (blockIdx.x == 0 && threadIdx.x == 0) { flags[KF_ITS] = iter; flags[KF_REASON] = (dp == dp) ? 0 : 1; flags[KF_DONE] = 1; }
    return;
  }
  double beta = 0.0, omega = 0.0;
  if (iter > 0) {
    const double alpha_old = prv[KS_RHO] / prv[KS_D1];
    omega = (prv[KS_TT] == 0.0) ? 0.0 : prv[KS_ST] / prv[KS_TT];
    beta = (cur[KS_RHO] / prv[KS_RHO]) * (alpha_old / omega);
  }
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (long long)gridDim.x * blockDim.x)
    P_own[i] = (iter == 0) ? R[i] : R[i] + beta * (P_own[i] - omega * V[i]);
}

__global__ void k_ks_update_s(int m, int iter, const double* __restrict__ R, const double* __restrict__ V,
                              double* __restrict__ S_own, const double* __restrict__ scal, const int* flags) {
  if (flags[KF_DONE]) return;
  const double* cur = scal + (iter & 1) * KS_BANK;
  const double alpha = cur[KS_RHO] / cur[KS_D1];
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (long long)gridDim.x * blockDim.x)
    S_own[i] = R[i] - alpha * V[i];
}

__global__ void k_ks_update_xr(int m, int iter, double* __restrict__ x_own, const double* __restrict__ P_own,
                               const double* __restrict__ S_own, const double* __restrict__ T, double* __restrict__ R,
                               const double* __restrict__ RP, double* scal, double* partials, unsigned* counter,
                               const int* flags) {
  if (flags[KF_DONE]) return;
  const int p = iter & 1;
  const double* cur = scal + p * KS_BANK;
  const double alpha = cur[KS_RHO] / cur[KS_D1];
  const double omega = (cur[KS_TT] == 0.0) ? 0.0 : cur[KS_ST] / cur[KS_TT];
  // breakdown (rho, (v, r~) or (t, t) vanished): leave x at the last finite iterate and hand a NaN residual norm to the next
  // convergence test, which stops the iteration with reason 1; the host restarts from x (krylov_impl.cuh)
  const bool bad = !(isfinite(alpha) && isfinite(omega));
  double acc[2] = {0.0, 0.0};  // RHO, RR of the next iteration (adjacent slots)
  if (!bad) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (long long)gridDim.x * blockDim.x) {
      const double s = S_own[i];
      x_own[i] += alpha * P_own[i] + omega * s;
      const double r = s - omega * T[i];
      R[i] = r;
      acc[0] += r * RP[i];
      acc[1] += r * r;
    }
  } else {
    acc[0] = acc[1] = nan("");
  }
  grid_sum<2>(acc, partials, counter, scal + (1 - p) * KS_BANK + KS_RHO);
}

__global__ void k_ks_final_check(int iter, const double* __restrict__ scal, int* flags) {
  if (flags[KF_DONE]) return;
  const double dp = sqrt(scal[(iter & 1) * KS_BANK + KS_RR]);
  flags[KF_ITS] = iter;
  flags[KF_REASON] = (dp <= ks_ttol(scal)) ? 0 : ((dp == dp) ? 2 : 1);  // 2 = hit maxit (silent in the reference)
  flags[KF_DONE] = 1;
}

}  // namespace ufe
