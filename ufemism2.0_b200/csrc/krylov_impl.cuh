// Host driver of the on-device BiCGStab, templated on the matrix functor (see krylov.cuh).
#pragma once
#include "krylov.cuh"

namespace ufe {

// d_b: local right-hand side (m_loc); d_dinv: local inverse diagonal (Jacobi) or ones;
// d_x_ext: extended solution vector, owned part at [rr.off()..]; halos are NOT refreshed on return.
// launcher for matrices given as a per-row functor (krylov.cuh: CsrOp<W>)
template <class Op>
struct RowOpMat {
  Op op;
  void spmv(Ctx* c, int mode, int m, int off, int iter, const double* x_ext, const double* dinv, double* y, const double* rp, double* scal,
            double* partials, unsigned* counter, const int* flags, int tc) const {
    const int gs = grid_for(c, (long long)m * Op::W, kBlock);
    UFE_LAUNCH(c, tc);
    if (mode == 0) k_spmv_fused<Op, 0><<<gs, kBlock, 0, c->stream>>>(op, m, off, iter, x_ext, dinv, y, rp, scal, partials, counter, flags);
    else if (mode == 1) k_spmv_fused<Op, 1><<<gs, kBlock, 0, c->stream>>>(op, m, off, iter, x_ext, dinv, y, rp, scal, partials, counter, flags);
    else k_spmv_fused<Op, 2><<<gs, kBlock, 0, c->stream>>>(op, m, off, iter, x_ext, dinv, y, rp, scal, partials, counter, flags);
  }
};

template <class Mat>
int bicgstab_solve_mat(Ctx* c, const Mat& A, const Range& rr, const double* d_b, const double* d_dinv, double* d_x_ext,
                       const KrylovParams& kp, int tc_spmv, int* its);

template <class Op>
int bicgstab_solve(Ctx* c, const Op& op, const Range& rr, const double* d_b, const double* d_dinv, double* d_x_ext,
                   const KrylovParams& kp, int tc_spmv, int* its) {
  return bicgstab_solve_mat(c, RowOpMat<Op>{op}, rr, d_b, d_dinv, d_x_ext, kp, tc_spmv, its);
}

template <class Mat>
int bicgstab_solve_mat(Ctx* c, const Mat& A, const Range& rr, const double* d_b, const double* d_dinv, double* d_x_ext,
                       const KrylovParams& kp, int tc_spmv, int* its) {
  if (!c->krylov) c->krylov = new Krylov();
  Krylov& K = *c->krylov;
  const int m = rr.n_loc(), off = rr.off();
  UFE_TRY(K.ensure(c, m, rr.n_ext()));
  cudaStream_t st = c->stream;
  const int gv = grid_for(c, m, kBlock);                       // vector kernels
  double* x_own = d_x_ext + off;
  double* P_own = K.P.p + off;
  double* S_own = K.S.p + off;
  if (rr.n_ext() > m) {  // halos of P/S start from zero so that a missing neighbour never injects garbage
    UFE_CUDA(cudaMemsetAsync(K.P.p, 0, sizeof(double) * rr.n_ext(), st));
    UFE_CUDA(cudaMemsetAsync(K.S.p, 0, sizeof(double) * rr.n_ext(), st));
  }
  // A breakdown of the recurrence leaves x at the last finite iterate (k_ks_update_xr); the solver then restarts from it, at most three
  // times.  A solve that still breaks down returns silently with that iterate and its iteration count, like non-convergence: the
  // reference never looks at PETSc's converged reason (petsc_basic.f90:237-242).
  int* h_flags = reinterpret_cast<int*>(c->h_pinned);
  int total_its = 0, use_guess = kp.use_guess;
  for (int attempt = 0;; ++attempt) {
    const int maxit = kp.maxit - total_its;
    const double* d_Ax = nullptr;
    if (use_guess) {  // T = A x0 (plain product), then R = B (b - T)
      UFE_TRY(halo_exchange(c, rr, d_x_ext));
      A.spmv(c, 0, m, off, 0, d_x_ext, nullptr, K.T.p, nullptr, nullptr, nullptr, nullptr, nullptr, tc_spmv);
      d_Ax = K.T.p;
    }
    {
      UFE_LAUNCH(c, TC_KRYLOV_VEC);
      k_ks_init<<<gv, kBlock, 0, st>>>(m, d_b, d_Ax, d_dinv, K.R.p, K.RP.p, x_own, P_own, K.V.p, K.scal.p, K.partials.p, K.counter.p,
                                       K.flags.p, kp.rtol, kp.abstol);
    }
    UFE_TRY(allreduce_sum(c, K.scal.p + KS_RR0, 2));
    {
      UFE_LAUNCH(c, TC_KRYLOV_VEC);
      k_ks_init_finish<<<1, 1, 0, st>>>(K.scal.p);
    }
    int iter = 0;
    int check_every = 8;
    bool done = false;
    while (!done && iter < maxit) {
      const int stop_at = (iter + check_every < maxit) ? iter + check_every : maxit;
      for (; iter < stop_at; ++iter) {
        const int p = iter & 1;
        {
          UFE_LAUNCH(c, TC_KRYLOV_VEC);
          k_ks_update_p<<<gv, kBlock, 0, st>>>(m, iter, K.R.p, K.V.p, P_own, K.scal.p, K.flags.p);
        }
        UFE_TRY(halo_exchange(c, rr, K.P.p));
        A.spmv(c, 1, m, off, iter, K.P.p, d_dinv, K.V.p, K.RP.p, K.scal.p, K.partials.p, K.counter.p + 1, K.flags.p, tc_spmv);
        UFE_TRY(allreduce_sum(c, K.scal.p + p * KS_BANK + KS_D1, 1));
        {
          UFE_LAUNCH(c, TC_KRYLOV_VEC);
          k_ks_update_s<<<gv, kBlock, 0, st>>>(m, iter, K.R.p, K.V.p, S_own, K.scal.p, K.flags.p);
        }
        UFE_TRY(halo_exchange(c, rr, K.S.p));
        A.spmv(c, 2, m, off, iter, K.S.p, d_dinv, K.T.p, nullptr, K.scal.p, K.partials.p, K.counter.p + 2, K.flags.p, tc_spmv);
        UFE_TRY(allreduce_sum(c, K.scal.p + p * KS_BANK + KS_ST, 2));
        {
          UFE_LAUNCH(c, TC_KRYLOV_VEC);
          k_ks_update_xr<<<gv, kBlock, 0, st>>>(m, iter, x_own, P_own, S_own, K.T.p, K.R.p, K.RP.p, K.scal.p, K.partials.p,
                                                K.counter.p + 3, K.flags.p);
        }
        UFE_TRY(allreduce_sum(c, K.scal.p + (1 - p) * KS_BANK + KS_RHO, 2));
      }
      if (iter >= maxit) {
        UFE_LAUNCH(c, TC_KRYLOV_VEC);
        k_ks_final_check<<<1, 1, 0, st>>>(iter, K.scal.p, K.flags.p);
      } else {
        // the convergence test of iteration `iter` runs inside its update_p; peek at it without advancing
        UFE_LAUNCH(c, TC_KRYLOV_VEC);
        k_ks_update_p<<<1, 32, 0, st>>>(0, iter, K.R.p, K.V.p, P_own, K.scal.p, K.flags.p);
      }
      UFE_CUDA(cudaMemcpyAsync(h_flags, K.flags.p, KF_COUNT * sizeof(int), cudaMemcpyDeviceToHost, st));
      UFE_CUDA(cudaStreamSynchronize(st));
      done = h_flags[KF_DONE] != 0;
      if (check_every < 32) check_every *= 2;
    }
    UFE_CUDA(cudaGetLastError());
    total_its += h_flags[KF_ITS];
    if (h_flags[KF_REASON] != 1 || attempt == 3 || total_its >= kp.maxit) break;
    use_guess = 1;
  }
  *its = total_its;
  return 0;
}

}  // namespace ufe
