// Generic device CSR: upload from the reference's layout, SpMV (vector-per-row / warp-per-row), Jacobi, BiCGStab.
#pragma once
#include "common.cuh"
#include "krylov.cuh"

namespace ufe {

int setup_range_halos(Ctx* c, Range& r);
int csr_upload(Ctx* c, DevCsr& A, int m_glob, int n_glob, int m_loc, int i1, int n_loc, int j1, const int32_t* ptr,
               const int32_t* ind, const double* val, bool square, const Range* force_cols = nullptr);
int csr_spmv(Ctx* c, const DevCsr& A, const double* x_ext, double* y, int timing_class);
int csr_spmv_multicol(Ctx* c, const DevCsr& A, const double* x_ext, double* y, int ncols, int timing_class);
// sliced-ELL copy of A (see DevCsr): pattern + values / values only (after A.val changed); csr_spmv* and csr_ksp_solve use it when present
int csr_sell_build(Ctx* c, DevCsr& A);
int csr_sell_refresh(Ctx* c, DevCsr& A);
int csr_make_dinv(Ctx* c, const DevCsr& A, double* dinv, int pc_type);
int csr_ksp_solve(Ctx* c, DevCsr& A, const double* d_b, double* d_x_ext, const KrylovParams& kp, int timing_class, int* its);

}  // namespace ufe
