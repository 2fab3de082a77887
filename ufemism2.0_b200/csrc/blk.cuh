// Blocked sliced-ELL storage of the SSA/DIVA stiffness matrix (internal format of the hot SpMV).
//
// The two matrix rows of a triangle (u and v equation) reference the same set of neighbour triangles, so the matrix
// is a sparse matrix of 2x2 blocks: one int32 column (the neighbour triangle) + 4 doubles per block = 36 B per 4
// non-zeros instead of 48 B in CSR.  Triangles are grouped in slices of 32 (one warp); inside a slice the k-th block of
// all 32 triangles is stored contiguously, so every load of a warp is one fully coalesced 128 B / 256 B transaction and
// one thread owns one triangle (no shuffles, no idle lanes).  Slices are padded to their longest row; x-sorted meshes
// give nearly uniform row lengths (padding is a few %).
//
// The CSR matrix (bit-exact reference pattern) stays the assembly target and the thing parity tests download; a
// precomputed map scatters its values into the blocks once per viscosity iteration (~1 ms at 4 M vertices against
// hundreds of SpMVs).
#pragma once
#include "common.cuh"
#include "krylov.cuh"

namespace ufe {

struct BlkMat {
  int n_tri = 0, n_slices = 0;
  long long n_slots = 0;        // sum over slices of width * 32
  DBuf<int> slice_ptr;          // [n_slices + 1] offsets in k-steps
  DBuf<int> col;                // [n_slots]   neighbour triangle (extended-local); padding points at the own triangle
  DBuf<double> val;             // [n_slots * 4] planes q = 0..3 (uu, uv, vu, vv) per k-step: ((base + k) * 4 + q) * 32 + lane
  DBuf<int> map;                // [nnz of the CSR matrix] -> position in val
  bool valid = false;
};

// MODE as in k_spmv_fused: 0 plain, 1 (y = B A x, D1 = (y, RP)), 2 (y = B A x, ST = (x, y), TT = (y, y))
template <int MODE>
__global__ void __launch_bounds__(kBlock) k_spmv_blk(int n_tri, int off_tri, int iter, const int* __restrict__ slice_ptr,
                                                     const int* __restrict__ col, const double* __restrict__ val,
                                                     const double* __restrict__ x_ext, const double* __restrict__ dinv,
                                                     double* __restrict__ y, const double* __restrict__ rp, double* scal, double* partials,
                                                     unsigned* counter, const int* flags) {
  if (MODE != 0 && flags[KF_DONE]) return;
  const int lane = threadIdx.x & 31;
  const int warps_per_block = kBlock / 32;
  const int n_slices = (n_tri + 31) >> 5;
  double acc[2] = {0.0, 0.0};
  const double2* __restrict__ x2 = reinterpret_cast<const double2*>(x_ext);
  for (int s = blockIdx.x * warps_per_block + (threadIdx.x >> 5); s < n_slices; s += gridDim.x * warps_per_block) {
    const int t = (s << 5) + lane;
    const int k0 = slice_ptr[s], k1 = slice_ptr[s + 1];
    double yu = 0.0, yv = 0.0;
    const int* __restrict__ cp = col + (size_t)k0 * 32 + lane;
    const double* __restrict__ vp = val + (size_t)k0 * 128 + lane;
#pragma unroll 2
    for (int k = k0; k < k1; ++k, cp += 32, vp += 128) {
      const int tj = *cp;
      const double a0 = vp[0], a1 = vp[32], a2 = vp[64], a3 = vp[96];
      const double2 xv = x2[tj];
      yu += a0 * xv.x + a1 * xv.y;
      yv += a2 * xv.x + a3 * xv.y;
    }
    if (t < n_tri) {
      if (MODE != 0) { yu *= dinv[2 * t]; yv *= dinv[2 * t + 1]; }
      reinterpret_cast<double2*>(y)[t] = make_double2(yu, yv);
      if (MODE == 1) acc[0] += yu * rp[2 * t] + yv * rp[2 * t + 1];
      if (MODE == 2) {
        const double2 xo = x2[off_tri + t];
        acc[0] += xo.x * yu + xo.y * yv;
        acc[1] += yu * yu + yv * yv;
      }
    }
  }
  if (MODE == 1) { double v[1] = {acc[0]}; grid_sum<1>(v, partials, counter, scal + (iter & 1) * KS_BANK + KS_D1); }
  if (MODE == 2) grid_sum<2>(acc, partials, counter, scal + (iter & 1) * KS_BANK + KS_ST);
}

struct BlkLauncher {
  const BlkMat* B;
  int off_tri;   // position of the first owned triangle in extended b-vectors
  void spmv(Ctx* c, int mode, int m, int off, int iter, const double* x_ext, const double* dinv, double* y, const double* rp, double* scal,
            double* partials, unsigned* counter, const int* flags, int tc) const {
    (void)m; (void)off;
    const int grid = grid_for(c, (long long)B->n_slices * 32, kBlock);
    UFE_LAUNCH(c, tc);
    if (mode == 0) k_spmv_blk<0><<<grid, kBlock, 0, c->stream>>>(B->n_tri, off_tri, iter, B->slice_ptr.p, B->col.p, B->val.p, x_ext, dinv, y, rp, scal, partials, counter, flags);
    else if (mode == 1) k_spmv_blk<1><<<grid, kBlock, 0, c->stream>>>(B->n_tri, off_tri, iter, B->slice_ptr.p, B->col.p, B->val.p, x_ext, dinv, y, rp, scal, partials, counter, flags);
    else k_spmv_blk<2><<<grid, kBlock, 0, c->stream>>>(B->n_tri, off_tri, iter, B->slice_ptr.p, B->col.p, B->val.p, x_ext, dinv, y, rp, scal, partials, counter, flags);
  }
};

// build the block pattern + scatter map from a CSR matrix whose rows come in (u, v) pairs per triangle
int blk_build(Ctx* c, BlkMat& B, const DevCsr& A, int n_tri, int off_tri);
// val_blk[map[k]] = val_csr[k]
int blk_update_values(Ctx* c, BlkMat& B, const DevCsr& A);

}  // namespace ufe
