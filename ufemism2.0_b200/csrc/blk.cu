// Construction of the blocked sliced-ELL stiffness matrix from its CSR pattern (see blk.cuh).
#include "blk.cuh"

#include <cub/device/device_scan.cuh>

namespace ufe {

namespace {

constexpr int kMaxBlk = 48;  // most distinct neighbour triangles one triangle's two rows may reference

// distinct neighbour triangles of the two rows of triangle t, in order of first appearance
__device__ __forceinline__ int collect_cols(const int* __restrict__ ptr, const int* __restrict__ ind, int t, int* list, bool* overflow) {
  int n = 0;
  for (int row = 2 * t; row < 2 * t + 2; ++row)
    for (int k = ptr[row]; k < ptr[row + 1]; ++k) {
      const int tj = ind[k] >> 1;
      bool seen = false;
      for (int i = 0; i < n; ++i) if (list[i] == tj) { seen = true; break; }
      if (!seen) {
        if (n >= kMaxBlk) { *overflow = true; return n; }
        list[n++] = tj;
      }
    }
  return n;
}

// one warp per slice: width = longest row of the slice
__global__ void k_blk_count(int n_tri, int n_slices, const int* __restrict__ ptr, const int* __restrict__ ind, int* __restrict__ width, int* err) {
  const int lane = threadIdx.x & 31;
  int list[kMaxBlk];
  for (int s = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); s < n_slices; s += gridDim.x * (blockDim.x >> 5)) {
    const int t = (s << 5) + lane;
    bool ovf = false;
    int n = (t < n_tri) ? collect_cols(ptr, ind, t, list, &ovf) : 0;
    if (ovf) atomicExch(err, t + 1);
    for (int o = 16; o > 0; o >>= 1) n = max(n, __shfl_xor_sync(0xffffffffu, n, o));
    if (lane == 0) width[s] = n;
  }
}

__global__ void k_blk_fill(int n_tri, int n_slices, int off_tri, const int* __restrict__ ptr, const int* __restrict__ ind,
                           const int* __restrict__ slice_ptr, int* __restrict__ col, int* __restrict__ map) {
  const int lane = threadIdx.x & 31;
  int list[kMaxBlk];
  for (int s = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); s < n_slices; s += gridDim.x * (blockDim.x >> 5)) {
    const int t = (s << 5) + lane;
    const int k0 = slice_ptr[s], k1 = slice_ptr[s + 1];
    bool ovf = false;
    const int n = (t < n_tri) ? collect_cols(ptr, ind, t, list, &ovf) : 0;
    // padding points at a valid x entry (the own triangle, or triangle 0 of the extended vector for the tail lanes)
    const int pad = (t < n_tri) ? off_tri + t : 0;
    for (int k = k0; k < k1; ++k) col[(size_t)k * 32 + lane] = (k - k0 < n) ? list[k - k0] : pad;
    if (t < n_tri)
      for (int uvr = 0; uvr < 2; ++uvr) {
        const int row = 2 * t + uvr;
        for (int k = ptr[row]; k < ptr[row + 1]; ++k) {
          const int c = ind[k], tj = c >> 1, uvc = c & 1;
          int j = 0;
          while (list[j] != tj) ++j;
          map[k] = (int)((((size_t)(k0 + j)) * 4 + uvr * 2 + uvc) * 32 + lane);
        }
      }
  }
}

__global__ void k_blk_scatter(long long nnz, const int* __restrict__ map, const double* __restrict__ v, double* __restrict__ out) {
  for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += (long long)gridDim.x * blockDim.x) out[map[k]] = v[k];
}

}  // namespace

int blk_build(Ctx* c, BlkMat& B, const DevCsr& A, int n_tri, int off_tri) {
  UFE_CHECK(A.m_loc == 2 * n_tri, "blk_build: the matrix must have two rows per triangle");
  UFE_CHECK((A.cols.lo & 1) == 1, "blk_build: extended column range must start at a u-unknown");
  B.valid = false;
  B.n_tri = n_tri; B.n_slices = (n_tri + 31) / 32;
  UFE_TRY(B.slice_ptr.alloc((size_t)B.n_slices + 1));
  UFE_CUDA(cudaMemsetAsync(B.slice_ptr.p, 0, sizeof(int) * ((size_t)B.n_slices + 1), c->stream));
  DBuf<int> err;
  UFE_TRY(err.alloc(1));
  UFE_CUDA(cudaMemsetAsync(err.p, 0, sizeof(int), c->stream));
  const int grid = grid_for(c, (long long)B.n_slices * 32, 128, 16);
  { UFE_LAUNCH(c, TC_ASSEMBLY); k_blk_count<<<grid, 128, 0, c->stream>>>(n_tri, B.n_slices, A.ptr, A.ind, B.slice_ptr.p, err.p); }
  size_t tmp_bytes = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, B.slice_ptr.p, B.slice_ptr.p, B.n_slices + 1, c->stream);
  DBuf<char> tmp;
  UFE_TRY(tmp.alloc(tmp_bytes));
  c->launches++;
  UFE_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, tmp_bytes, B.slice_ptr.p, B.slice_ptr.p, B.n_slices + 1, c->stream));
  int h[2] = {0, 0};
  UFE_CUDA(cudaMemcpyAsync(&h[0], B.slice_ptr.p + B.n_slices, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  UFE_CUDA(cudaMemcpyAsync(&h[1], err.p, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  UFE_CHECK(h[1] == 0, "blk_build: triangle %d references more than %d neighbour triangles", h[1], kMaxBlk);
  B.n_slots = (long long)h[0] * 32;
  UFE_CHECK(B.n_slots * 4 < 2147483647LL, "blk_build: matrix too large for 32-bit value offsets (%lld block slots)", B.n_slots);
  UFE_TRY(B.col.alloc(B.n_slots)); UFE_TRY(B.val.alloc(B.n_slots * 4)); UFE_TRY(B.map.alloc(A.nnz));
  UFE_CUDA(cudaMemsetAsync(B.val.p, 0, sizeof(double) * B.n_slots * 4, c->stream));
  { UFE_LAUNCH(c, TC_ASSEMBLY); k_blk_fill<<<grid, 128, 0, c->stream>>>(n_tri, B.n_slices, off_tri, A.ptr, A.ind, B.slice_ptr.p, B.col.p, B.map.p); }
  UFE_CUDA(cudaGetLastError());
  B.valid = true;
  return 0;
}

int blk_update_values(Ctx* c, BlkMat& B, const DevCsr& A) {
  UFE_LAUNCH(c, TC_ASSEMBLY);
  k_blk_scatter<<<grid_for(c, A.nnz, 256), 256, 0, c->stream>>>(A.nnz, B.map.p, A.val, B.val.p);
  UFE_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace ufe
