// Deterministic grid-wide sum reductions fused into the producing kernel (K8 of SURVEY 2e):
// every block writes its partial sums, the last block to arrive (ticket counter) adds the partials in a fixed
// order and stores the totals.  For a fixed launch configuration the result is bit-reproducible.
#pragma once
#include <cuda_runtime.h>

namespace ufe {

constexpr int kBlock = 256;

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Sum N values over the whole grid. partials: [N][gridDim.x] doubles; counter: one unsigned (zero before first use);
// out[k] receives the totals (written by the last block only).  All threads of all blocks must call this.
template <int N>
__device__ __forceinline__ void grid_sum(double (&v)[N], double* __restrict__ partials, unsigned* __restrict__ counter,
                                         double* __restrict__ out) {
  __shared__ double s_w[N][kBlock / 32];
  __shared__ bool s_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < N; ++k) {
    double w = warp_sum(v[k]);
    if (lane == 0) s_w[k][warp] = w;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < N; ++k) {
      double s = 0.0;
      for (int w = 0; w < kBlock / 32; ++w) s += s_w[k][w];
      partials[(size_t)k * gridDim.x + blockIdx.x] = s;
    }
    __threadfence();
    const unsigned ticket = atomicAdd(counter, 1u);
    s_last = (ticket == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
#pragma unroll
  for (int k = 0; k < N; ++k) {
    double s = 0.0;
    for (unsigned b = threadIdx.x; b < gridDim.x; b += kBlock) s += __ldcg(&partials[(size_t)k * gridDim.x + b]);
    double w = warp_sum(s);
    __syncthreads();
    if (lane == 0) s_w[k][warp] = w;
    __syncthreads();
    if (threadIdx.x == 0) {
      double t = 0.0;
      for (int w2 = 0; w2 < kBlock / 32; ++w2) t += s_w[k][w2];
      out[k] = t;
    }
  }
  if (threadIdx.x == 0) *counter = 0u;
}

}  // namespace ufe
