// Narrow seam: generic distributed CSR on the device (a1-a4, a14 of SURVEY 8a).
//   type_sparse_matrix_CSR_dp                      src/UPSY/basic/CSR_sparse_matrix_type.f90:15-38
//   multiply_CSR_matrix_with_vector_{1D,2D}         src/UPSY/basic/CSR_matrix_algebra/CSR_matrix_vector_multiplication.f90:199-365
//   solve_matrix_equation_CSR_PETSc                 src/UPSY/basic/petsc_basic.f90:32-215
#include "nccl_dl.cuh"

#include <algorithm>

#include "csr.cuh"
#include "krylov_impl.cuh"

namespace ufe {

// Exchange (lo, hi, i1, i2) with the neighbours and derive what they need from me.
int setup_range_halos(Ctx* c, Range& r) {
  r.send_l = r.send_r = 0;
  if (c->nranks == 1) return 0;
  DBuf<int> d_all;
  UFE_TRY(d_all.alloc((size_t)4 * c->nranks));
  int mine[4] = {r.lo, r.hi, r.i1, r.i2};
  UFE_CUDA(cudaMemcpyAsync(d_all.p + 4 * c->rank, mine, sizeof(mine), cudaMemcpyHostToDevice, c->stream));
  NcclApi* N = nccl_api();
  if (!N) return 3;
  ncclResult_t rc = N->AllGather(d_all.p + 4 * c->rank, d_all.p, 4, ncclInt32, (ncclComm_t)c->nccl, c->stream);
  UFE_CHECK(rc == ncclSuccess, "ncclAllGather: %s", N->GetErrorString(rc));
  std::vector<int> all((size_t)4 * c->nranks);
  UFE_CUDA(cudaMemcpyAsync(all.data(), d_all.p, sizeof(int) * all.size(), cudaMemcpyDeviceToHost, c->stream));
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  if (c->rank > 0) {
    const int* L = &all[4 * (c->rank - 1)];
    UFE_CHECK(r.lo >= L[2], "halo of rank %d reaches beyond its left neighbour (needs %d, neighbour owns from %d): too many ranks for this mesh",
              c->rank, r.lo, L[2]);
    r.send_l = std::max(0, L[1] - r.i1 + 1);
    UFE_CHECK(r.send_l <= r.n_loc(), "left neighbour's halo is wider than the owned range of rank %d", c->rank);
  } else {
    UFE_CHECK(r.lo == r.i1 || r.lo >= 1, "bad extended range");
  }
  if (c->rank < c->nranks - 1) {
    const int* Rn = &all[4 * (c->rank + 1)];
    UFE_CHECK(r.hi <= Rn[3], "halo of rank %d reaches beyond its right neighbour: too many ranks for this mesh", c->rank);
    r.send_r = std::max(0, r.i2 - Rn[0] + 1);
    UFE_CHECK(r.send_r <= r.n_loc(), "right neighbour's halo is wider than the owned range of rank %d", c->rank);
  }
  return 0;
}

__global__ void k_csr_find_diag(int m, int row_glob0, int col_lo, const int* __restrict__ ptr, const int* __restrict__ ind,
                                int* __restrict__ diag) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < m; r += (long long)gridDim.x * blockDim.x) {
    const int target = row_glob0 + (int)r - col_lo;  // extended-column position of the diagonal
    int d = -1;
    for (int k = ptr[r]; k < ptr[r + 1]; ++k) if (ind[k] == target) { d = k; break; }
    diag[r] = d;
  }
}

// inverse of the (summed) diagonal; zero diagonal -> 1 (PCJacobi behaviour)
__global__ void k_csr_dinv(int m, int row_glob0, int col_lo, const int* __restrict__ ptr, const int* __restrict__ ind,
                           const double* __restrict__ val, double* __restrict__ dinv, int pc_none) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < m; r += (long long)gridDim.x * blockDim.x) {
    if (pc_none) { dinv[r] = 1.0; continue; }
    const int target = row_glob0 + (int)r - col_lo;
    double d = 0.0;
    for (int k = ptr[r]; k < ptr[r + 1]; ++k) if (ind[k] == target) d += val[k];
    dinv[r] = (d != 0.0) ? 1.0 / d : 1.0;
  }
}

// Upload rows i1..i1+m_loc-1 given in the reference's layout (1-based relative offsets, 1-based global columns).
int csr_upload(Ctx* c, DevCsr& A, int m_glob, int n_glob, int m_loc, int i1, int n_loc, int j1, const int32_t* ptr,
               const int32_t* ind, const double* val, bool square, const Range* force_cols) {
  UFE_CHECK(m_loc >= 0 && ptr && (m_loc == 0 || (ind && val)), "csr_upload: NULL arrays");
  const long long nnz = (m_loc > 0) ? (long long)ptr[m_loc] - ptr[0] : 0;
  UFE_CHECK(nnz >= 0, "csr_upload: ptr is not ascending");
  A.m_loc = m_loc; A.nnz = nnz;
  A.rows.n_glob = m_glob; A.rows.i1 = i1; A.rows.i2 = i1 + m_loc - 1; A.rows.lo = i1; A.rows.hi = i1 + m_loc - 1;
  A.cols.n_glob = n_glob; A.cols.i1 = j1; A.cols.i2 = j1 + n_loc - 1;
  int jmin = j1, jmax = j1 + n_loc - 1;
  for (long long k = 0; k < nnz; ++k) {
    const int j = ind[k];
    UFE_CHECK(j >= 1 && j <= n_glob, "csr_upload: column index %d out of range 1..%d", j, n_glob);
    jmin = std::min(jmin, j); jmax = std::max(jmax, j);
  }
  if (c->nranks == 1) UFE_CHECK(jmin >= j1 && jmax <= j1 + n_loc - 1, "csr_upload: on one rank every column must lie in the owned x range");
  if (force_cols) {
    UFE_CHECK(jmin >= force_cols->lo && jmax <= force_cols->hi, "csr_upload: columns %d..%d outside the extended range %d..%d", jmin, jmax,
              force_cols->lo, force_cols->hi);
    A.cols = *force_cols;
  } else {
    A.cols.lo = jmin; A.cols.hi = jmax;
    UFE_TRY(setup_range_halos(c, A.cols));
  }
  std::vector<int> p0((size_t)m_loc + 1), i0((size_t)nnz);
  for (int r = 0; r <= m_loc; ++r) p0[r] = ptr[r] - ptr[0];
  for (long long k = 0; k < nnz; ++k) i0[k] = ind[k] - A.cols.lo;
  UFE_TRY(A.ptr_buf.alloc(m_loc + 1)); UFE_TRY(A.ind_buf.alloc(nnz)); UFE_TRY(A.val_buf.alloc(nnz));
  A.ptr = A.ptr_buf.p; A.ind = A.ind_buf.p; A.val = A.val_buf.p;
  UFE_CUDA(cudaMemcpyAsync(A.ptr, p0.data(), sizeof(int) * p0.size(), cudaMemcpyHostToDevice, c->stream));
  if (nnz) {
    UFE_CUDA(cudaMemcpyAsync(A.ind, i0.data(), sizeof(int) * nnz, cudaMemcpyHostToDevice, c->stream));
    UFE_CUDA(cudaMemcpyAsync(A.val, val, sizeof(double) * nnz, cudaMemcpyHostToDevice, c->stream));
  }
  UFE_CUDA(cudaStreamSynchronize(c->stream));  // p0/i0 go out of scope
  const double avg = m_loc > 0 ? (double)nnz / m_loc : 0.0;
  A.vec_width = avg <= 2.5 ? 2 : avg <= 6.0 ? 4 : avg <= 14.0 ? 8 : avg <= 48.0 ? 16 : 32;
  (void)square;
  A.valid = true;
  return 0;
}

// ---- sliced-ELL copy (DevCsr::sell_*) --------------------------------------------------------------------------------------------------------
// width of every slice of 32 rows = its longest row
__global__ void k_sell_width(int m, int nslices, const int* __restrict__ ptr, int* __restrict__ width) {
  for (long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x; s < nslices; s += (long long)gridDim.x * blockDim.x) {
    int w = 0;
    const int r1 = (int)s * 32, r2 = min(m, r1 + 32);
    for (int r = r1; r < r2; ++r) w = max(w, ptr[r + 1] - ptr[r]);
    width[s] = w;
  }
}
// one thread per (slice, lane): entry k of row r sits at slice[s] + 32 k + lane; padding = value 0 at a column of the same row
template <bool PATTERN>
__global__ void k_sell_fill(int m, int nslices, const int* __restrict__ ptr, const int* __restrict__ ind, const double* __restrict__ val,
                            const int* __restrict__ slice, int* __restrict__ sind, double* __restrict__ sval) {
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < (long long)nslices * 32; t += (long long)gridDim.x * blockDim.x) {
    const int r = (int)t, s = r >> 5, b = slice[s], w = (slice[s + 1] - b) >> 5;
    const int k1 = r < m ? ptr[r] : 0, len = r < m ? ptr[r + 1] - k1 : 0;
    const int pad = len > 0 ? ind[k1 + len - 1] : 0;
    for (int k = 0; k < w; ++k) {
      const size_t slot = (size_t)b + (size_t)k * 32 + (r & 31);
      sval[slot] = k < len ? val[k1 + k] : 0.0;
      if (PATTERN) sind[slot] = k < len ? ind[k1 + k] : pad;
    }
  }
}

// UFE_B200_CSR_KERNEL=vector keeps the vector-per-row kernels everywhere (the comparison of profiles/r02_spmv_probe_*.log)
static bool sell_disabled() {
  static const bool off = [] { const char* e = getenv("UFE_B200_CSR_KERNEL"); return e && !strcmp(e, "vector"); }();
  return off;
}

int csr_sell_build(Ctx* c, DevCsr& A) {
  if (A.m_loc == 0 || sell_disabled()) return 0;
  const int nslices = (A.m_loc + 31) / 32;
  DBuf<int> width;
  UFE_TRY(width.alloc(nslices));
  {
    UFE_LAUNCH(c, TC_OPERATOR_SPMV);
    k_sell_width<<<grid_for(c, nslices, kBlock), kBlock, 0, c->stream>>>(A.m_loc, nslices, A.ptr, width.p);
  }
  UFE_CUDA(cudaGetLastError());
  std::vector<int> w(nslices), sl((size_t)nslices + 1);     // once per pattern: the prefix sum runs on the host
  UFE_CUDA(cudaMemcpyAsync(w.data(), width.p, sizeof(int) * nslices, cudaMemcpyDeviceToHost, c->stream));
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  long long slots = 0;
  for (int s = 0; s < nslices; ++s) { sl[s] = (int)slots; slots += 32LL * w[s]; }
  UFE_CHECK(slots < (1LL << 31), "sliced-ELL copy: %lld slots do not fit 32-bit offsets", slots);
  sl[nslices] = (int)slots;
  UFE_TRY(A.sell_slice.alloc((size_t)nslices + 1)); UFE_TRY(A.sell_ind.alloc((size_t)std::max(slots, 1LL))); UFE_TRY(A.sell_val.alloc((size_t)std::max(slots, 1LL)));
  UFE_CUDA(cudaMemcpyAsync(A.sell_slice.p, sl.data(), sizeof(int) * sl.size(), cudaMemcpyHostToDevice, c->stream));
  {
    UFE_LAUNCH(c, TC_OPERATOR_SPMV);
    k_sell_fill<true><<<grid_for(c, (long long)nslices * 32, kBlock), kBlock, 0, c->stream>>>(A.m_loc, nslices, A.ptr, A.ind, A.val, A.sell_slice.p,
                                                                                            A.sell_ind.p, A.sell_val.p);
  }
  UFE_CUDA(cudaGetLastError());
  UFE_CUDA(cudaStreamSynchronize(c->stream));   // sl goes out of scope
  A.sell_src_ptr = A.ptr; A.sell_src_nnz = A.nnz;
  return 0;
}

int csr_sell_refresh(Ctx* c, DevCsr& A) {
  if (!A.has_sell()) return csr_sell_build(c, A);
  if (A.m_loc == 0) return 0;
  const int nslices = (A.m_loc + 31) / 32;
  UFE_LAUNCH(c, TC_OPERATOR_SPMV);
  k_sell_fill<false><<<grid_for(c, (long long)nslices * 32, kBlock), kBlock, 0, c->stream>>>(A.m_loc, nslices, A.ptr, A.ind, A.val, A.sell_slice.p,
                                                                                           nullptr, A.sell_val.p);
  UFE_CUDA(cudaGetLastError());
  return 0;
}

template <int W>
static void launch_spmv_plain(Ctx* c, const DevCsr& A, const double* x_ext, double* y, int tc) {
  CsrOp<W> op{A.ptr, A.ind, A.val};
  UFE_LAUNCH(c, tc);
  k_spmv_fused<CsrOp<W>, 0><<<grid_for(c, (long long)A.m_loc * W, kBlock), kBlock, 0, c->stream>>>(
      op, A.m_loc, 0, 0, x_ext, nullptr, y, nullptr, nullptr, nullptr, nullptr, nullptr);
}

int csr_spmv(Ctx* c, const DevCsr& A, const double* x_ext, double* y, int tc) {
  if (A.m_loc == 0) return 0;
  if (A.has_sell()) {
    SellOp op{A.sell_slice.p, A.sell_ind.p, A.sell_val.p};
    UFE_LAUNCH(c, tc);
    k_spmv_fused<SellOp, 0><<<grid_for(c, A.m_loc, kBlock), kBlock, 0, c->stream>>>(op, A.m_loc, 0, 0, x_ext, nullptr, y, nullptr, nullptr, nullptr,
                                                                                 nullptr, nullptr);
    UFE_CUDA(cudaGetLastError());
    return 0;
  }
  switch (A.vec_width) {
    case 2: launch_spmv_plain<2>(c, A, x_ext, y, tc); break;
    case 4: launch_spmv_plain<4>(c, A, x_ext, y, tc); break;
    case 8: launch_spmv_plain<8>(c, A, x_ext, y, tc); break;
    case 16: launch_spmv_plain<16>(c, A, x_ext, y, tc); break;
    default: launch_spmv_plain<32>(c, A, x_ext, y, tc); break;
  }
  UFE_CUDA(cudaGetLastError());
  return 0;
}

// multi-column product: thread per row keeps the row in registers and streams the ncols right-hand sides
// (K2 of SURVEY 2e: the matrix is read once instead of nz times as in CSR_matrix_vector_multiplication.f90:356-360)
__global__ void k_spmv_multicol(int m, int n_ext, int ncols, const int* __restrict__ ptr, const int* __restrict__ ind,
                                const double* __restrict__ val, const double* __restrict__ x, double* __restrict__ y) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < m; r += (long long)gridDim.x * blockDim.x) {
    const int k1 = ptr[r], k2 = ptr[r + 1];
    for (int c0 = 0; c0 < ncols; ++c0) {
      const double* xc = x + (size_t)c0 * n_ext;
      double s = 0.0;
      for (int k = k1; k < k2; ++k) s += val[k] * xc[ind[k]];
      y[(size_t)c0 * m + r] = s;
    }
  }
}

// the same product on the sliced-ELL copy: one row per thread, the row's entries read once per chunk of NC columns (coalesced, see SellOp), NC
// accumulators and NC independent gathers per entry in flight; blockIdx.y = column chunk
template <int NC>
__global__ void __launch_bounds__(kBlock) k_sell_multicol(int m, int n_ext, int ncols, const int* __restrict__ slice, const int* __restrict__ ind,
                                                          const double* __restrict__ val, const double* __restrict__ x, double* __restrict__ y) {
  const int c0 = blockIdx.y * NC, nc = min(NC, ncols - c0);
  const double* __restrict__ xc = x + (size_t)c0 * n_ext;
  for (long long rr = (long long)blockIdx.x * blockDim.x + threadIdx.x; rr < m; rr += (long long)gridDim.x * blockDim.x) {
    const int r = (int)rr, s = r >> 5, b = slice[s], w = (slice[s + 1] - b) >> 5;
    const int* __restrict__ ip = ind + b + (r & 31);
    const double* __restrict__ vp = val + b + (r & 31);
    double acc[NC];
#pragma unroll
    for (int q = 0; q < NC; ++q) acc[q] = 0.0;
    for (int k = 0; k < w; ++k) {
      const int j = ip[k * 32];
      const double v = vp[k * 32];
#pragma unroll
      for (int q = 0; q < NC; ++q)
        if (q < nc) acc[q] += v * xc[(size_t)q * n_ext + j];
    }
#pragma unroll
    for (int q = 0; q < NC; ++q)
      if (q < nc) y[(size_t)(c0 + q) * m + r] = acc[q];
  }
}

// [k][i] -> [i][k] through a shared-memory tile (reads and writes in full segments)
__global__ void __launch_bounds__(128) k_interleave_columns(int n, int ncols, const double* __restrict__ in, double* __restrict__ out) {
  extern __shared__ double tile[];   // [128][ncols + 1]
  const int pitch = ncols + 1;
  for (long long base = (long long)blockIdx.x * 128; base < n; base += (long long)gridDim.x * 128) {
    const int cnt = (int)min((long long)128, n - base);
    if ((int)threadIdx.x < cnt)
      for (int k = 0; k < ncols; ++k) tile[threadIdx.x * pitch + k] = in[(size_t)k * n + base + threadIdx.x];
    __syncthreads();
    for (int t = threadIdx.x; t < cnt * ncols; t += 128) out[(size_t)base * ncols + t] = tile[(t / ncols) * pitch + t % ncols];
    __syncthreads();
  }
}
// the multi-column product on a column-interleaved copy xT[j][c] of the right-hand sides: the NC values an entry needs are one contiguous segment
// (8 NC bytes) instead of NC sectors 8 n_ext bytes apart -- the ncu capture of k_sell_multicol showed the product bound by exactly those sector-sized
// gathers (DRAM traffic = algorithmic bytes at 20 % of the bandwidth, profiles/r02_generic_spmv_4M_ncu.md)
template <int NC>
__global__ void __launch_bounds__(kBlock) k_sell_multicol_T(int m, int ncols, const int* __restrict__ slice, const int* __restrict__ ind,
                                                            const double* __restrict__ val, const double* __restrict__ xT, double* __restrict__ y) {
  const int c0 = blockIdx.y * NC, nc = min(NC, ncols - c0);
  for (long long rr = (long long)blockIdx.x * blockDim.x + threadIdx.x; rr < m; rr += (long long)gridDim.x * blockDim.x) {
    const int r = (int)rr, s = r >> 5, b = slice[s], w = (slice[s + 1] - b) >> 5;
    const int* __restrict__ ip = ind + b + (r & 31);
    const double* __restrict__ vp = val + b + (r & 31);
    double acc[NC];
#pragma unroll
    for (int q = 0; q < NC; ++q) acc[q] = 0.0;
    for (int k = 0; k < w; ++k) {
      const double* __restrict__ xr = xT + (size_t)ip[k * 32] * ncols + c0;
      const double v = vp[k * 32];
#pragma unroll
      for (int q = 0; q < NC; ++q)
        if (q < nc) acc[q] += v * xr[q];
    }
#pragma unroll
    for (int q = 0; q < NC; ++q)
      if (q < nc) y[(size_t)(c0 + q) * m + r] = acc[q];
  }
}

template <int NC>
static void launch_sell_multicol_T(Ctx* c, const DevCsr& A, const double* xT, double* y, int ncols) {
  dim3 grid(grid_for(c, A.m_loc, kBlock), (ncols + NC - 1) / NC);
  k_sell_multicol_T<NC><<<grid, kBlock, 0, c->stream>>>(A.m_loc, ncols, A.sell_slice.p, A.sell_ind.p, A.sell_val.p, xT, y);
}

template <int NC>
static void launch_sell_multicol(Ctx* c, const DevCsr& A, const double* x_ext, double* y, int ncols) {
  dim3 grid(grid_for(c, A.m_loc, kBlock), (ncols + NC - 1) / NC);
  k_sell_multicol<NC><<<grid, kBlock, 0, c->stream>>>(A.m_loc, A.cols.n_ext(), ncols, A.sell_slice.p, A.sell_ind.p, A.sell_val.p, x_ext, y);
}

int csr_spmv_multicol(Ctx* c, const DevCsr& A, const double* x_ext, double* y, int ncols, int tc) {
  if (ncols == 1) return csr_spmv(c, A, x_ext, y, tc);
  if (A.m_loc == 0) return 0;
  // Measured on the operators of the 4 M-vertex mesh (profiles/r02_spmv_probe_sell_v2.log): the interleaved copy does not pay -- 0.94 / 1.01 / 1.14 ms
  // against 0.90 / 0.85 / 0.93 ms for map_a_b / ddx_a_a / ddx_b_a with 12 columns -- so the product is bound by the dependent index -> value chain,
  // not by the sectors; the path stays selectable (UFE_B200_MULTICOL_INTERLEAVE=1) and tested.
  if (A.has_sell() && ncols >= 4 && ncols <= 40 && getenv("UFE_B200_MULTICOL_INTERLEAVE") && atoi(getenv("UFE_B200_MULTICOL_INTERLEAVE")) == 1) {
    const int n_ext = A.cols.n_ext();
    UFE_TRY(c->xT.alloc((size_t)n_ext * ncols));
    {
      UFE_LAUNCH(c, tc);
      k_interleave_columns<<<grid_for(c, n_ext, 128), 128, sizeof(double) * 128 * (size_t)(ncols + 1), c->stream>>>(n_ext, ncols, x_ext, c->xT.p);
    }
    UFE_LAUNCH(c, tc);
    if (ncols <= 4) launch_sell_multicol_T<4>(c, A, c->xT.p, y, ncols);
    else if (ncols <= 8) launch_sell_multicol_T<8>(c, A, c->xT.p, y, ncols);
    else if (ncols <= 12 || (ncols > 16 && ncols <= 24)) launch_sell_multicol_T<12>(c, A, c->xT.p, y, ncols);
    else launch_sell_multicol_T<16>(c, A, c->xT.p, y, ncols);
    UFE_CUDA(cudaGetLastError());
    return 0;
  }
  if (A.has_sell()) {
    UFE_LAUNCH(c, tc);
    if (ncols <= 4) launch_sell_multicol<4>(c, A, x_ext, y, ncols);
    else if (ncols <= 8) launch_sell_multicol<8>(c, A, x_ext, y, ncols);
    else if (ncols <= 12 || (ncols > 16 && ncols <= 24)) launch_sell_multicol<12>(c, A, x_ext, y, ncols);
    else launch_sell_multicol<16>(c, A, x_ext, y, ncols);
    UFE_CUDA(cudaGetLastError());
    return 0;
  }
  UFE_LAUNCH(c, tc);
  k_spmv_multicol<<<grid_for(c, A.m_loc, kBlock), kBlock, 0, c->stream>>>(A.m_loc, A.cols.n_ext(), ncols, A.ptr, A.ind,
                                                                         A.val, x_ext, y);
  UFE_CUDA(cudaGetLastError());
  return 0;
}

int csr_make_dinv(Ctx* c, const DevCsr& A, double* dinv, int pc_type) {
  UFE_LAUNCH(c, TC_KRYLOV_VEC);
  k_csr_dinv<<<grid_for(c, A.m_loc, kBlock), kBlock, 0, c->stream>>>(A.m_loc, A.rows.i1, A.cols.lo, A.ptr, A.ind, A.val,
                                                                    dinv, pc_type == UFE_PC_NONE ? 1 : 0);
  UFE_CUDA(cudaGetLastError());
  return 0;
}

int csr_ksp_solve(Ctx* c, DevCsr& A, const double* d_b, double* d_x_ext, const KrylovParams& kp, int tc, int* its) {
  UFE_CHECK(kp.ksp_type == UFE_KSP_BICGSTAB, "only BiCGStab is available for general (non-symmetric) matrices");
  UFE_CHECK(kp.pc_type == UFE_PC_JACOBI || kp.pc_type == UFE_PC_NONE, "generic CSR solve supports pc = jacobi | none");
  if (!c->krylov) c->krylov = new Krylov();
  UFE_TRY(c->krylov->dinv.alloc(A.m_loc));
  UFE_TRY(csr_make_dinv(c, A, c->krylov->dinv.p, kp.pc_type));
  const double* dinv = c->krylov->dinv.p;
  // A Krylov solve multiplies with A at least twice per iteration: one pass that (re)fills the sliced-ELL copy pays off from the first iterations on.
  if (!sell_disabled() && A.m_loc > 0) {
    UFE_TRY(csr_sell_refresh(c, A));
    return bicgstab_solve(c, SellOp{A.sell_slice.p, A.sell_ind.p, A.sell_val.p}, A.cols, d_b, dinv, d_x_ext, kp, tc, its);
  }
  switch (A.vec_width) {
    case 2: return bicgstab_solve(c, CsrOp<2>{A.ptr, A.ind, A.val}, A.cols, d_b, dinv, d_x_ext, kp, tc, its);
    case 4: return bicgstab_solve(c, CsrOp<4>{A.ptr, A.ind, A.val}, A.cols, d_b, dinv, d_x_ext, kp, tc, its);
    case 8: return bicgstab_solve(c, CsrOp<8>{A.ptr, A.ind, A.val}, A.cols, d_b, dinv, d_x_ext, kp, tc, its);
    case 16: return bicgstab_solve(c, CsrOp<16>{A.ptr, A.ind, A.val}, A.cols, d_b, dinv, d_x_ext, kp, tc, its);
    default: return bicgstab_solve(c, CsrOp<32>{A.ptr, A.ind, A.val}, A.cols, d_b, dinv, d_x_ext, kp, tc, its);
  }
}

}  // namespace ufe

using namespace ufe;

extern "C" {

int ufe_b200_ksp_solve(ufe_b200_handle h, int m_glob, int m_loc, int i1, const int32_t* ptr, const int32_t* ind,
                       const double* val, const double* bb, double* xx, double rtol, double abstol, int maxit, int ksp_type,
                       int pc_type, int* n_Axb_its) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c, "NULL handle");
  UFE_CUDA(cudaSetDevice(c->device));
  DevCsr A;
  UFE_TRY(csr_upload(c, A, m_glob, m_glob, m_loc, i1, m_loc, i1, ptr, ind, val, true));
  DBuf<double> b, x;
  UFE_TRY(b.alloc(m_loc)); UFE_TRY(x.alloc(A.cols.n_ext()));
  UFE_CUDA(cudaMemcpyAsync(b.p, bb, sizeof(double) * m_loc, cudaMemcpyHostToDevice, c->stream));
  UFE_CUDA(cudaMemsetAsync(x.p, 0, sizeof(double) * A.cols.n_ext(), c->stream));
  KrylovParams kp{rtol, abstol, maxit, ksp_type, pc_type};
  int its = 0;
  UFE_TRY(csr_ksp_solve(c, A, b.p, x.p, kp, TC_SPMV_STIFFNESS, &its));
  UFE_CUDA(cudaMemcpyAsync(xx, x.p + A.cols.off(), sizeof(double) * m_loc, cudaMemcpyDeviceToHost, c->stream));
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  if (n_Axb_its) *n_Axb_its = its;
  return 0;
}

int ufe_b200_csr_register(ufe_b200_handle h, int id, int m_glob, int n_glob, int m_loc, int i1, int n_loc, int j1,
                          const int32_t* ptr, const int32_t* ind, const double* val) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c, "NULL handle");
  UFE_CUDA(cudaSetDevice(c->device));
  auto it = c->csr.find(id);
  if (it != c->csr.end()) { delete it->second; c->csr.erase(it); }
  DevCsr* A = new DevCsr();
  int rc = csr_upload(c, *A, m_glob, n_glob, m_loc, i1, n_loc, j1, ptr, ind, val, false);
  if (!rc) rc = csr_sell_build(c, *A);      // registered matrices are multiplied repeatedly (the SpMV seam)
  if (rc) { delete A; return rc; }
  c->csr[id] = A;
  return 0;
}

int ufe_b200_csr_update_values(ufe_b200_handle h, int id, const double* val) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  auto it = c->csr.find(id);
  UFE_CHECK(it != c->csr.end(), "ufe_b200_csr_update_values: matrix %d is not registered", id);
  UFE_CUDA(cudaSetDevice(c->device));
  UFE_CUDA(cudaMemcpyAsync(it->second->val, val, sizeof(double) * it->second->nnz, cudaMemcpyHostToDevice, c->stream));
  UFE_TRY(csr_sell_refresh(c, *it->second));
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

int ufe_b200_spmv(ufe_b200_handle h, int id, const double* xx, double* yy, int ncols) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  auto it = c->csr.find(id);
  UFE_CHECK(it != c->csr.end(), "ufe_b200_spmv: matrix %d is not registered", id);
  UFE_CHECK(ncols >= 1, "ufe_b200_spmv: ncols must be >= 1");
  UFE_CUDA(cudaSetDevice(c->device));
  DevCsr& A = *it->second;
  const int n_loc = A.cols.n_loc(), n_ext = A.cols.n_ext();
  DBuf<double> x, y;
  UFE_TRY(x.alloc((size_t)n_ext * ncols)); UFE_TRY(y.alloc((size_t)A.m_loc * ncols));
  if (n_ext > n_loc) UFE_CUDA(cudaMemsetAsync(x.p, 0, sizeof(double) * n_ext * ncols, c->stream));
  UFE_CUDA(cudaMemcpy2DAsync(x.p + A.cols.off(), sizeof(double) * n_ext, xx, sizeof(double) * n_loc, sizeof(double) * n_loc, ncols,
                             cudaMemcpyHostToDevice, c->stream));
  UFE_TRY(halo_exchange(c, A.cols, x.p, ncols));
  UFE_TRY(csr_spmv_multicol(c, A, x.p, y.p, ncols, TC_OPERATOR_SPMV));
  UFE_CUDA(cudaMemcpyAsync(yy, y.p, sizeof(double) * A.m_loc * ncols, cudaMemcpyDeviceToHost, c->stream));
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

int ufe_b200_csr_release(ufe_b200_handle h, int id) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  auto it = c->csr.find(id);
  if (it != c->csr.end()) { delete it->second; c->csr.erase(it); }
  return 0;
}

}  // extern "C"
