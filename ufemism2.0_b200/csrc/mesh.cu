// Mesh handle: upload of type_mesh, ownership ranges, operator access (wide seam).
//   type_mesh                         src/UPSY/types/mesh_types.f90:17-284
//   partition_list                    src/UPSY/basic/mpi_parallelisation/mpi_distributed_memory.f90:47-73
//   setup_mesh_parallelisation        src/UPSY/mesh/mesh_parallelisation.f90:69-172
//   ddx/ddy/map_{a,b}_{a,b}_{2D,3D}   src/UPSY/mesh/discretisation/mesh_disc_apply_operators.f90:18-431
#include <algorithm>

#include "csr.cuh"
#include "mesh.cuh"
#include "thickness.cuh"
#include "velocity.cuh"

namespace ufe {

int mesh_calc_operators(Ctx* c);  // operators.cu

static void partition_list(int ntot, int i, int n, int& i1, int& i2) {
  if (ntot > n * 2) {
    const int remainder = ntot % n, slice = ntot / n;
    i1 = slice * i + std::min(i, remainder) + 1;
    i2 = slice * (i + 1) + std::min(i + 1, remainder);
  } else if (i == 0) { i1 = 1; i2 = ntot; } else { i1 = 1; i2 = 0; }
}

template <typename T>
static int up(Ctx* c, DBuf<T>& d, const T* h, size_t n, const char* name) {
  UFE_CHECK(h != nullptr, "ufe_b200_mesh_create: mesh array %s is NULL", name);
  UFE_TRY(d.alloc(n));
  UFE_CUDA(cudaMemcpyAsync(d.p, h, sizeof(T) * n, cudaMemcpyHostToDevice, c->stream));
  return 0;
}

__global__ void k_minmax_int(long long n, const int* __restrict__ v, int add, int* __restrict__ mm) {
  int lo = 0x7fffffff, hi = -0x7fffffff;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int x = v[i] + add;
    lo = min(lo, x); hi = max(hi, x);
  }
  for (int o = 16; o > 0; o >>= 1) { lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o)); hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o)); }
  if ((threadIdx.x & 31) == 0) { atomicMin(&mm[0], lo); atomicMax(&mm[1], hi); }
}

// neighbours the thickness stencil of the owned vertices touches: vertices C(vi,:) and the triangles of their edges
__global__ void k_minmax_thickness(int vi1, int n_loc, int nV, int nE, const int* __restrict__ nC, const int* __restrict__ C,
                                   const int* __restrict__ VE, const int* __restrict__ ETri, int* __restrict__ mm_a,
                                   int* __restrict__ mm_b) {
  int alo = 0x7fffffff, ahi = -0x7fffffff, blo = 0x7fffffff, bhi = -0x7fffffff;
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n_loc; r += (long long)gridDim.x * blockDim.x) {
    const int vi = vi1 + (int)r;
    for (int ci = 0; ci < nC[vi - 1]; ++ci) {
      const size_t k = (size_t)ci * nV + vi - 1;
      const int vj = C[k], ei = VE[k];
      alo = min(alo, vj); ahi = max(ahi, vj);
      for (int s = 0; s < 2; ++s) {
        const int t = ETri[(size_t)s * nE + ei - 1];
        if (t > 0) { blo = min(blo, t); bhi = max(bhi, t); }
      }
    }
  }
  atomicMin(&mm_a[0], alo); atomicMax(&mm_a[1], ahi); atomicMin(&mm_b[0], blo); atomicMax(&mm_b[1], bhi);
}

__global__ void k_shift_int(long long n, int* __restrict__ v, int sub) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) v[i] -= sub;
}

// Extended ranges = everything any kernel of this rank gathers from; then convert operator columns to extended-local.
int mesh_finalize_ranges(Ctx* c) {
  Mesh& M = *c->mesh;
  if (c->nranks > 1) {
    DBuf<int> mm;
    UFE_TRY(mm.alloc(4));
    const int init[4] = {M.a.i1, M.a.i2, M.b.i1, M.b.i2};
    UFE_CUDA(cudaMemcpyAsync(mm.p, init, sizeof(init), cudaMemcpyHostToDevice, c->stream));
    for (int g = 0; g < Mesh::G_COUNT; ++g) {
      const bool cols_a = (g == Mesh::G_AA || g == Mesh::G_AB);
      if (M.grp[g].nnz == 0) continue;
      c->launches++;
      k_minmax_int<<<grid_for(c, M.grp[g].nnz, 256), 256, 0, c->stream>>>(M.grp[g].nnz, M.grp[g].ind.p, 1, mm.p + (cols_a ? 0 : 2));
    }
    c->launches++;
    k_minmax_thickness<<<grid_for(c, M.a.n_loc(), 256), 256, 0, c->stream>>>(M.a.i1, M.a.n_loc(), M.nV, M.nE, M.nC.p, M.C.p, M.VE.p,
                                                                             M.ETri.p, mm.p, mm.p + 2);
    int h[4];
    UFE_CUDA(cudaMemcpyAsync(h, mm.p, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
    UFE_CUDA(cudaStreamSynchronize(c->stream));
    M.a.lo = h[0]; M.a.hi = h[1]; M.b.lo = h[2]; M.b.hi = h[3];
  } else {
    M.a.lo = 1; M.a.hi = M.nV; M.b.lo = 1; M.b.hi = M.nTri;
  }
  UFE_TRY(setup_range_halos(c, M.a));
  UFE_TRY(setup_range_halos(c, M.b));
  M.buv.n_glob = 2 * M.nTri; M.buv.i1 = 2 * M.b.i1 - 1; M.buv.i2 = 2 * M.b.i2; M.buv.lo = 2 * M.b.lo - 1; M.buv.hi = 2 * M.b.hi;
  M.buv.send_l = 2 * M.b.send_l; M.buv.send_r = 2 * M.b.send_r;
  for (int g = 0; g < Mesh::G_COUNT; ++g) {
    const bool cols_a = (g == Mesh::G_AA || g == Mesh::G_AB);
    const Range& cr = cols_a ? M.a : M.b;
    if (cr.lo > 1 && M.grp[g].nnz > 0) {
      c->launches++;
      k_shift_int<<<grid_for(c, M.grp[g].nnz, 256), 256, 0, c->stream>>>(M.grp[g].nnz, M.grp[g].ind.p, cr.lo - 1);
    }
    for (int q = 0; q < M.grp[g].nmat; ++q) M.op[M.grp[g].first_op + q].cols = cr;
  }
  UFE_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace ufe

using namespace ufe;

extern "C" {

int ufe_b200_mesh_create(ufe_b200_handle h, const ufe_b200_mesh_desc* d) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && d, "ufe_b200_mesh_create: NULL argument");
  UFE_CUDA(cudaSetDevice(c->device));
  ufe_b200_mesh_destroy(h);
  UFE_CHECK(d->nV > 0 && d->nTri > 0 && d->nE > 0 && d->nC_mem > 0 && d->nz >= 2, "ufe_b200_mesh_create: bad sizes");
  Mesh* M = new Mesh();
  c->mesh = M;
  M->nV = d->nV; M->nTri = d->nTri; M->nE = d->nE; M->nC_mem = d->nC_mem; M->nz = d->nz;
  M->xmin = d->xmin; M->xmax = d->xmax; M->ymin = d->ymin; M->ymax = d->ymax;
  const size_t nV = d->nV, nTri = d->nTri, nE = d->nE, ncm = d->nC_mem;
  UFE_TRY(up(c, M->V, d->V, nV * 2, "V")); UFE_TRY(up(c, M->nC, d->nC, nV, "nC")); UFE_TRY(up(c, M->C, d->C, nV * ncm, "C"));
  UFE_TRY(up(c, M->niTri, d->niTri, nV, "niTri")); UFE_TRY(up(c, M->iTri, d->iTri, nV * ncm, "iTri"));
  UFE_TRY(up(c, M->VBI, d->VBI, nV, "VBI")); UFE_TRY(up(c, M->Tri, d->Tri, nTri * 3, "Tri")); UFE_TRY(up(c, M->TriC, d->TriC, nTri * 3, "TriC"));
  UFE_TRY(up(c, M->TriBI, d->TriBI, nTri, "TriBI")); UFE_TRY(up(c, M->TriGC, d->TriGC, nTri * 2, "TriGC"));
  UFE_TRY(up(c, M->A, d->A, nV, "A")); UFE_TRY(up(c, M->Cw, d->Cw, nV * ncm, "Cw")); UFE_TRY(up(c, M->D_x, d->D_x, nV * ncm, "D_x"));
  UFE_TRY(up(c, M->D_y, d->D_y, nV * ncm, "D_y")); UFE_TRY(up(c, M->D, d->D, nV * ncm, "D")); UFE_TRY(up(c, M->VE, d->VE, nV * ncm, "VE"));
  UFE_TRY(up(c, M->ETri, d->ETri, nE * 2, "ETri")); UFE_TRY(up(c, M->zeta, d->zeta, (size_t)d->nz, "zeta"));
  if (d->Tricc) UFE_TRY(up(c, M->Tricc, d->Tricc, nTri * 2, "Tricc"));
  M->h_zeta.assign(d->zeta, d->zeta + d->nz);
  auto setr = [&](Range& r, int n) {
    r.n_glob = n;
    partition_list(n, c->rank, c->nranks, r.i1, r.i2);
    r.lo = r.i1; r.hi = r.i2;
  };
  setr(M->a, d->nV); setr(M->b, d->nTri); setr(M->c, d->nE);
  UFE_CHECK(M->a.n_loc() > 0 && M->b.n_loc() > 0, "rank %d owns no vertices/triangles: mesh too small for %d ranks", c->rank, c->nranks);
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

int ufe_b200_mesh_destroy(ufe_b200_handle h) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  if (!c) return 0;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  velocity_free(c->vel); c->vel = nullptr;
  thickness_free(c->thick); c->thick = nullptr;
  secondary_free(c->sec); c->sec = nullptr;
  geometry_free(c->geo); c->geo = nullptr;
  pc_state_free(c->pc); c->pc = nullptr;
  delete c->mesh; c->mesh = nullptr;
  return 0;
}

int ufe_b200_mesh_ranges(ufe_b200_handle h, int32_t* out6) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && c->mesh, "ufe_b200_mesh_ranges: no mesh");
  const Mesh& M = *c->mesh;
  out6[0] = M.a.i1; out6[1] = M.a.i2; out6[2] = M.b.i1; out6[3] = M.b.i2; out6[4] = M.c.i1; out6[5] = M.c.i2;
  return 0;
}

int ufe_b200_calc_matrix_operators(ufe_b200_handle h) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && c->mesh, "ufe_b200_calc_matrix_operators: no mesh");
  UFE_CUDA(cudaSetDevice(c->device));
  velocity_free(c->vel); c->vel = nullptr;  // patterns derived from the operators are rebuilt on demand
  return mesh_calc_operators(c);
}

int ufe_b200_get_operator_size(ufe_b200_handle h, int op, int* m_loc, int* nnz_loc) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && c->mesh && op >= 0 && op < UFE_OP_COUNT && c->mesh->op[op].valid, "ufe_b200_get_operator_size: operator %d not available", op);
  *m_loc = c->mesh->op[op].m_loc; *nnz_loc = (int)c->mesh->op[op].nnz;
  return 0;
}

// returned in the reference's layout: ptr = 1-based offsets, ind = 1-based global columns
int ufe_b200_get_operator(ufe_b200_handle h, int op, int32_t* ptr, int32_t* ind, double* val) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && c->mesh && op >= 0 && op < UFE_OP_COUNT && c->mesh->op[op].valid, "ufe_b200_get_operator: operator %d not available", op);
  UFE_CUDA(cudaSetDevice(c->device));
  const DevCsr& A = c->mesh->op[op];
  UFE_CUDA(cudaMemcpyAsync(ptr, A.ptr, sizeof(int) * (A.m_loc + 1), cudaMemcpyDeviceToHost, c->stream));
  UFE_CUDA(cudaMemcpyAsync(ind, A.ind, sizeof(int) * A.nnz, cudaMemcpyDeviceToHost, c->stream));
  UFE_CUDA(cudaMemcpyAsync(val, A.val, sizeof(double) * A.nnz, cudaMemcpyDeviceToHost, c->stream));
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  for (int r = 0; r <= A.m_loc; ++r) ptr[r] += 1;
  for (long long k = 0; k < A.nnz; ++k) ind[k] += A.cols.lo;
  return 0;
}

int ufe_b200_set_operator(ufe_b200_handle h, int op, int m_loc, const int32_t* ptr, const int32_t* ind, const double* val) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && c->mesh && op >= 0 && op < UFE_OP_COUNT, "ufe_b200_set_operator: bad operator %d", op);
  UFE_CHECK(c->nranks == 1, "ufe_b200_set_operator is only available on one rank; use ufe_b200_calc_matrix_operators");
  UFE_CUDA(cudaSetDevice(c->device));
  Mesh& M = *c->mesh;
  const bool rows_a = (op <= UFE_OP_DDY_A_A) || (op >= UFE_OP_MAP_B_A && op <= UFE_OP_DDY_B_A);
  const bool cols_a = (op <= UFE_OP_DDY_A_B);
  const Range& rr = rows_a ? M.a : M.b;
  const Range& cr = cols_a ? M.a : M.b;
  UFE_CHECK(m_loc == rr.n_loc(), "ufe_b200_set_operator: operator %d needs %d rows, got %d", op, rr.n_loc(), m_loc);
  DevCsr& A = M.op[op];
  A = DevCsr();
  UFE_TRY(csr_upload(c, A, rr.n_glob, cr.n_glob, m_loc, rr.i1, cr.n_loc(), cr.i1, ptr, ind, val, false, &cr));
  return 0;
}

int ufe_b200_apply_operator(ufe_b200_handle h, int op, const double* d_in, double* d_out, int ncols) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && c->mesh && op >= 0 && op < UFE_OP_COUNT && c->mesh->op[op].valid, "ufe_b200_apply_operator: operator %d not available", op);
  UFE_CHECK(ncols >= 1, "ufe_b200_apply_operator: ncols must be >= 1");
  UFE_CUDA(cudaSetDevice(c->device));
  DevCsr& A = c->mesh->op[op];
  if (!A.has_sell()) UFE_TRY(csr_sell_build(c, A));     // first product with this operator: the sliced-ELL copy every later one streams
  const int n_loc = A.cols.n_loc(), n_ext = A.cols.n_ext();
  DBuf<double> x, y;
  UFE_TRY(x.alloc((size_t)n_ext * ncols)); UFE_TRY(y.alloc((size_t)A.m_loc * ncols));
  if (n_ext > n_loc) UFE_CUDA(cudaMemsetAsync(x.p, 0, sizeof(double) * n_ext * ncols, c->stream));
  UFE_CUDA(cudaMemcpy2DAsync(x.p + A.cols.off(), sizeof(double) * n_ext, d_in, sizeof(double) * n_loc, sizeof(double) * n_loc, ncols,
                             cudaMemcpyHostToDevice, c->stream));
  UFE_TRY(halo_exchange(c, A.cols, x.p, ncols));
  UFE_TRY(csr_spmv_multicol(c, A, x.p, y.p, ncols, TC_OPERATOR_SPMV));
  UFE_CUDA(cudaMemcpyAsync(d_out, y.p, sizeof(double) * A.m_loc * ncols, cudaMemcpyDeviceToHost, c->stream));
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

}  // extern "C"
