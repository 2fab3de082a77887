// On-device construction of the 15 mesh operator matrices (a6 of SURVEY 8a; "next" row f4 done early).
// Replaces calc_all_matrix_operators_mesh and its five group builders
//   src/UPSY/mesh/discretisation/mesh_disc_calc_matrix_operators_2D.f90:26-764
// with the flood fill of extend_group_single_iteration_{a,b} (src/UPSY/mesh/mesh_utilities.f90:1856-1935).
// One thread owns one matrix row: it grows the neighbour group in exactly the reference's order (so the CSR
// column order is bit-identical), solves the small weighted least-squares problem and writes the row.
// Two passes (count, exclusive scan, fill) replace the reference's append-only add_entry_CSR_dist.
#include <cub/device/device_scan.cuh>

#include "mesh.cuh"
#include "shape_functions.cuh"

namespace ufe {

namespace {

constexpr int MAXN = 40;  // largest neighbour group a row may collect before it is reported as an error

struct MV {
  int nV, nTri, nC_mem;
  const double *Vx, *Vy, *Gx, *Gy;
  const int *nC, *C, *niTri, *iTri, *Tri, *TriC;
};

__device__ __forceinline__ bool in_stack(const int* st, int n, int v) {
  for (int i = 0; i < n; ++i)
    if (st[i] == v) return true;
  return false;
}

// extend_group_single_iteration_a (mesh_utilities.f90:1856-1894)
__device__ int extend_a(const MV& m, int* st, int n, bool* overflow) {
  const int n0 = n;
  for (int i = 0; i < n0; ++i) {
    const int vi = st[i];
    const int nc = m.nC[vi - 1];
    for (int ci = 0; ci < nc; ++ci) {
      const int vj = m.C[(size_t)ci * m.nV + vi - 1];
      if (!in_stack(st, n, vj)) {
        if (n >= MAXN) { *overflow = true; return n; }
        st[n++] = vj;
      }
    }
  }
  return n;
}
// extend_group_single_iteration_b (mesh_utilities.f90:1896-1935)
__device__ int extend_b(const MV& m, int* st, int n, bool* overflow) {
  const int n0 = n;
  for (int i = 0; i < n0; ++i) {
    const int ti = st[i];
    for (int n2 = 0; n2 < 3; ++n2) {
      const int tj = m.TriC[(size_t)n2 * m.nTri + ti - 1];
      if (tj == 0) continue;
      if (!in_stack(st, n, tj)) {
        if (n >= MAXN) { *overflow = true; return n; }
        st[n++] = tj;
      }
    }
  }
  return n;
}

// G: Mesh::G_AA .. G_M2.  FILL = false: write the row length to ptr[r]; FILL = true: ptr holds offsets, write the row.
template <int G, bool FILL>
__global__ void __launch_bounds__(128) k_build_group(MV m, int row1, int m_loc, int* __restrict__ ptr, int* __restrict__ ind,
                                                     double* __restrict__ v0, double* __restrict__ v1, double* __restrict__ v2,
                                                     double* __restrict__ v3, double* __restrict__ v4, int* __restrict__ err) {
  constexpr bool cols_are_vertices = (G == Mesh::G_AA || G == Mesh::G_AB);
  constexpr bool has_home = (G == Mesh::G_AA || G == Mesh::G_BB || G == Mesh::G_M2);
  constexpr int n_min = (G == Mesh::G_M2) ? 5 : (G == Mesh::G_AA || G == Mesh::G_BB) ? 2 : 3;
  constexpr int NMAT = (G == Mesh::G_M2) ? 5 : (G == Mesh::G_AA || G == Mesh::G_BB) ? 2 : 3;
  int st[MAXN];
  double xc[MAXN], yc[MAXN], Nc[NMAT * MAXN], Ni[5];
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < m_loc; r += (long long)gridDim.x * blockDim.x) {
    const int row = row1 + (int)r;  // global 1-based row = vertex (AA, BA) or triangle (AB, BB, M2)
    bool overflow = false;
    int n = 0;
    double x, y;
    if (G == Mesh::G_AA || G == Mesh::G_BA) { x = m.Vx[row - 1]; y = m.Vy[row - 1]; } else { x = m.Gx[row - 1]; y = m.Gy[row - 1]; }
    if (G == Mesh::G_AA || G == Mesh::G_BB || G == Mesh::G_M2) st[n++] = row;
    if (G == Mesh::G_AB) for (int k = 0; k < 3; ++k) st[n++] = m.Tri[(size_t)k * m.nTri + row - 1];
    if (G == Mesh::G_BA) {
      const int nt = m.niTri[row - 1];
      for (int k = 0; k < nt && n < MAXN; ++k) st[n++] = m.iTri[(size_t)k * m.nV + row - 1];
    }
    int guard = 0;
    while ((has_home ? n - 1 : n) < n_min && !overflow && guard++ < 16) n = cols_are_vertices ? extend_a(m, st, n, &overflow) : extend_b(m, st, n, &overflow);
    bool ok = false;
    int n_c = 0;
    while (!ok && !overflow && guard++ < 32) {
      n_c = 0;
      for (int i = 0; i < n; ++i) {
        const int j = st[i];
        if (has_home && j == row) continue;
        st[n_c] = j;  // compact in place: the home element is always st[0], so this only shifts left by one
        xc[n_c] = cols_are_vertices ? m.Vx[j - 1] : m.Gx[j - 1];
        yc[n_c] = cols_are_vertices ? m.Vy[j - 1] : m.Gy[j - 1];
        ++n_c;
      }
      if (G == Mesh::G_AA || G == Mesh::G_BB) ok = sf_reg_1st(x, y, n_c, xc, yc, Ni, Nc);
      else if (G == Mesh::G_M2) ok = sf_reg_2nd(x, y, n_c, xc, yc, Ni, Nc);
      else ok = sf_stag_1st(x, y, n_c, xc, yc, Nc);
      if (!ok) {
        // singular normal matrix: restore the stack (home first) and add another ring of neighbours
        if (has_home) { for (int i = n_c; i > 0; --i) st[i] = st[i - 1]; st[0] = row; }
        n = cols_are_vertices ? extend_a(m, st, n, &overflow) : extend_b(m, st, n, &overflow);
      }
    }
    if (!ok) { atomicExch(err, row); if (!FILL) ptr[r] = 0; continue; }
    if (!FILL) { ptr[r] = n_c + (has_home ? 1 : 0); continue; }
    int k = ptr[r];
    double* vv[5] = {v0, v1, v2, v3, v4};
    if (has_home) {
      ind[k] = row - 1;
      for (int q = 0; q < NMAT; ++q) vv[q][k] = Ni[q];
      ++k;
    }
    for (int i = 0; i < n_c; ++i, ++k) {
      ind[k] = st[i] - 1;
      for (int q = 0; q < NMAT; ++q) vv[q][k] = Nc[q * n_c + i];
    }
  }
}

template <int G>
int build_group(Ctx* c, Mesh& M, const MV& mv, int first_op, int nmat, const Range& rows, const Range& cols, int* d_err) {
  Mesh::OpGroup& g = M.grp[G];
  g.nmat = nmat; g.first_op = first_op;
  const int m_loc = rows.n_loc();
  UFE_TRY(g.ptr.alloc((size_t)m_loc + 1));
  UFE_CUDA(cudaMemsetAsync(g.ptr.p, 0, sizeof(int) * ((size_t)m_loc + 1), c->stream));
  const int grid = grid_for(c, m_loc, 128, 16);
  {
    UFE_LAUNCH(c, TC_OPERATOR_SPMV);
    k_build_group<G, false><<<grid, 128, 0, c->stream>>>(mv, rows.i1, m_loc, g.ptr.p, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, d_err);
  }
  UFE_CUDA(cudaGetLastError());
  // exclusive scan of the row lengths -> offsets (ptr[m_loc] = nnz)
  size_t tmp_bytes = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, g.ptr.p, g.ptr.p, m_loc + 1, c->stream);
  DBuf<char> tmp;
  UFE_TRY(tmp.alloc(tmp_bytes));
  c->launches++;
  UFE_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, tmp_bytes, g.ptr.p, g.ptr.p, m_loc + 1, c->stream));
  int nnz = 0;
  UFE_CUDA(cudaMemcpyAsync(&nnz, g.ptr.p + m_loc, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  g.nnz = nnz;
  UFE_TRY(g.ind.alloc(nnz));
  for (int q = 0; q < nmat; ++q) UFE_TRY(g.val[q].alloc(nnz));
  {
    UFE_LAUNCH(c, TC_OPERATOR_SPMV);
    k_build_group<G, true><<<grid, 128, 0, c->stream>>>(mv, rows.i1, m_loc, g.ptr.p, g.ind.p, g.val[0].p, g.val[1].p, g.val[2].p,
                                                        g.val[3].p, g.val[4].p, d_err);
  }
  UFE_CUDA(cudaGetLastError());
  for (int q = 0; q < nmat; ++q) {
    DevCsr& A = M.op[first_op + q];
    A = DevCsr();
    A.m_loc = m_loc; A.nnz = nnz; A.rows = rows; A.cols = cols;
    A.ptr = g.ptr.p; A.ind = g.ind.p; A.val = g.val[q].p;
    const double avg = m_loc > 0 ? (double)nnz / m_loc : 0.0;
    A.vec_width = avg <= 2.5 ? 2 : avg <= 6.0 ? 4 : avg <= 14.0 ? 8 : 16;
    A.valid = true;
  }
  return 0;
}

}  // namespace

int mesh_calc_operators(Ctx* c) {
  Mesh& M = *c->mesh;
  MV mv{M.nV, M.nTri, M.nC_mem, M.V.p, M.V.p + M.nV, M.TriGC.p, M.TriGC.p + M.nTri, M.nC.p, M.C.p, M.niTri.p, M.iTri.p, M.Tri.p, M.TriC.p};
  DBuf<int> err;
  UFE_TRY(err.alloc(1));
  UFE_CUDA(cudaMemsetAsync(err.p, 0, sizeof(int), c->stream));
  // during construction columns are global (lo = 1); mesh_finalize_ranges() narrows them to the extended ranges
  Range ga = M.a, gb = M.b;
  ga.lo = 1; ga.hi = M.nV; gb.lo = 1; gb.hi = M.nTri;
  UFE_TRY((build_group<Mesh::G_AA>(c, M, mv, UFE_OP_DDX_A_A, 2, M.a, ga, err.p)));
  UFE_TRY((build_group<Mesh::G_AB>(c, M, mv, UFE_OP_MAP_A_B, 3, M.b, ga, err.p)));
  UFE_TRY((build_group<Mesh::G_BA>(c, M, mv, UFE_OP_MAP_B_A, 3, M.a, gb, err.p)));
  UFE_TRY((build_group<Mesh::G_BB>(c, M, mv, UFE_OP_DDX_B_B, 2, M.b, gb, err.p)));
  UFE_TRY((build_group<Mesh::G_M2>(c, M, mv, UFE_OP2_DDX_B_B, 5, M.b, gb, err.p)));
  int h_err = 0;
  UFE_CUDA(cudaMemcpyAsync(&h_err, err.p, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  UFE_CHECK(h_err == 0, "calc_matrix_operators: could not build shape functions for row %d (expanded local neighbourhood too far)", h_err);
  M.have_ops = true;
  return mesh_finalize_ranges(c);
}

}  // namespace ufe
