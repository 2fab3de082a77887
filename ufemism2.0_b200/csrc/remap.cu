// Remapping of the velocity solver's state after a mesh update (the SpMV part):
//   remap_DIVA_solver( mesh_old, mesh_new, DIVA)   src/UFEMISM/ice_dynamics/conservation_of_momentum/SSA_DIVA/DIVA_main.f90:115-222
//   remap_SSA_solver ( mesh_old, mesh_new, SSA)    .../SSA_DIVA/SSA_main.f90:247-311
//   apply_map_mesh_to_mesh_{2D,3D}                 src/UPSY/mesh/remapping/apply_maps.f90:531-583  (MatMult with the Atlas map, then
//   set_border_vertices_to_interior_mean_dp_{2D,3D} src/UPSY/mesh/mesh_utilities.f90:1648-1757)
// Every re-used field goes  b-grid(old) --map_b_a(old)--> a-grid(old) --M_map--> a-grid(new) --border fix--> --map_a_b(new)--> b-grid(new);
// everything else of the solver object is re-allocated (zero here), u_b_prev / v_b_prev are cleaned.  The map matrix itself (conservative
// remapping: line tracing through both meshes, create_map_from_mesh_to_mesh_2nd_order_conservative) is mesh generation, out of scope: the host
// passes it as the rank's rows of a CSR matrix, exactly what the reference keeps in its Atlas (type_map%M).
#include "csr.cuh"
#include "mesh.cuh"
#include "velocity.cuh"

namespace ufe {
namespace {

// out := in, then on border vertices the mean of the interior neighbours' OLD values (or, without interior neighbours, of all neighbours' old values):
// both passes of the reference read the gathered copy d_tot, so one out-of-place kernel reproduces them.  Fields are [ncols][n_ext].
__global__ void k_border_mean(int n_loc, int off, int lo, int n_ext, int ncols, int nV, const int* __restrict__ nC, const int* __restrict__ Cn,
                              const int* __restrict__ VBI, int vi1, const double* __restrict__ in, double* __restrict__ out) {
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < (long long)n_loc * ncols; t += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(t % n_loc), col = (int)(t / n_loc);
    const int vi = vi1 + r;                          // 1-based global vertex
    const double* f = in + (size_t)col * n_ext;
    double val = f[off + r];
    if (VBI[vi - 1] > 0) {
      int n_int = 0; double s_int = 0.0, s_all = 0.0;
      const int nc = nC[vi - 1];
      for (int ci = 0; ci < nc; ++ci) {
        const int vj = Cn[(size_t)ci * nV + (vi - 1)];   // C( vi, ci), column-major
        const double d = f[vj - lo];
        s_all += d;
        if (VBI[vj - 1] == 0) { ++n_int; s_int += d; }
      }
      val = n_int > 0 ? s_int / (double)n_int : s_all / (double)nc;
    }
    out[(size_t)col * n_ext + off + r] = val;
  }
}

struct RemapWork {
  DBuf<double> a_old, full_old, a_new, a_new2, b_new;
};

// one field: src = [ncols][bo.n_ext] on the old handle, dst = [ncols][bn.n_ext] on the new one
int remap_field(Ctx* co, Ctx* cn, const DevCsr& map, RemapWork& W, const double* src, double* dst, int ncols) {
  Mesh& Mo = *co->mesh; Mesh& Mn = *cn->mesh;
  const int nao = Mo.a.n_loc(), nan = Mn.a.n_loc(), nbn = Mn.b.n_loc();
  // (1) triangles -> vertices on the old mesh (map_b_a_{2D,3D})
  UFE_TRY(halo_exchange(co, Mo.b, const_cast<double*>(src), ncols));
  UFE_TRY(csr_spmv_multicol(co, Mo.op[UFE_OP_MAP_B_A], src, W.a_old.p, ncols, TC_OPERATOR_SPMV));
  // (2) the whole old a-grid field on every rank (the map's columns are global: the new row blocks do not follow the old ones)
  if (co->nranks > 1) {
    UFE_CUDA(cudaMemsetAsync(W.full_old.p, 0, sizeof(double) * (size_t)Mo.nV * ncols, co->stream));
    UFE_CUDA(cudaMemcpy2DAsync(W.full_old.p + (Mo.a.i1 - 1), sizeof(double) * Mo.nV, W.a_old.p, sizeof(double) * nao, sizeof(double) * nao, ncols,
                               cudaMemcpyDeviceToDevice, co->stream));
    UFE_TRY(allreduce_sum_large(co, W.full_old.p, (size_t)Mo.nV * ncols, false));   // disjoint supports: a gather, exact
  }
  UFE_CUDA(cudaStreamSynchronize(co->stream));      // the two handles own different streams
  const double* x = co->nranks > 1 ? W.full_old.p : W.a_old.p;
  // (3) M_map (multiply_PETSc_matrix_with_vector_{1D,2D}); result in the new mesh's extended a-layout
  UFE_TRY(csr_spmv_multicol(cn, map, x, W.b_new.p, ncols, TC_OPERATOR_SPMV));          // W.b_new doubles as the [ncols][nan] scratch (nbn >= nan)
  UFE_CUDA(cudaMemcpy2DAsync(W.a_new.p + Mn.a.off(), sizeof(double) * Mn.a.n_ext(), W.b_new.p, sizeof(double) * nan, sizeof(double) * nan, ncols,
                             cudaMemcpyDeviceToDevice, cn->stream));
  UFE_TRY(halo_exchange(cn, Mn.a, W.a_new.p, ncols));
  // (4) set_border_vertices_to_interior_mean_dp
  {
    UFE_LAUNCH(cn, TC_OPERATOR_SPMV);
    k_border_mean<<<grid_for(cn, (long long)nan * ncols, 256), 256, 0, cn->stream>>>(nan, Mn.a.off(), Mn.a.lo, Mn.a.n_ext(), ncols, Mn.nV, Mn.nC.p, Mn.C.p,
                                                                                    Mn.VBI.p, Mn.a.i1, W.a_new.p, W.a_new2.p);
  }
  UFE_CUDA(cudaGetLastError());
  UFE_TRY(halo_exchange(cn, Mn.a, W.a_new2.p, ncols));
  // (5) vertices -> triangles on the new mesh (map_a_b_{2D,3D}), into the extended b-layout of the solver state
  UFE_TRY(csr_spmv_multicol(cn, Mn.op[UFE_OP_MAP_A_B], W.a_new2.p, W.b_new.p, ncols, TC_OPERATOR_SPMV));
  UFE_CUDA(cudaMemcpy2DAsync(dst + Mn.b.off(), sizeof(double) * Mn.b.n_ext(), W.b_new.p, sizeof(double) * nbn, sizeof(double) * nbn, ncols,
                             cudaMemcpyDeviceToDevice, cn->stream));
  return 0;
}

}  // namespace
}  // namespace ufe

using namespace ufe;

extern "C" int ufe_b200_remap_velocity_solver(ufe_b200_handle h_old, ufe_b200_handle h_new, int approximation, int m_loc, const int32_t* map_ptr,
                                              const int32_t* map_ind, const double* map_val) {
  Ctx* co = reinterpret_cast<Ctx*>(h_old); Ctx* cn = reinterpret_cast<Ctx*>(h_new);
  UFE_CHECK(co && cn && co != cn && map_ptr, "ufe_b200_remap_velocity_solver: two different handles and the map are required");
  UFE_CHECK(approximation == 0 || approximation == 1, "ufe_b200_remap_velocity_solver: approximation must be 0 (SSA) or 1 (DIVA)");
  UFE_CHECK(co->device == cn->device && co->rank == cn->rank && co->nranks == cn->nranks, "ufe_b200_remap_velocity_solver: the handles must belong to the same rank and GPU");
  UFE_CHECK(co->mesh && co->mesh->have_ops && cn->mesh && cn->mesh->have_ops, "ufe_b200_remap_velocity_solver: both meshes need their operators");
  UFE_CHECK(co->vel && co->vel->uploaded, "ufe_b200_remap_velocity_solver: the old handle holds no velocity state");
  UFE_CUDA(cudaSetDevice(cn->device));
  Mesh& Mo = *co->mesh; Mesh& Mn = *cn->mesh;
  UFE_CHECK(Mo.nz == Mn.nz, "ufe_b200_remap_velocity_solver: the meshes have different numbers of layers (%d, %d)", Mo.nz, Mn.nz);
  UFE_CHECK(m_loc == Mn.a.n_loc(), "ufe_b200_remap_velocity_solver: the map must hold the rank's %d new vertices as rows (got %d)", Mn.a.n_loc(), m_loc);
  UFE_TRY(ensure_velocity(cn));
  Velocity& Vo = *co->vel; Velocity& Vn = *cn->vel;
  const int nz = Mn.nz;
  // the Atlas map: rows vi1:vi2 of the new mesh, columns = all vertices of the old one
  DevCsr map;
  Range all_old; all_old.n_glob = Mo.nV; all_old.i1 = 1; all_old.i2 = Mo.nV; all_old.lo = 1; all_old.hi = Mo.nV;
  UFE_TRY(csr_upload(cn, map, Mn.nV, Mo.nV, m_loc, Mn.a.i1, Mo.nV, 1, map_ptr, map_ind, map_val, false, &all_old));
  UFE_TRY(csr_sell_build(cn, map));
  if (!Mo.op[UFE_OP_MAP_B_A].has_sell()) UFE_TRY(csr_sell_build(co, Mo.op[UFE_OP_MAP_B_A]));
  if (!Mn.op[UFE_OP_MAP_A_B].has_sell()) UFE_TRY(csr_sell_build(cn, Mn.op[UFE_OP_MAP_A_B]));
  RemapWork W;
  const size_t cols = (size_t)nz;
  UFE_TRY(W.a_old.alloc((size_t)Mo.a.n_loc() * cols));
  if (co->nranks > 1) UFE_TRY(W.full_old.alloc((size_t)Mo.nV * cols));
  UFE_TRY(W.a_new.alloc((size_t)Mn.a.n_ext() * cols)); UFE_TRY(W.a_new2.alloc((size_t)Mn.a.n_ext() * cols));
  UFE_TRY(W.b_new.alloc((size_t)std::max(Mn.b.n_loc(), Mn.a.n_loc()) * cols));
  UFE_CUDA(cudaMemsetAsync(W.a_new.p, 0, sizeof(double) * W.a_new.n, cn->stream));
  UFE_CUDA(cudaMemsetAsync(W.a_new2.p, 0, sizeof(double) * W.a_new2.n, cn->stream));
  // everything that is only re-allocated in the reference starts from zero; u_b_prev / v_b_prev are cleaned (reallocate_clean)
  for (DBuf<double>* f : {&Vn.u, &Vn.v, &Vn.u_base, &Vn.v_base, &Vn.tau_bx, &Vn.tau_by, &Vn.eta_3D_b, &Vn.u_prev, &Vn.v_prev, &Vn.tau_dx, &Vn.tau_dy,
                          &Vn.u_3D, &Vn.v_3D})
    if (f->p) UFE_CUDA(cudaMemsetAsync(f->p, 0, sizeof(double) * f->n, cn->stream));
  UFE_TRY(remap_field(co, cn, map, W, Vo.u.p, Vn.u.p, 1));
  UFE_TRY(remap_field(co, cn, map, W, Vo.v.p, Vn.v.p, 1));
  if (approximation == 1) {
    UFE_TRY(remap_field(co, cn, map, W, Vo.tau_bx.p, Vn.tau_bx.p, 1));
    UFE_TRY(remap_field(co, cn, map, W, Vo.tau_by.p, Vn.tau_by.p, 1));
    UFE_TRY(remap_field(co, cn, map, W, Vo.eta_3D_b.p, Vn.eta_3D_b.p, nz));
    UFE_TRY(remap_field(co, cn, map, W, Vo.u_3D.p, Vn.u_3D.p, nz));
    UFE_TRY(remap_field(co, cn, map, W, Vo.v_3D.p, Vn.v_3D.p, nz));
  }
  UFE_CUDA(cudaStreamSynchronize(cn->stream));
  Vn.uploaded = true;           // ice inputs still have to follow (ufe_b200_velocity_upload with a NULL state)
  Vn.pattern_valid = false;
  return 0;
}
