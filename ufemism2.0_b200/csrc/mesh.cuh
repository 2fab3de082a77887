// Device mirror of the members of type_mesh the hot path reads (src/UPSY/types/mesh_types.f90:40-232).
#pragma once
#include "common.cuh"

namespace ufe {

// Mesh arrays keep the reference's layout (column-major, 1-based values, global size) so that the kernels read
// exactly what the Fortran host owns; fields are distributed (Range a/b per rank), mesh arrays are global.
struct Mesh {
  int nV = 0, nTri = 0, nE = 0, nC_mem = 0, nz = 0;
  double xmin = 0, xmax = 0, ymin = 0, ymax = 0;
  DBuf<double> V, TriGC, Tricc, A, Cw, D_x, D_y, D, zeta;
  DBuf<int> nC, C, niTri, iTri, VBI, Tri, TriC, TriBI, VE, ETri;
  std::vector<double> h_zeta;
  Range a, b, c, buv;     // vertices, triangles, edges, interleaved (u,v) unknowns on triangles
  // The 15 operators come in 5 groups that share one sparsity pattern each (built by the same flood fill):
  //   AA: ddx,ddy a->a | AB: map,ddx,ddy a->b | BA: map,ddx,ddy b->a | BB: ddx,ddy b->b | M2: ddx,ddy,d2dx2,d2dxdy,d2dy2 b->b
  struct OpGroup {
    int nmat = 0, first_op = 0;
    long long nnz = 0;
    DBuf<int> ptr, ind;
    DBuf<double> val[5];
  };
  enum { G_AA = 0, G_AB, G_BA, G_BB, G_M2, G_COUNT };
  OpGroup grp[G_COUNT];
  DevCsr op[UFE_OP_COUNT];   // views into the groups (or own storage when set through ufe_b200_set_operator)
  bool have_ops = false;
};

int mesh_finalize_ranges(Ctx* c);   // after the operators exist: extended ranges + halo plans
int mesh_build_groups(Ctx* c);

}  // namespace ufe
