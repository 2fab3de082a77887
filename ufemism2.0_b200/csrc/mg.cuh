// Aggregation multigrid preconditioner for the SSA/DIVA stiffness matrix (pc_type = UFE_PC_MG).
//
// Why: the reference hands the linearised system to PETSc (solve_matrix_equation_CSR_PETSc, src/UPSY/basic/petsc_basic.f90:32-242),
// whose default one-level preconditioners (block-Jacobi/ILU(0), or Jacobi as north_star prescribes) need O(1/h) Krylov iterations:
// ~270 at 60 k vertices, ~610 at 250 k and more than the reference's 2000-iteration cap (:166) at 4 M.  A multilevel
// preconditioner keeps the count nearly mesh-independent, so the solve stays a few dozen streamed passes over the matrix.
//
// What: plain (unsmoothed) aggregation on the graph of 2x2 triangle blocks.
//   * hierarchy (pattern) built once per (mesh, BC) on the host from the blocked sliced-ELL pattern: greedy neighbourhood
//     aggregation (root + all free neighbours, leftovers join the most common neighbouring aggregate), coarse patterns in the same
//     sliced-ELL layout, and per coarse block the list of fine blocks it sums ("Galerkin gather list");
//   * values rebuilt on the device every viscosity iteration, deterministically (no atomics): A_c = P^T A P with piecewise-constant
//     P per velocity component, i.e. coarse block (I,J) = sum of the fine blocks (i,j), i in I, j in J;
//   * cycle: damped 2x2-block-Jacobi smoothing, V-cycle with an optional second coarse visit on the first `w_levels` coarse levels,
//     dense inverse (Gauss-Jordan on the device) on the coarsest level;
//   * optional FP32 storage of the matrix values inside the preconditioner (the outer BiCGStab, its residuals and the convergence
//     test stay FP64, so the solution meets the same tolerance).
// Multi-GPU: level 0 is the rank's row block (smoothing sweeps exchange halos like the Krylov SpMV); aggregates never cross ranks, but the
// coarse levels are global: every rank coarsens its own block, the coarse block rows are all-gathered once per pattern, and levels >= 1 are
// replicated -- each rank restricts its residual into its slice of the level-1 right-hand side, one all-reduce completes it, and every rank
// runs the same (latency-bound, graph-replayed) coarse correction redundantly.  Level-1 values are completed by one all-reduce per
// viscosity iteration.  The iteration count therefore does not depend on the number of ranks beyond the aggregate boundaries.
#pragma once
#include "blk.cuh"
#include "common.cuh"

namespace ufe {

// one step of the recorded coarse-level program (mg.cu: k_mg_program)
enum MgOp { MG_OP_SMOOTH0 = 0, MG_OP_SWEEP = 1, MG_OP_RESID = 2, MG_OP_RESTRICT = 3, MG_OP_PROLONG = 4, MG_OP_AXPY = 5, MG_OP_DENSE = 6 };
struct MgStep {
  int op, n, off, pad;
  const int* ip0;      // slice_ptr | mem_ptr | agg
  const int* ip1;      // col | mem
  const void* val;     // matrix values | dense inverse
  const double* x;     // iterate (extended) | residual to restrict | coarse solution | axpy operand
  const double* b;     // right-hand side
  const double* d;     // inverse diagonal blocks
  double* out;
  double w;
};


struct MgLevel {
  int n = 0, n_ext = 0, off = 0;          // nodes, size of the extended vector (level 0: owned + halo triangles), offset of node 0
  int n_slices = 0;
  long long n_slots = 0;
  const int* slice_ptr = nullptr;         // views of the pattern (level 0: the Krylov solver's BlkMat)
  const int* col = nullptr;
  const double* val = nullptr;            // FP64 values (level 0: BlkMat::val)
  DBuf<int> slice_ptr_buf, col_buf;
  DBuf<double> val_buf;
  DBuf<float> valf;                       // FP32 copy (fp32 = 1)
  DBuf<double> dblk;                      // [n][4] omega * inverse of the diagonal 2x2 block
  // transfer to the next coarser level
  int nc = 0;                             // nodes of the next level (all ranks)
  int nc_loc = 0, c_offset = 0;           // this rank's aggregates are the coarse nodes c_offset .. c_offset + nc_loc - 1 (level 0 only; else all)
  DBuf<int> agg;                          // [n] coarse node (global id) of each node
  DBuf<int> mem_ptr, mem;                 // coarse node -> member nodes
  DBuf<int> gal_ptr, gal_src;             // coarse block slot -> fine block slots
  // work vectors (2 doubles per node)
  DBuf<double> xa, xb;                    // ping-pong iterates over the extended range
  DBuf<double> b, b2, r, sol, sol2;       // restricted right-hand sides, residual, coarse solutions (levels >= 1)
};

struct Mg {
  ~Mg() { drop_graph(); }
  void drop_graph() {
    if (graph_exec) cudaGraphExecDestroy(graph_exec);
    if (graph) cudaGraphDestroy(graph);
    graph_exec = nullptr; graph = nullptr; graph_launches = 0; program_n = 0; tail_len[0] = tail_len[1] = 0;
  }
  std::vector<MgLevel> lv;
  DBuf<unsigned long long> lmax_bits;     // bits of the bound of the spectral radius of Dw A on the coarsest level (Chebyshev steps)
  size_t coarsest_smem = 0;               // shared memory per CTA of k_mg_coarsest_cluster for the last level (0: does not fit / not wanted)
  DBuf<double> dense_w, dense_inv;        // Gauss-Jordan workspace [N][2N], inverse (column-major) [N][N]
  int dense_n = 0;                        // unknowns of the dense coarsest solve (0: smoothing only)
  MgParams p;
  bool distributed = false;               // nranks > 1: level 0 is this rank's row block, levels >= 1 are replicated on every rank
  Range rb;                               // triangle range of level 0 (halo exchange of the level-0 iterates)
  // the coarse-grid correction (every launch on levels >= 1) captured once as a CUDA graph and replayed per cycle
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t graph_exec = nullptr;
  long long graph_launches = 0;
  // ... or, use_graph = 2, recorded as a step list that one persistent cooperative kernel walks (mg.cu: k_mg_program)
  std::vector<MgStep>* rec = nullptr;   // non-null while recording
  DBuf<char> program;
  int program_n = 0, program_grid = 0;
  // the levels from tail_level down run as one single-CTA kernel per visit (mg.cu: k_mg_tail); two step lists: (b, sol) and (b2, sol2)
  int tail_level = 0, tail_off[2] = {0, 0}, tail_len[2] = {0, 0};
  DBuf<char> tail_prog;
  DBuf<double> mr_q, mr_scal, mr_partials;   // level-0 correction with its own step length (MgParams::adaptive)
  DBuf<unsigned> mr_counter;
  bool valid = false;
  long long gather_entries = 0;
  double setup_host_s = 0.0;
};

// pattern phase (once per stiffness pattern); B is the level-0 matrix the Krylov solver streams
int mg_build(Ctx* c, Mg& G, const BlkMat& B, const Range& rb);
// numeric phase (every viscosity iteration, after blk_update_values)
int mg_update_values(Ctx* c, Mg& G, const BlkMat& B, const DevCsr& A);
// z = M^-1 r for owned rows (r, z: 2 doubles per owned triangle)
int mg_apply(Ctx* c, Mg& G, const double* r, double* z);
// left-preconditioned BiCGStab (same recurrence, stopping rule and scalars as bicgstab_solve_mat) with M^-1 = one multigrid cycle
int bicgstab_solve_mg(Ctx* c, Mg& G, const BlkLauncher& A, const Range& rr, const double* d_b, double* d_x_ext, const KrylovParams& kp, int* its);

}  // namespace ufe
