// Shared declarations of libufe_b200.so (sm_100a only; no CPU fallback anywhere in this directory).
#pragma once
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "../../include/ufe_b200.h"

namespace ufe {

// ---- errors ------------------------------------------------------------------------------------
void set_error(const char* fmt, ...);
const char* get_error();

#define UFE_CUDA(call)                                                                              \
  do {                                                                                              \
    cudaError_t e__ = (call);                                                                       \
    if (e__ != cudaSuccess) {                                                                       \
      ufe::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__));        \
      return 1;                                                                                     \
    }                                                                                               \
  } while (0)

#define UFE_CHECK(cond, ...)                                                                        \
  do {                                                                                              \
    if (!(cond)) { ufe::set_error(__VA_ARGS__); return 2; }                                         \
  } while (0)

#define UFE_TRY(expr)                                                                               \
  do { int rc__ = (expr); if (rc__ != 0) return rc__; } while (0)

// ---- device buffers ------------------------------------------------------------------------------
template <typename T>
struct DBuf {
  T* p = nullptr;
  size_t n = 0;
  int alloc(size_t count) {
    if (count <= n && p) return 0;
    release();
    if (count == 0) return 0;
    cudaError_t e = cudaMalloc((void**)&p, count * sizeof(T));
    if (e != cudaSuccess) { set_error("cudaMalloc(%zu bytes) failed: %s", count * sizeof(T), cudaGetErrorString(e)); p = nullptr; return 1; }
    n = count;
    return 0;
  }
  void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
  ~DBuf() { release(); }
  DBuf() = default;
  DBuf(const DBuf&) = delete;
  DBuf& operator=(const DBuf&) = delete;
  DBuf(DBuf&& o) noexcept : p(o.p), n(o.n) { o.p = nullptr; o.n = 0; }
  DBuf& operator=(DBuf&& o) noexcept { if (this != &o) { release(); p = o.p; n = o.n; o.p = nullptr; o.n = 0; } return *this; }
};

// ---- index ranges ---------------------------------------------------------------------------------
// One grid (a = vertices, b = triangles, c = edges, buv = interleaved triangle u/v) on one rank.
// Owned global indices i1..i2 (1-based); device vectors cover the extended range lo..hi (owned + halos),
// element g lives at [g - lo].
struct Range {
  int n_glob = 0, i1 = 1, i2 = 0, lo = 1, hi = 0;
  int send_l = 0, send_r = 0;  // how many of my first / last owned elements the left / right neighbour's halo needs
  int n_loc() const { return i2 - i1 + 1; }
  int n_ext() const { return hi - lo + 1; }
  int off() const { return i1 - lo; }  // position of the first owned element in an extended vector
};

// ---- CSR matrix on the device -----------------------------------------------------------------------
// Rows = the rank's owned rows (local 0..m_loc-1); ptr 0-based; ind = 0-based position in the *extended*
// column vector (global column - col.lo).  The reference's construction order of entries is preserved.
struct DevCsr {
  int m_loc = 0;
  long long nnz = 0;
  Range rows, cols;
  int* ptr = nullptr;  // [m_loc+1]   (may alias another matrix of the same operator group)
  int* ind = nullptr;  // [nnz]
  double* val = nullptr;
  DBuf<int> ptr_buf, ind_buf;   // owning storage (empty when the pattern is shared)
  DBuf<double> val_buf;
  int vec_width = 4;   // lanes per row of the vector-per-row kernel (1,2,4,8,16,32)
  bool valid = false;
  // Sliced-ELL copy (slices of 32 rows, entries of a slice stored column-major so that a warp with one row per lane reads val/ind in full
  // 256 B / 128 B segments; csr.cu: csr_sell_build / csr_sell_refresh).  Built on demand for matrices that are multiplied many times (Krylov
  // solves, registered matrices, mesh operators); empty otherwise.  sell_slice[s] = first slot of slice s (a multiple of 32).
  DBuf<int> sell_slice, sell_ind;
  DBuf<double> sell_val;
  const int* sell_src_ptr = nullptr;   // the ptr array the pattern was built from (a rebuilt CSR pattern invalidates the copy)
  long long sell_src_nnz = -1;
  bool has_sell() const { return sell_val.p != nullptr && sell_src_ptr == ptr && sell_src_nnz == nnz; }
};

// ---- timing classes ----------------------------------------------------------------------------------
enum TimingClass { TC_SPMV_STIFFNESS = 0, TC_KRYLOV_VEC, TC_ASSEMBLY, TC_COLUMN, TC_TRI, TC_EPILOGUE, TC_THICKNESS, TC_OPERATOR_SPMV,
                   TC_MG_FINE, TC_MG_VEC, TC_MG_COARSE, TC_MG_SETUP, TC_H2D, TC_D2H, TC_COUNT };
const char* timing_class_name(int c);

struct Mesh;      // mesh.cuh
struct Velocity;  // velocity.cuh
struct Krylov;    // krylov.cuh
struct Thickness; // thickness.cuh
struct Secondary; // secondary.cu
void secondary_free(Secondary* s);
struct Geometry;  // geometry.cu
void geometry_free(Geometry* g);

// multigrid preconditioner knobs (mg.cuh); set through ufe_b200_set_mg_params
struct MgParams {
  int nu_pre = 1, nu_post = 1;   // block-Jacobi sweeps before / after the coarse-grid correction
  double omega = 0.8;            // damping
  double overcorr = 1.5;         // scaling of the coarse-grid correction (plain aggregation under-corrects; > 1.7 is not robust)
  int adaptive = 0;              // 1: on level 0 the correction is scaled by the factor that minimises the residual instead of `overcorr` (mg.cu: k_mg_aq)
  double adaptive_max = 4.0;     // ... kept inside [0, adaptive_max]
  int w_levels = 3;              // coarse levels 1..w_levels are visited twice per cycle (W-cycle on top, V-cycle below)
  int dense_max = 128;           // coarsest level: at most this many nodes -> dense inverse; otherwise `coarse_sweeps` sweeps
  int coarse_sweeps = 8;
  int coarsest_cluster = 0;      // > 1: the sweeps on the coarsest level run as one kernel of this many CTAs in one thread-block cluster (mg.cu: k_mg_coarsest_cluster)
  int coarsest_cheb = 0;         // 1: the `coarse_sweeps` corrections of that kernel are Chebyshev steps for [lmax / cheb_ratio, lmax] of Dw A instead of stationary sweeps
  double cheb_ratio = 30.0;
  int coarsest_threads = 256;    // threads per CTA of that kernel (a multiple of 32, at most 256)
  int coarsest_dsmem = 0;        // ... with the iterate exchanged through L2 (0) or held in distributed shared memory (1); measured equal (2 us per sweep either way)
  int stop_nodes = 0;            // > 0: coarsening stops at the first level with at most this many nodes; above dense_max that level is solved by `coarse_sweeps` sweeps
  int tail_max = 0;              // > 0: levels from the first one with <= tail_max nodes down run as one single-CTA kernel per visit (k_mg_tail).
                                 // Measured slower than graph nodes on B200 at 4 M vertices (0.82 vs 0.72 ms per cycle with 2048, 1.98 ms with
                                 // 16384: one SM's L2 bandwidth and latency cost more than ~80 kernel boundaries save;
                                 // profiles/r01_mg_probe_tail.log), so off by default; UFE_B200_MG_TAIL=<nodes> enables it
  int tail_cluster = 0;          // > 0: the tail kernel runs as one thread-block cluster of this many CTAs (cluster barrier between operations)
  int max_levels = 12;
  int fuse = 0;                  // 1: fuse neighbouring operations on the coarse levels (3 launches per level visit instead of 5); measured no
                                 // gain on B200 (the fused kernels have longer dependent-load chains: profiles/r01_mg_probe_fuse.log)
  int double_from = 0;           // > 0: transfers from this level on aggregate twice (ratio ~60): fewer latency-bound coarse levels
  int fp32 = 0;                  // 1: matrix values of the preconditioner stored as float
  int use_graph = 1;             // coarse-grid correction (levels >= 1): 0 plain launches, 1 one CUDA graph (default), 2 one persistent
                                 // cooperative kernel (measured slower than the graph on B200: profiles/r01_mg_probe_modes.log)
};

struct PcState;                      // timestep.cu
void pc_state_free(PcState*);

struct Ctx {
  int device = 0, rank = 0, nranks = 1;
  cudaStream_t stream = nullptr;
  DBuf<double> xT;                      // column-interleaved copy of the right-hand sides of a multi-column product (csr.cu)
  DBuf<unsigned char> l2_flush;         // ufe_b200_l2_flush: a buffer larger than L2, written between timed iterations of small workloads
  cudaStream_t copy_stream = nullptr;   // host<->device transfers that overlap compute (ufe_b200_pc_step)
  cudaEvent_t ev_copy_a = nullptr, ev_copy_b = nullptr;
  void* nccl = nullptr;             // ncclComm_t when nranks > 1
  long long launches = 0;
  int sm_count = 148;
  // timing
  bool timing = false;
  int timing_level = 1;             // 1: every launch; 2: only the host<->device transfer regions (TC_H2D / TC_D2H)
  int timing_suppress = 0;          // > 0: launches are counted but not individually timed (inside a timed region / graph capture)
  struct Ev { cudaEvent_t a, b; int cls; };
  std::vector<Ev> ev_pool; size_t ev_used = 0;
  double tc_ms[TC_COUNT] = {0}; long long tc_n[TC_COUNT] = {0};
  // registered generic matrices (narrow seam)
  std::map<int, DevCsr*> csr;
  Mesh* mesh = nullptr;
  Velocity* vel = nullptr;
  Krylov* krylov = nullptr;
  Thickness* thick = nullptr;
  Secondary* sec = nullptr;
  Geometry* geo = nullptr;
  PcState* pc = nullptr;         // state of the predictor-corrector stepper between calls (released with the mesh)
  cudaEvent_t t0 = nullptr, t1 = nullptr;   // ufe_b200_timer_start / _stop
  unsigned long long bc_version = 0;        // bumped whenever the prescribed-velocity mask of the stiffness pattern changes
  int state_resident = 0;        // ufe_b200_set_state_resident: ufe_b200_pc_step keeps the solver state on the device between calls
  MgParams mg_params;
  // pinned scratch for small D2H reads
  double* h_pinned = nullptr;       // 64 doubles
};

// Launch bookkeeping: every kernel launch of the library goes through one of these.
struct LaunchTimer {
  Ctx* c; int cls; int idx = -1;
  LaunchTimer(Ctx* c_, int cls_, bool count = true);   // count = false: a timed region around launches that count themselves
  ~LaunchTimer();
};
#define UFE_LAUNCH(ctx, cls) ufe::LaunchTimer lt__((ctx), (cls))
#define UFE_REGION(ctx, cls) ufe::LaunchTimer region__((ctx), (cls), false)
void timing_collect(Ctx* c);

inline int grid_for(const Ctx* c, long long n, int block, int max_waves = 8) {
  long long g = (n + block - 1) / block;
  long long cap = (long long)c->sm_count * max_waves;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}

// halo exchange of an extended vector (no-op on one rank); ncols vectors of stride r.n_ext()
int halo_exchange(Ctx* c, const Range& r, double* v, int ncols = 1, int elems_per_index = 1);
int halo_exchange_int(Ctx* c, const Range& r, int* v);
int allreduce_sum(Ctx* c, double* d_vals, int n);
int allreduce_min(Ctx* c, double* d_vals, int n);
int allreduce_sum_large(Ctx* c, void* d_vals, size_t n, bool is_float);   // in place, double or float
// all-gather of variable-length int lists through device buffers (setup paths only): counts[r], concatenation in rank order
int allgatherv_int(Ctx* c, const std::vector<int>& mine, std::vector<int>& counts, std::vector<int>& all);

}  // namespace ufe
