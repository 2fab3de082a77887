// Aggregation multigrid preconditioner of the stiffness-matrix solve (see mg.cuh for the design and its reference context).
#include "mg.cuh"

#include <algorithm>
#include <chrono>
#include <cstdlib>

#include <cooperative_groups.h>

#include "krylov_impl.cuh"

namespace ufe {

namespace {

// =====================================================================================================================
// host: hierarchy (pattern) construction
// =====================================================================================================================
struct HostLevel {
  int n = 0, off = 0;
  std::vector<int> slice_ptr, col;   // sliced-ELL block pattern, as BlkMat
};

// adjacency of G u G^T restricted to owned nodes, no self loops, no duplicates
void build_sym_graph(const HostLevel& H, std::vector<size_t>& ptr, std::vector<int>& adj) {
  const int n = H.n;
  std::vector<int> deg(n, 0);
  auto for_each_edge = [&](auto&& f) {
    for (int t = 0; t < n; ++t) {
      const int s = t >> 5, lane = t & 31;
      for (int k = H.slice_ptr[s]; k < H.slice_ptr[s + 1]; ++k) {
        const int cj = H.col[(size_t)k * 32 + lane] - H.off;
        if (cj < 0 || cj >= n || cj == t) continue;
        f(t, cj);
      }
    }
  };
  for_each_edge([&](int i, int j) { deg[i]++; deg[j]++; });
  std::vector<size_t> p0(n + 1, 0);
  for (int i = 0; i < n; ++i) p0[i + 1] = p0[i] + deg[i];
  std::vector<int> raw(p0[n]);
  std::vector<size_t> fill(p0.begin(), p0.end() - 1);
  for_each_edge([&](int i, int j) { raw[fill[i]++] = j; raw[fill[j]++] = i; });
  // dedupe (keeps first occurrence, i.e. out-edges in pattern order first)
  std::vector<int> stamp(n, -1);
  ptr.assign(n + 1, 0);
  adj.clear();
  adj.reserve(raw.size() / 2 + n);
  for (int i = 0; i < n; ++i) {
    for (size_t k = p0[i]; k < p0[i + 1]; ++k) {
      const int j = raw[k];
      if (stamp[j] != i) { stamp[j] = i; adj.push_back(j); }
    }
    ptr[i + 1] = adj.size();
  }
}

// greedy neighbourhood aggregation; returns the number of aggregates
int aggregate(int n, const std::vector<size_t>& ptr, const std::vector<int>& adj, std::vector<int>& agg) {
  agg.assign(n, -1);
  int na = 0;
  // pass 1: a node whose whole neighbourhood is still free becomes a root and takes the neighbourhood
  for (int i = 0; i < n; ++i) {
    if (agg[i] >= 0 || ptr[i + 1] == ptr[i]) continue;
    bool is_free = true;
    for (size_t k = ptr[i]; k < ptr[i + 1]; ++k) if (agg[adj[k]] >= 0) { is_free = false; break; }
    if (!is_free) continue;
    agg[i] = na;
    for (size_t k = ptr[i]; k < ptr[i + 1]; ++k) agg[adj[k]] = na;
    ++na;
  }
  // pass 2: leftovers join the aggregate most of their neighbours belong to (pass-1 state, so no chains form)
  std::vector<int> agg1(agg);
  for (int i = 0; i < n; ++i) {
    if (agg1[i] >= 0) continue;
    int best = -1, best_cnt = 0;
    for (size_t k = ptr[i]; k < ptr[i + 1]; ++k) {
      const int a = agg1[adj[k]];
      if (a < 0 || a == best) continue;
      int cnt = 0;
      for (size_t k2 = ptr[i]; k2 < ptr[i + 1]; ++k2) cnt += (agg1[adj[k2]] == a);
      if (cnt > best_cnt || (cnt == best_cnt && a < best)) { best = a; best_cnt = cnt; }
    }
    if (best >= 0) agg[i] = best;
  }
  // pass 3: anything still free (cannot happen on a symmetric graph, kept for safety) starts an aggregate with its free neighbours
  for (int i = 0; i < n; ++i) {
    if (agg[i] >= 0 || ptr[i + 1] == ptr[i]) continue;
    agg[i] = na;
    for (size_t k = ptr[i]; k < ptr[i + 1]; ++k) if (agg[adj[k]] < 0) agg[adj[k]] = na;
    ++na;
  }
  // pass 4: isolated nodes (rows that only hold their diagonal and that nobody references) are grouped 8 at a time
  int in_group = 0;
  for (int i = 0; i < n; ++i) {
    if (agg[i] >= 0) continue;
    if (in_group == 0) ++na;
    agg[i] = na - 1;
    in_group = (in_group + 1) & 7;
  }
  return na;
}

// ---- one coarsening step, written so that the row blocks of several ranks can be coarsened independently and then assembled into
// ---- one coarse level that every rank holds (single rank: one part) ------------------------------------------------------------

// aggregation of one rank's own nodes: agg = local aggregate ids 0..nc_loc-1, members in ascending fine order
struct LocalAgg {
  int nc_loc = 0;
  std::vector<int> agg, mem_ptr, mem;
};

void aggregate_part(const HostLevel& F, LocalAgg& A) {
  std::vector<size_t> gptr;
  std::vector<int> gadj;
  build_sym_graph(F, gptr, gadj);
  A.nc_loc = aggregate(F.n, gptr, gadj, A.agg);
  A.mem_ptr.assign(A.nc_loc + 1, 0);
  for (int i = 0; i < F.n; ++i) A.mem_ptr[A.agg[i] + 1]++;
  for (int I = 0; I < A.nc_loc; ++I) A.mem_ptr[I + 1] += A.mem_ptr[I];
  A.mem.resize(F.n);
  std::vector<int> fill(A.mem_ptr.begin(), A.mem_ptr.end() - 1);
  for (int i = 0; i < F.n; ++i) A.mem[fill[A.agg[i]]++] = i;
}

// coarse block rows of this part's aggregates with GLOBAL coarse column ids: own aggregate first, then neighbours in order of first
// appearance.  agg_ext[c] = global coarse id of extended node c (own nodes: offset + local id; halo nodes: the neighbour rank's id;
// -1: unknown, the coupling is dropped).  slotpos[fine slot] = position of its coarse block inside the row (-1: dropped).
struct LocalRows {
  std::vector<int> cptr, ccol, slotpos;
};

void build_local_rows(const HostLevel& F, const LocalAgg& A, const std::vector<int>& agg_ext, int offset, int nc_glob, LocalRows& R) {
  R.slotpos.assign(F.col.size(), -1);
  std::vector<int> stamp(nc_glob, -1), posn(nc_glob, 0);
  R.cptr.assign(A.nc_loc + 1, 0);
  R.ccol.clear();
  R.ccol.reserve((size_t)A.nc_loc * 10);
  const int n_ext = (int)agg_ext.size();
  for (int I = 0; I < A.nc_loc; ++I) {
    const int Ig = offset + I;
    const int start = (int)R.ccol.size();
    stamp[Ig] = Ig; posn[Ig] = 0; R.ccol.push_back(Ig);
    for (int m = A.mem_ptr[I]; m < A.mem_ptr[I + 1]; ++m) {
      const int i = A.mem[m], s = i >> 5, lane = i & 31;
      for (int k = F.slice_ptr[s]; k < F.slice_ptr[s + 1]; ++k) {
        const size_t fs = (size_t)k * 32 + lane;
        const int ce = F.col[fs];
        if (ce < 0 || ce >= n_ext) continue;
        const int J = agg_ext[ce];
        if (J < 0) continue;
        if (stamp[J] != Ig) { stamp[J] = Ig; posn[J] = (int)R.ccol.size() - start; R.ccol.push_back(J); }
        R.slotpos[fs] = posn[J];
      }
    }
    R.cptr[I + 1] = (int)R.ccol.size();
  }
}

// sliced-ELL pattern of the assembled coarse level from its rows (cptr/ccol over all nc coarse nodes, global ids)
void ell_from_rows(int nc, const std::vector<int>& cptr, const std::vector<int>& ccol, HostLevel& C) {
  C.n = nc; C.off = 0;
  const int nsl = (nc + 31) / 32;
  C.slice_ptr.assign(nsl + 1, 0);
  for (int s = 0; s < nsl; ++s) {
    int w = 0;
    for (int I = s * 32; I < std::min(nc, s * 32 + 32); ++I) w = std::max(w, cptr[I + 1] - cptr[I]);
    C.slice_ptr[s + 1] = C.slice_ptr[s] + w;
  }
  C.col.assign((size_t)C.slice_ptr[nsl] * 32, 0);
  for (int s = 0; s < nsl; ++s)
    for (int lane = 0; lane < 32; ++lane) {
      const int I = s * 32 + lane;
      const int len = I < nc ? cptr[I + 1] - cptr[I] : 0;
      const int pad = I < nc ? I : 0;
      for (int k = C.slice_ptr[s]; k < C.slice_ptr[s + 1]; ++k) {
        const int j = k - C.slice_ptr[s];
        C.col[(size_t)k * 32 + lane] = j < len ? ccol[cptr[I] + j] : pad;
      }
    }
}

// Galerkin gather lists of this part: coarse slot (of the assembled level) -> this part's fine slots, ascending (deterministic sums)
void build_gather(const HostLevel& F, const LocalAgg& A, int offset, const LocalRows& R, const HostLevel& C, std::vector<int>& gal_ptr,
                  std::vector<int>& gal_src) {
  const size_t coarse_slots = C.col.size();
  gal_ptr.assign(coarse_slots + 1, 0);
  auto coarse_slot = [&](int i, size_t fs) -> size_t {
    const int Ig = offset + A.agg[i];
    return ((size_t)C.slice_ptr[Ig >> 5] + R.slotpos[fs]) * 32 + (Ig & 31);
  };
  const int n = F.n, nslf = (n + 31) / 32;
  for (int i = 0; i < n; ++i) {
    const int s = i >> 5, lane = i & 31;
    for (int k = F.slice_ptr[s]; k < F.slice_ptr[s + 1]; ++k) {
      const size_t fs = (size_t)k * 32 + lane;
      if (R.slotpos[fs] >= 0) gal_ptr[coarse_slot(i, fs) + 1]++;
    }
  }
  for (size_t q = 0; q < coarse_slots; ++q) gal_ptr[q + 1] += gal_ptr[q];
  gal_src.resize((size_t)gal_ptr[coarse_slots]);
  std::vector<int> fill(gal_ptr.begin(), gal_ptr.end() - 1);
  for (int s = 0; s < nslf; ++s)
    for (int k = F.slice_ptr[s]; k < F.slice_ptr[s + 1]; ++k)
      for (int lane = 0; lane < 32; ++lane) {
        const int i = s * 32 + lane;
        if (i >= n) break;
        const size_t fs = (size_t)k * 32 + lane;
        if (R.slotpos[fs] >= 0) gal_src[fill[coarse_slot(i, fs)]++] = (int)fs;
      }
}

struct HostTransfer {
  int nc = 0;                 // coarse nodes of the assembled level (all parts)
  int nc_loc = 0, offset = 0; // this part's aggregates are the coarse nodes offset .. offset + nc_loc - 1
  std::vector<int> agg;       // GLOBAL coarse id of each own node
  std::vector<int> mem_ptr, mem, gal_ptr, gal_src;
  HostLevel coarse;
  long long coarse_slots = 0;
};

// single part (one rank, and every level >= 1): the extended range holds nothing but the own nodes' aggregates.
// passes = 2 aggregates the aggregates once more (coarsening ratio ~60 instead of ~8): fewer, latency-bound coarse levels.
void build_transfer(const HostLevel& F, HostTransfer& T, int passes = 1) {
  LocalAgg A;
  aggregate_part(F, A);
  auto own_agg_ext = [&](const LocalAgg& a) {
    int n_ext = F.off + F.n;
    for (int v : F.col) n_ext = std::max(n_ext, v + 1);
    std::vector<int> e(n_ext, -1);                      // halo triangles (other ranks' rows) are dropped
    for (int i = 0; i < F.n; ++i) e[F.off + i] = a.agg[i];
    return e;
  };
  for (int pass = 1; pass < passes && A.nc_loc > 1; ++pass) {
    // pattern of the intermediate level, aggregated again; the two maps are composed and the intermediate level is dropped
    LocalRows R1;
    HostLevel C1;
    build_local_rows(F, A, own_agg_ext(A), 0, A.nc_loc, R1);
    ell_from_rows(A.nc_loc, R1.cptr, R1.ccol, C1);
    LocalAgg A2;
    aggregate_part(C1, A2);
    for (int i = 0; i < F.n; ++i) A.agg[i] = A2.agg[A.agg[i]];
    A.nc_loc = A2.nc_loc;
    A.mem_ptr.assign(A.nc_loc + 1, 0);
    for (int i = 0; i < F.n; ++i) A.mem_ptr[A.agg[i] + 1]++;
    for (int I = 0; I < A.nc_loc; ++I) A.mem_ptr[I + 1] += A.mem_ptr[I];
    std::vector<int> fill(A.mem_ptr.begin(), A.mem_ptr.end() - 1);
    for (int i = 0; i < F.n; ++i) A.mem[fill[A.agg[i]]++] = i;
  }
  LocalRows R;
  build_local_rows(F, A, own_agg_ext(A), 0, A.nc_loc, R);
  ell_from_rows(A.nc_loc, R.cptr, R.ccol, T.coarse);
  build_gather(F, A, 0, R, T.coarse, T.gal_ptr, T.gal_src);
  T.nc = T.nc_loc = A.nc_loc; T.offset = 0;
  T.coarse_slots = (long long)T.coarse.col.size();
  T.agg = std::move(A.agg); T.mem_ptr = std::move(A.mem_ptr); T.mem = std::move(A.mem);
}

// assembled coarse rows of all parts: lens (one per coarse node, rank order) and the concatenated global column ids
void rows_from_parts(const std::vector<int>& all_lens, std::vector<int>& cptr) {
  cptr.assign(all_lens.size() + 1, 0);
  for (size_t I = 0; I < all_lens.size(); ++I) cptr[I + 1] = cptr[I] + all_lens[I];
}

template <typename T>
int upload(Ctx* c, DBuf<T>& d, const std::vector<T>& h) {
  UFE_TRY(d.alloc(h.size() ? h.size() : 1));
  if (!h.empty()) UFE_CUDA(cudaMemcpyAsync(d.p, h.data(), sizeof(T) * h.size(), cudaMemcpyHostToDevice, c->stream));
  return 0;
}

// level 0 -> 1 when the matrix rows are split over ranks: aggregate the own block, exchange the aggregate ids of the halo triangles, build
// the own coarse rows with global column ids, all-gather the rows and assemble the (replicated) level 1
int build_transfer_distributed(Ctx* c, const HostLevel& F, const Range& rb, HostTransfer& T) {
  LocalAgg A;
  aggregate_part(F, A);
  std::vector<int> counts, dummy;
  UFE_TRY(allgatherv_int(c, std::vector<int>{A.nc_loc}, counts, dummy));
  int offset = 0, nc_glob = 0;
  for (int q = 0; q < c->nranks; ++q) { if (q < c->rank) offset += dummy[q]; nc_glob += dummy[q]; }
  // global aggregate id of every triangle of the extended range
  std::vector<int> agg_ext(rb.n_ext(), -1);
  for (int i = 0; i < F.n; ++i) agg_ext[F.off + i] = offset + A.agg[i];
  {
    DBuf<int> d_ext;
    UFE_TRY(upload(c, d_ext, agg_ext));
    UFE_TRY(halo_exchange_int(c, rb, d_ext.p));
    UFE_CUDA(cudaMemcpyAsync(agg_ext.data(), d_ext.p, sizeof(int) * agg_ext.size(), cudaMemcpyDeviceToHost, c->stream));
    UFE_CUDA(cudaStreamSynchronize(c->stream));
  }
  LocalRows R;
  build_local_rows(F, A, agg_ext, offset, nc_glob, R);
  std::vector<int> lens(A.nc_loc), all_lens, all_cols, cptr;
  for (int I = 0; I < A.nc_loc; ++I) lens[I] = R.cptr[I + 1] - R.cptr[I];
  UFE_TRY(allgatherv_int(c, lens, counts, all_lens));
  UFE_TRY(allgatherv_int(c, R.ccol, counts, all_cols));
  UFE_CHECK((int)all_lens.size() == nc_glob, "multigrid: assembled %zu coarse rows, expected %d", all_lens.size(), nc_glob);
  rows_from_parts(all_lens, cptr);
  ell_from_rows(nc_glob, cptr, all_cols, T.coarse);
  build_gather(F, A, offset, R, T.coarse, T.gal_ptr, T.gal_src);
  T.nc = nc_glob; T.nc_loc = A.nc_loc; T.offset = offset;
  T.coarse_slots = (long long)T.coarse.col.size();
  T.agg.resize(F.n);
  for (int i = 0; i < F.n; ++i) T.agg[i] = offset + A.agg[i];
  T.mem_ptr = std::move(A.mem_ptr); T.mem = std::move(A.mem);
  return 0;
}

// =====================================================================================================================
// device kernels
// =====================================================================================================================

// Every cycle operation is a __device__ function over (block id, number of blocks), so that the same code runs as a stand-alone kernel
// (level 0, and levels >= 1 in launch / graph mode) and as one step of the persistent coarse-level program (k_mg_program).

// MODE 0: out = x + Dw (b - A x)            (damped block-Jacobi sweep; Dw = omega * inverse diagonal block)
// MODE 1: out = b - A x                     (residual)
// MODE 2: out = b - A x, out2 = Dw out      (residual that becomes the right-hand side of a second coarse visit + its first sweep)
// MODE 3: x' = x + w P e, out = x' + Dw (b - A x')   (coarse-grid correction fused into the first post-smoothing sweep; the
//         neighbours' corrections are gathered through agg, so the prolongated iterate is never written; coarse levels only)
// x_ext: iterate over the extended range (node t at x2[off + t]); out, out2, b: node t at [t]
// Pointer flavours of the cycle operations.  R = true: stand-alone kernels -- inputs are read-only for the whole launch (__restrict__, so the
// compiler may use the non-coherent load path).  R = false: inside k_mg_tail one launch runs many operations and an operation reads what an
// earlier one wrote, so every vector goes through ordinary coherent loads (one CTA: the SM's L1 plus __syncthreads order them).
template <bool R> struct MgPtr {
  typedef const double* __restrict__ cd; typedef double* __restrict__ d; typedef const double2* __restrict__ cd2;
};
template <> struct MgPtr<false> {
  typedef const double* cd; typedef double* d; typedef const double2* cd2;
};

// U: slots of a row in flight at once.  Level 0 streams gigabytes and is bandwidth-bound (U = 2 keeps its registers low); the coarse levels are a
// chain of dependent loads (column -> iterate) with only a few warps per SM, so their time is load latency x chain length: U = 8 issues the
// columns, values and gathers of 8 slots before the first use and shortens the chain ~4x.  The summation order is the same for every U.
template <typename VT, int MODE, int U = 2, bool R = true>
__device__ __forceinline__ void d_mg_sweep(int n, int off, const int* __restrict__ slice_ptr, const int* __restrict__ col, const VT* __restrict__ val,
                                           typename MgPtr<R>::cd x_ext, typename MgPtr<R>::cd b, const double* __restrict__ dblk,
                                           typename MgPtr<R>::d out, int bid, int nb, typename MgPtr<R>::d out2 = nullptr,
                                           const int* __restrict__ agg = nullptr, typename MgPtr<R>::cd ec = nullptr, double w = 0.0) {
  const int lane = threadIdx.x & 31;
  const int warps_per_block = blockDim.x >> 5;
  const int n_slices = (n + 31) >> 5;
  typename MgPtr<R>::cd2 x2 = reinterpret_cast<const double2*>(x_ext);
  typename MgPtr<R>::cd2 e2 = reinterpret_cast<const double2*>(ec);
  for (int s = bid * warps_per_block + (threadIdx.x >> 5); s < n_slices; s += nb * warps_per_block) {
    const int t = (s << 5) + lane;
    const int k0 = slice_ptr[s], k1 = slice_ptr[s + 1];
    double yu = 0.0, yv = 0.0;
    const int* __restrict__ cp = col + (size_t)k0 * 32 + lane;
    const VT* __restrict__ vp = val + (size_t)k0 * 128 + lane;
    if (U <= 2) {
#pragma unroll 2
      for (int k = k0; k < k1; ++k, cp += 32, vp += 128) {
        const int tj = *cp;
        const double a0 = (double)vp[0], a1 = (double)vp[32], a2 = (double)vp[64], a3 = (double)vp[96];
        double2 xv = x2[tj];
        if (MODE == 3) { const double2 ev = e2[agg[tj]]; xv.x += w * ev.x; xv.y += w * ev.y; }
        yu += a0 * xv.x + a1 * xv.y;
        yv += a2 * xv.x + a3 * xv.y;
      }
    } else {
      for (int k = k0; k < k1; k += U, cp += U * 32, vp += U * 128) {
        int tj[U], ta[U];
        VT a[U][4];
        double2 xv[U], ev[U];
#pragma unroll
        for (int u = 0; u < U; ++u) tj[u] = (k + u < k1) ? cp[u * 32] : -1;
#pragma unroll
        for (int u = 0; u < U; ++u)
          if (tj[u] >= 0) { a[u][0] = vp[u * 128]; a[u][1] = vp[u * 128 + 32]; a[u][2] = vp[u * 128 + 64]; a[u][3] = vp[u * 128 + 96]; }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          if (tj[u] >= 0) xv[u] = x2[tj[u]];
          if (MODE == 3 && tj[u] >= 0) ta[u] = agg[tj[u]];
        }
        if (MODE == 3) {
#pragma unroll
          for (int u = 0; u < U; ++u)
            if (tj[u] >= 0) ev[u] = e2[ta[u]];
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
          if (tj[u] >= 0) {
            double2 x = xv[u];
            if (MODE == 3) { x.x += w * ev[u].x; x.y += w * ev[u].y; }
            yu += (double)a[u][0] * x.x + (double)a[u][1] * x.y;
            yv += (double)a[u][2] * x.x + (double)a[u][3] * x.y;
          }
      }
    }
    if (t < n) {
      const double2 bv = reinterpret_cast<const double2*>(b)[t];
      const double ru = bv.x - yu, rv = bv.y - yv;
      if (MODE == 1 || MODE == 2) reinterpret_cast<double2*>(out)[t] = make_double2(ru, rv);
      if (MODE != 1) {
        const double2 d01 = reinterpret_cast<const double2*>(dblk)[2 * t], d23 = reinterpret_cast<const double2*>(dblk)[2 * t + 1];
        double2 xo = make_double2(0.0, 0.0);
        if (MODE == 0 || MODE == 3) xo = x2[off + t];
        if (MODE == 3) { const double2 e = e2[agg[off + t]]; xo.x += w * e.x; xo.y += w * e.y; }
        double* dst = (MODE == 2) ? out2 : out;
        reinterpret_cast<double2*>(dst)[t] = make_double2(xo.x + d01.x * ru + d01.y * rv, xo.y + d23.x * ru + d23.y * rv);
      }
    }
  }
}
template <typename VT, int MODE, int U>
__global__ void __launch_bounds__(kBlock) k_mg_sweep(int n, int off, const int* __restrict__ slice_ptr, const int* __restrict__ col,
                                                     const VT* __restrict__ val, const double* __restrict__ x_ext, const double* __restrict__ b,
                                                     const double* __restrict__ dblk, double* __restrict__ out, double* __restrict__ out2,
                                                     const int* __restrict__ agg, const double* __restrict__ ec, double w) {
  d_mg_sweep<VT, MODE, U>(n, off, slice_ptr, col, val, x_ext, b, dblk, out, blockIdx.x, gridDim.x, out2, agg, ec, w);
}

// x = Dw b  (first sweep from a zero iterate)
template <bool R = true>
__device__ __forceinline__ void d_mg_smooth0(int n, typename MgPtr<R>::cd b, const double* __restrict__ dblk, typename MgPtr<R>::d x_own, int bid, int nb) {
  for (int t = bid * blockDim.x + threadIdx.x; t < n; t += nb * blockDim.x) {
    const double2 bv = reinterpret_cast<const double2*>(b)[t];
    const double2 d01 = reinterpret_cast<const double2*>(dblk)[2 * t], d23 = reinterpret_cast<const double2*>(dblk)[2 * t + 1];
    reinterpret_cast<double2*>(x_own)[t] = make_double2(d01.x * bv.x + d01.y * bv.y, d23.x * bv.x + d23.y * bv.y);
  }
}
__global__ void k_mg_smooth0(int n, const double* __restrict__ b, const double* __restrict__ dblk, double* __restrict__ x_own) {
  d_mg_smooth0(n, b, dblk, x_own, blockIdx.x, gridDim.x);
}

// bc_I = sum of r over the members of aggregate I (fixed order); with dblk_c also the coarse level's first sweep x_I = Dw_I bc_I
template <bool R = true>
__device__ __forceinline__ void d_mg_restrict(int nc, const int* __restrict__ mem_ptr, const int* __restrict__ mem, typename MgPtr<R>::cd r,
                                              typename MgPtr<R>::d bc, int bid, int nb, const double* __restrict__ dblk_c = nullptr,
                                              typename MgPtr<R>::d xc = nullptr) {
  constexpr int U = 8;   // members in flight (same order of summation; see d_mg_sweep)
  for (int I = bid * blockDim.x + threadIdx.x; I < nc; I += nb * blockDim.x) {
    double su = 0.0, sv = 0.0;
    const int k1 = mem_ptr[I + 1];
    for (int k = mem_ptr[I]; k < k1; k += U) {
      int mj[U];
      double2 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) mj[u] = (k + u < k1) ? mem[k + u] : -1;
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (mj[u] >= 0) v[u] = reinterpret_cast<const double2*>(r)[mj[u]];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (mj[u] >= 0) { su += v[u].x; sv += v[u].y; }
    }
    reinterpret_cast<double2*>(bc)[I] = make_double2(su, sv);
    if (dblk_c) {
      const double2 d01 = reinterpret_cast<const double2*>(dblk_c)[2 * I], d23 = reinterpret_cast<const double2*>(dblk_c)[2 * I + 1];
      reinterpret_cast<double2*>(xc)[I] = make_double2(d01.x * su + d01.y * sv, d23.x * su + d23.y * sv);
    }
  }
}
__global__ void k_mg_restrict(int nc, const int* __restrict__ mem_ptr, const int* __restrict__ mem, const double* __restrict__ r, double* __restrict__ bc,
                              const double* __restrict__ dblk_c, double* __restrict__ xc) {
  d_mg_restrict(nc, mem_ptr, mem, r, bc, blockIdx.x, gridDim.x, dblk_c, xc);
}

// x_i += w * ec_agg(i)
template <bool R = true>
__device__ __forceinline__ void d_mg_prolong(int n, const int* __restrict__ agg, typename MgPtr<R>::cd ec, double w, typename MgPtr<R>::d x_own,
                                             int bid, int nb) {
  for (int t = bid * blockDim.x + threadIdx.x; t < n; t += nb * blockDim.x) {
    const double2 e = reinterpret_cast<const double2*>(ec)[agg[t]];
    double2 x = reinterpret_cast<double2*>(x_own)[t];
    x.x += w * e.x; x.y += w * e.y;
    reinterpret_cast<double2*>(x_own)[t] = x;
  }
}
__global__ void k_mg_prolong(int n, const int* __restrict__ agg, const double* __restrict__ ec, double w, double* __restrict__ x_own) {
  d_mg_prolong(n, agg, ec, w, x_own, blockIdx.x, gridDim.x);
}

template <bool R = true>
__device__ __forceinline__ void d_mg_axpy(int n2, typename MgPtr<R>::cd a, typename MgPtr<R>::d y, int bid, int nb) {
  for (int i = bid * blockDim.x + threadIdx.x; i < n2; i += nb * blockDim.x) y[i] += a[i];
}
__global__ void k_mg_axpy(int n2, const double* __restrict__ a, double* __restrict__ y) { d_mg_axpy(n2, a, y, blockIdx.x, gridDim.x); }

// coarse block values: sum of the fine blocks listed for the slot
// ---- level 0: the coarse-grid correction with its own step length --------------------------------------------------------------------------
// Plain aggregation under-corrects, and the factor that repairs it depends on the problem (1.5 is safe everywhere, 2.0 - 2.3 halves the Krylov count
// on the 4 M-vertex MISMIP_mod mesh but diverges on MISMIP+; profiles/r01_mg_probe_4M_b.log).  MgParams::adaptive replaces the fixed factor on level 0
// by the one that minimises the residual of the corrected iterate: with r = b - A x (the pre-smoothing residual that was restricted) and q = A (P e),
// w = (r, q) / (q, q), x <- x + w P e, and the post-smoothing sweep x + Dw (b - A x) follows from r - w q without another pass over the matrix.
// So the cycle still streams the level-0 matrix twice; what is new is one more vector (q) and two dot products, fused into the pass that forms q.
__global__ void k_mg_prolong_set(int n, const int* __restrict__ agg, const double* __restrict__ ec, double* __restrict__ pe_own) {
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x)
    reinterpret_cast<double2*>(pe_own)[t] = reinterpret_cast<const double2*>(ec)[agg[t]];
}
// q = A pe (owned rows), scal[0] = (r, q), scal[1] = (q, q)
template <typename VT>
__global__ void __launch_bounds__(kBlock) k_mg_aq(int n, const int* __restrict__ slice_ptr, const int* __restrict__ col, const VT* __restrict__ val,
                                                  const double* __restrict__ pe_ext, const double* __restrict__ r, double* __restrict__ q,
                                                  double* scal, double* partials, unsigned* counter) {
  const int lane = threadIdx.x & 31;
  const int warps_per_block = blockDim.x >> 5;
  const int n_slices = (n + 31) >> 5;
  const double2* __restrict__ x2 = reinterpret_cast<const double2*>(pe_ext);
  double acc[2] = {0.0, 0.0};
  for (int s = blockIdx.x * warps_per_block + (threadIdx.x >> 5); s < n_slices; s += gridDim.x * warps_per_block) {
    const int t = (s << 5) + lane;
    const int k0 = slice_ptr[s], k1 = slice_ptr[s + 1];
    double yu = 0.0, yv = 0.0;
    const int* __restrict__ cp = col + (size_t)k0 * 32 + lane;
    const VT* __restrict__ vp = val + (size_t)k0 * 128 + lane;
#pragma unroll 2
    for (int k = k0; k < k1; ++k, cp += 32, vp += 128) {
      const int tj = *cp;
      const double a0 = (double)vp[0], a1 = (double)vp[32], a2 = (double)vp[64], a3 = (double)vp[96];
      const double2 xv = x2[tj];
      yu += a0 * xv.x + a1 * xv.y;
      yv += a2 * xv.x + a3 * xv.y;
    }
    if (t < n) {
      reinterpret_cast<double2*>(q)[t] = make_double2(yu, yv);
      const double2 rv = reinterpret_cast<const double2*>(r)[t];
      acc[0] += rv.x * yu + rv.y * yv;
      acc[1] += yu * yu + yv * yv;
    }
  }
  grid_sum<2>(acc, partials, counter, scal);
}
// out = x + w pe + Dw (r - w q);  w = (r, q) / (q, q), kept inside [0, w_max]; (q, q) = 0 (nothing to correct): w = 0
__global__ void k_mg_mr_update(int n, int off, const double* __restrict__ x_ext, const double* __restrict__ pe_ext, const double* __restrict__ r,
                               const double* __restrict__ q, const double* __restrict__ dblk, const double* __restrict__ scal, double w_max,
                               double* __restrict__ out) {
  const double rq = scal[0], qq = scal[1];
  double w = (qq > 0.0) ? rq / qq : 0.0;
  w = fmin(fmax(w, 0.0), w_max);
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) {
    const double2 xv = reinterpret_cast<const double2*>(x_ext)[off + t], pv = reinterpret_cast<const double2*>(pe_ext)[off + t];
    const double2 rv = reinterpret_cast<const double2*>(r)[t], qv = reinterpret_cast<const double2*>(q)[t];
    const double2 d01 = reinterpret_cast<const double2*>(dblk)[2 * t], d23 = reinterpret_cast<const double2*>(dblk)[2 * t + 1];
    const double ru = rv.x - w * qv.x, rw = rv.y - w * qv.y;
    reinterpret_cast<double2*>(out)[t] = make_double2(xv.x + w * pv.x + d01.x * ru + d01.y * rw, xv.y + w * pv.y + d23.x * ru + d23.y * rw);
  }
}

template <typename VTI, typename VTO>
__global__ void k_mg_galerkin(long long n_cslots, const int* __restrict__ gal_ptr, const int* __restrict__ gal_src, const VTI* __restrict__ fv,
                              VTO* __restrict__ cv) {
  constexpr int U = 4;   // source slots in flight: indices first, then their 4 x U values (same order of summation)
  for (long long cs = (long long)blockIdx.x * blockDim.x + threadIdx.x; cs < n_cslots; cs += (long long)gridDim.x * blockDim.x) {
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    const int k1 = gal_ptr[cs + 1];
    for (int k = gal_ptr[cs]; k < k1; k += U) {
      int fs[U];
      VTI v[U][4];
#pragma unroll
      for (int u = 0; u < U; ++u) fs[u] = (k + u < k1) ? gal_src[k + u] : -1;
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (fs[u] >= 0) {
          const VTI* p = fv + ((size_t)(fs[u] >> 5) * 128) + (fs[u] & 31);
          v[u][0] = p[0]; v[u][1] = p[32]; v[u][2] = p[64]; v[u][3] = p[96];
        }
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (fs[u] >= 0) { a0 += (double)v[u][0]; a1 += (double)v[u][1]; a2 += (double)v[u][2]; a3 += (double)v[u][3]; }
    }
    VTO* q = cv + ((size_t)(cs >> 5) * 128) + (cs & 31);
    q[0] = (VTO)a0; q[32] = (VTO)a1; q[64] = (VTO)a2; q[96] = (VTO)a3;
  }
}

// FP32 copy of the level-0 values straight from the CSR matrix through the block scatter map
__global__ void k_mg_scatter_f32(long long nnz, const int* __restrict__ map, const double* __restrict__ v, float* __restrict__ out) {
  for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += (long long)gridDim.x * blockDim.x) out[map[k]] = (float)v[k];
}

// Dw = omega * inverse of the diagonal 2x2 block (sum of every slot that points at the own node; padding holds zeros)
template <typename VT>
__global__ void k_mg_diag(int n, int off, const int* __restrict__ slice_ptr, const int* __restrict__ col, const VT* __restrict__ val, double omega,
                          double* __restrict__ dblk) {
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) {
    const int s = t >> 5, lane = t & 31;
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    for (int k = slice_ptr[s]; k < slice_ptr[s + 1]; ++k) {
      if (col[(size_t)k * 32 + lane] != off + t) continue;
      const VT* p = val + (size_t)k * 128 + lane;
      a0 += (double)p[0]; a1 += (double)p[32]; a2 += (double)p[64]; a3 += (double)p[96];
    }
    const double det = a0 * a3 - a1 * a2;
    double i0, i1, i2, i3;
    if (det != 0.0 && fabs(det) > 1e-14 * (fabs(a0 * a3) + fabs(a1 * a2))) {
      const double id = omega / det;
      i0 = a3 * id; i1 = -a1 * id; i2 = -a2 * id; i3 = a0 * id;
    } else {  // singular block: plain Jacobi on whatever diagonal exists
      i0 = a0 != 0.0 ? omega / a0 : omega; i3 = a3 != 0.0 ? omega / a3 : omega; i1 = i2 = 0.0;
    }
    reinterpret_cast<double2*>(dblk)[2 * t] = make_double2(i0, i1);
    reinterpret_cast<double2*>(dblk)[2 * t + 1] = make_double2(i2, i3);
  }
}

// dense copy of the coarsest matrix into the left half of the Gauss-Jordan workspace W[N][2N]; right half = identity
template <typename VT>
__global__ void k_mg_dense_fill(int n, const int* __restrict__ slice_ptr, const int* __restrict__ col, const VT* __restrict__ val, double* __restrict__ W) {
  const int N = 2 * n;
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) {
    const int s = t >> 5, lane = t & 31;
    double* ru = W + (size_t)(2 * t) * 2 * N;
    double* rv = ru + 2 * N;
    for (int k = slice_ptr[s]; k < slice_ptr[s + 1]; ++k) {
      const int cj = col[(size_t)k * 32 + lane];
      const VT* p = val + (size_t)k * 128 + lane;
      ru[2 * cj] += (double)p[0]; ru[2 * cj + 1] += (double)p[32];
      rv[2 * cj] += (double)p[64]; rv[2 * cj + 1] += (double)p[96];
    }
    ru[N + 2 * t] = 1.0; rv[N + 2 * t + 1] = 1.0;
  }
}

// single-block Gauss-Jordan with partial pivoting on W[N][2N]; writes the inverse column-major: inv[j*N + i] = (A^-1)(i,j)
__global__ void __launch_bounds__(1024) k_mg_dense_invert(int N, double* __restrict__ W, double* __restrict__ inv) {
  extern __shared__ double sh[];   // N factors + reduction scratch
  double* fac = sh;
  __shared__ double s_max[32];
  __shared__ int s_arg[32];
  __shared__ int s_piv;
  __shared__ double s_pval;
  const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, warp = tid >> 5;
  const int ld = 2 * N;
  for (int p = 0; p < N; ++p) {
    // pivot search in column p, rows p..N-1
    double best = -1.0; int arg = p;
    for (int i = p + tid; i < N; i += nt) {
      const double a = fabs(W[(size_t)i * ld + p]);
      if (a > best) { best = a; arg = i; }
    }
    for (int o = 16; o > 0; o >>= 1) {
      const double ob = __shfl_xor_sync(0xffffffffu, best, o);
      const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
      if (ob > best || (ob == best && oa < arg)) { best = ob; arg = oa; }
    }
    if (lane == 0) { s_max[warp] = best; s_arg[warp] = arg; }
    __syncthreads();
    if (tid == 0) {
      double bb = s_max[0]; int aa = s_arg[0];
      for (int w = 1; w < (nt + 31) / 32; ++w)
        if (s_max[w] > bb || (s_max[w] == bb && s_arg[w] < aa)) { bb = s_max[w]; aa = s_arg[w]; }
      s_piv = aa;
      const double pv = W[(size_t)aa * ld + p];
      s_pval = (pv != 0.0) ? pv : 1.0;
    }
    __syncthreads();
    const int piv = s_piv;
    const double ipv = 1.0 / s_pval;
    // swap rows p <-> piv and scale the pivot row
    for (int j = tid; j < ld; j += nt) {
      const double a = W[(size_t)piv * ld + j];
      if (piv != p) W[(size_t)piv * ld + j] = W[(size_t)p * ld + j];
      W[(size_t)p * ld + j] = a * ipv;
    }
    __syncthreads();
    for (int i = tid; i < N; i += nt) fac[i] = (i == p) ? 0.0 : W[(size_t)i * ld + p];
    __syncthreads();
    // eliminate column p from every other row; columns < p of the left half are already unit vectors
    const int j0 = p, ncol = ld - j0;
    for (long long e = tid; e < (long long)N * ncol; e += nt) {
      const int i = (int)(e / ncol), j = j0 + (int)(e % ncol);
      const double f = fac[i];
      if (f != 0.0) W[(size_t)i * ld + j] -= f * W[(size_t)p * ld + j];
    }
    __syncthreads();
  }
  for (long long e = tid; e < (long long)N * N; e += nt) {
    const int i = (int)(e / N), j = (int)(e % N);
    inv[(size_t)j * N + i] = W[(size_t)i * ld + N + j];
  }
}

// x = A^-1 b on the coarsest level (one block; sb: N doubles of shared memory)
template <bool R = true>
__device__ __forceinline__ void d_mg_dense_solve(int N, const double* __restrict__ inv, typename MgPtr<R>::cd b, typename MgPtr<R>::d x, double* sb) {
  for (int j = threadIdx.x; j < N; j += blockDim.x) sb[j] = b[j];
  __syncthreads();
  for (int i = threadIdx.x; i < N; i += blockDim.x) {
    double s = 0.0;
    for (int j = 0; j < N; ++j) s += inv[(size_t)j * N + i] * sb[j];
    x[i] = s;
  }
}
__global__ void __launch_bounds__(1024) k_mg_dense_solve(int N, const double* __restrict__ inv, const double* __restrict__ b, double* __restrict__ x) {
  extern __shared__ double sb[];
  d_mg_dense_solve(N, inv, b, x, sb);
}

// ---- the coarse-grid correction as ONE persistent cooperative kernel ------------------------------------------------------------------
// Levels >= 1 are a fixed sequence of ~170 small dependent operations per cycle (W-cycle on top, V below): a few microseconds of work each,
// so stream launches and even CUDA-graph nodes are dominated by launch/drain latency.  k_mg_program keeps one grid resident (a few CTAs per
// SM, cooperative launch) and walks the recorded step list, separating steps by a grid-wide barrier instead of a kernel boundary.
// Measured on B200 at 4 M vertices (profiles/r01_mg_probe_modes.log): 0.91 ms per cycle with 4 CTAs/SM, 0.98 with 2, 1.27 with 1, against
// 0.74 ms for the CUDA graph -- cooperative_groups' grid.sync() across 148 SMs on two dies costs more than a graph-node boundary.  The graph
// stays the default (MgParams::use_graph = 1); this path is kept selectable (UFE_B200_MG_GRAPH=2) and is covered by the same tests.
template <typename VT>
__global__ void __launch_bounds__(kBlock) k_mg_program(const MgStep* __restrict__ steps, int nsteps) {
  cooperative_groups::grid_group grid = cooperative_groups::this_grid();
  __shared__ double sb[2 * 128];   // dense coarsest solve: at most 128 nodes (MgParams::dense_max)
  const int bid = blockIdx.x, nb = gridDim.x;
  for (int s = 0; s < nsteps; ++s) {
    const MgStep st = steps[s];
    switch (st.op) {
      case MG_OP_SMOOTH0: d_mg_smooth0(st.n, st.b, st.d, st.out, bid, nb); break;
      case MG_OP_SWEEP: d_mg_sweep<VT, 0, 8>(st.n, st.off, st.ip0, st.ip1, static_cast<const VT*>(st.val), st.x, st.b, st.d, st.out, bid, nb); break;
      case MG_OP_RESID: d_mg_sweep<VT, 1, 8>(st.n, st.off, st.ip0, st.ip1, static_cast<const VT*>(st.val), st.x, st.b, st.d, st.out, bid, nb); break;
      case MG_OP_RESTRICT: d_mg_restrict(st.n, st.ip0, st.ip1, st.x, st.out, bid, nb); break;
      case MG_OP_PROLONG: d_mg_prolong(st.n, st.ip0, st.x, st.w, st.out, bid, nb); break;
      case MG_OP_AXPY: d_mg_axpy(st.n, st.x, st.out, bid, nb); break;
      case MG_OP_DENSE: if (bid == 0) d_mg_dense_solve(st.n, static_cast<const double*>(st.val), st.b, st.out, sb); break;
    }
    if (s + 1 < nsteps) grid.sync();
  }
}

// ---- the small levels at the bottom of the hierarchy as ONE single-CTA kernel per visit ----------------------------------------------------
// From the first level with at most MgParams::tail_max nodes downwards a level visit is ~13 dependent operations on a few hundred to a few
// thousand nodes; as graph nodes each costs a kernel boundary (~2 us) plus its own load-latency chain, and with a W-cycle above them these
// levels are visited 8 times per cycle -- about half of the coarse-grid time at 4 M vertices.  One CTA of 512 threads walks the recorded
// operations of the whole sub-cycle instead, separated by __syncthreads(): no kernel boundaries, the matrices stay in this SM's L1/L2, and
// the arithmetic (hence every result bit) is the same as in the stand-alone kernels.
// Measured on B200 at 4 M vertices (profiles/r01_mg_probe_tail.log): 0.82 ms per cycle against 0.72 ms with plain graph nodes -- a kernel
// boundary inside a graph is cheaper than one SM streaming a 0.4 MB level from L2 three times per visit.  Kept selectable
// (MgParams::tail_max / UFE_B200_MG_TAIL, default off) and covered by tests/test_gpu_multigrid.py.
constexpr int kTailBlock = 512, kTailU = 4;
template <typename VT>
__global__ void __launch_bounds__(kTailBlock, 1) k_mg_tail(const MgStep* __restrict__ steps, int nsteps) {
  __shared__ double sb[2 * 128];   // dense coarsest solve: at most 128 nodes (MgParams::dense_max)
  for (int s = 0; s < nsteps; ++s) {
    const MgStep st = steps[s];
    switch (st.op) {
      case MG_OP_SMOOTH0: d_mg_smooth0<false>(st.n, st.b, st.d, st.out, 0, 1); break;
      case MG_OP_SWEEP: d_mg_sweep<VT, 0, kTailU, false>(st.n, st.off, st.ip0, st.ip1, static_cast<const VT*>(st.val), st.x, st.b, st.d, st.out, 0, 1); break;
      case MG_OP_RESID: d_mg_sweep<VT, 1, kTailU, false>(st.n, st.off, st.ip0, st.ip1, static_cast<const VT*>(st.val), st.x, st.b, st.d, st.out, 0, 1); break;
      case MG_OP_RESTRICT: d_mg_restrict<false>(st.n, st.ip0, st.ip1, st.x, st.out, 0, 1); break;
      case MG_OP_PROLONG: d_mg_prolong<false>(st.n, st.ip0, st.x, st.w, st.out, 0, 1); break;
      case MG_OP_AXPY: d_mg_axpy<false>(st.n, st.x, st.out, 0, 1); break;
      case MG_OP_DENSE: d_mg_dense_solve<false>(st.n, static_cast<const double*>(st.val), st.b, st.out, sb); break;
    }
    __syncthreads();
  }
}

// The same step list on one thread-block CLUSTER: the CTAs of the cluster (8, or 16 with the non-portable size) share every operation, separated by
// the cluster barrier (barrier.cluster, release / acquire at cluster scope: a few hundred nanoseconds in hardware, against ~3 us for a grid-wide
// cooperative barrier over 148 SMs and ~4 us for a graph-node boundary).  That makes the tail worthwhile from a larger level on (MgParams::tail_max
// ~ 10^4 nodes: level 3 of the 4 M-vertex hierarchy, 16 operations per visit, 8 visits per cycle).  Arithmetic and summation order per row are those
// of the stand-alone kernels.
template <typename VT>
__global__ void __launch_bounds__(kTailBlock, 1) k_mg_tail_cluster(const MgStep* __restrict__ steps, int nsteps) {
  cooperative_groups::cluster_group cl = cooperative_groups::this_cluster();
  const int bid = (int)cl.block_rank(), nb = (int)cl.num_blocks();
  __shared__ double sb[2 * 128];
  for (int s = 0; s < nsteps; ++s) {
    const MgStep st = steps[s];
    switch (st.op) {
      case MG_OP_SMOOTH0: d_mg_smooth0<false>(st.n, st.b, st.d, st.out, bid, nb); break;
      case MG_OP_SWEEP: d_mg_sweep<VT, 0, kTailU, false>(st.n, st.off, st.ip0, st.ip1, static_cast<const VT*>(st.val), st.x, st.b, st.d, st.out, bid, nb); break;
      case MG_OP_RESID: d_mg_sweep<VT, 1, kTailU, false>(st.n, st.off, st.ip0, st.ip1, static_cast<const VT*>(st.val), st.x, st.b, st.d, st.out, bid, nb); break;
      case MG_OP_RESTRICT: d_mg_restrict<false>(st.n, st.ip0, st.ip1, st.x, st.out, bid, nb); break;
      case MG_OP_PROLONG: d_mg_prolong<false>(st.n, st.ip0, st.x, st.w, st.out, bid, nb); break;
      case MG_OP_AXPY: d_mg_axpy<false>(st.n, st.x, st.out, bid, nb); break;
      case MG_OP_DENSE: if (bid == 0) d_mg_dense_solve<false>(st.n, static_cast<const double*>(st.val), st.b, st.out, sb); break;
    }
    cl.sync();
  }
}
template <typename VT>
int launch_tail_cluster(Ctx* c, int cluster, const MgStep* steps, int nsteps) {
  static bool allowed = false;
  if (cluster > 8 && !allowed) {
    UFE_CUDA(cudaFuncSetAttribute(k_mg_tail_cluster<VT>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    allowed = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(cluster); cfg.blockDim = dim3(kTailBlock); cfg.dynamicSmemBytes = 0; cfg.stream = c->stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  UFE_CUDA(cudaLaunchKernelEx(&cfg, k_mg_tail_cluster<VT>, steps, nsteps));
  return 0;
}

// ---- the sweeps on the coarsest level as ONE thread-block-cluster kernel per visit --------------------------------------------------------------
// A hierarchy that stops at ~10^3 nodes (MgParams::stop_nodes) solves that level with `coarse_sweeps` damped block-Jacobi sweeps.  As graph nodes
// each sweep costs ~3 us although it moves only ~0.5 MB: two dependent L2 round trips (column index -> iterate) plus a kernel boundary, 8 visits x 24
// sweeps per cycle.  Here every CTA of the cluster first copies the matrix slices of its rows into shared memory (the whole level is ~0.6 MB: 16 CTAs
// hold it), then runs all sweeps: columns and values come from shared memory, only the iterate is gathered from L2 (ld.global.cg: never from a stale
// L1 line), written back and published to the other CTAs by the cluster barrier (barrier.cluster, release / acquire at cluster scope).  Same row
// arithmetic and summation order as k_mg_smooth0 / k_mg_sweep<MODE 0>.
// upper bound of the spectral radius of Dw A on one level: the largest absolute row sum of the 2x2 blocks Dw_i A_ij (infinity norm), as the bits of a
// positive double through atomicMax.  Input of the Chebyshev variant of k_mg_coarsest_cluster; one pass per numeric set-up.
template <typename VT>
__global__ void k_mg_lmax_bound(int n, const int* __restrict__ slice_ptr, const int* __restrict__ col, const VT* __restrict__ val, const double* __restrict__ dblk,
                                unsigned long long* out_bits) {
  double worst = 0.0;
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) {
    const int s = t >> 5, lane = t & 31;
    const double2 d01 = reinterpret_cast<const double2*>(dblk)[2 * t], d23 = reinterpret_cast<const double2*>(dblk)[2 * t + 1];
    double r0 = 0.0, r1 = 0.0;
    for (int k = slice_ptr[s]; k < slice_ptr[s + 1]; ++k) {
      const VT* p = val + (size_t)k * 128 + lane;
      const double a0 = (double)p[0], a1 = (double)p[32], a2 = (double)p[64], a3 = (double)p[96];
      r0 += fabs(d01.x * a0 + d01.y * a2) + fabs(d01.x * a1 + d01.y * a3);
      r1 += fabs(d23.x * a0 + d23.y * a2) + fabs(d23.x * a1 + d23.y * a3);
    }
    worst = fmax(worst, fmax(r0, r1));
  }
  for (int o = 16; o > 0; o >>= 1) worst = fmax(worst, __shfl_xor_sync(0xffffffffu, worst, o));
  if ((threadIdx.x & 31) == 0 && worst > 0.0) atomicMax(out_bits, (unsigned long long)__double_as_longlong(worst));
}

constexpr int kCoarsestBlock = 256;
// DS = true: the iterate lives in every CTA's shared memory (two buffers of n double2); a CTA writes the new values of its rows into the next buffer of
// ALL CTAs through distributed shared memory (cluster.map_shared_rank) and the gathers of the next sweep are local shared-memory reads -- no L2 round
// trip left inside a sweep.  DS = false: the iterate is exchanged through L2 (ld.global.cg / st.global + cluster barrier).
template <typename VT, bool DS>
__global__ void __launch_bounds__(kCoarsestBlock, 1) k_mg_coarsest_cluster(int n, const int* __restrict__ slice_ptr, const int* __restrict__ col,
                                                                           const VT* __restrict__ val, const double* __restrict__ dblk,
                                                                           const double* __restrict__ b, double* xa, double* xb, double* out, int nsweeps,
                                                                           const double* __restrict__ lmax, double cheb_ratio) {
  // lmax != nullptr: the nsweeps corrections are Chebyshev steps for the eigenvalue interval [lmax / cheb_ratio, lmax] of Dw A instead of stationary
  // sweeps: x_{k+1} = x_k + d_k, d_k = rho_k rho_{k-1} d_{k-1} + (2 rho_k / delta) Dw (b - A x_k) -- a fixed polynomial of Dw A, i.e. still a linear,
  // iteration-independent preconditioner, which damps the whole interval by ((sqrt(kappa) - 1) / (sqrt(kappa) + 1))^k where a stationary sweep needs
  // (1 - 1 / kappa)^k.  d_k of a thread's own rows stays in registers.
  const bool cheb = lmax != nullptr;
  double theta = 1.0, delta = 1.0, sigma = 1.0, rho_old = 1.0;
  if (cheb) {
    const double hi = 1.05 * fmax(*lmax, 1e-300), lo = hi / cheb_ratio;
    theta = 0.5 * (hi + lo); delta = 0.5 * (hi - lo); sigma = theta / delta; rho_old = 1.0 / sigma;
  }
  constexpr int kMaxOwn = 4;      // slices a warp may own (launch_coarsest_cluster checks it)
  double2 dk[kMaxOwn];
  cooperative_groups::cluster_group cl = cooperative_groups::this_cluster();
  const int rank = (int)cl.block_rank(), C = (int)cl.num_blocks();
  extern __shared__ __align__(16) unsigned char s_raw[];
  const int n_slices = (n + 31) >> 5;
  const int per = (n_slices + C - 1) / C;
  const int s0 = min(n_slices, rank * per), s1 = min(n_slices, s0 + per);
  const int k_lo = slice_ptr[s0], k_hi = slice_ptr[s1];
  // layout: [iterate buffers: 2 x n_slices * 32 double2 (DS only)] [values: slot x 4 x 32] [columns: slot x 32]; the iterate region has the same
  // size in every CTA, so the same offset addresses it in every peer
  const size_t xbytes = DS ? sizeof(double2) * 2 * (size_t)n_slices * 32 : 0;
  double2* s_x = reinterpret_cast<double2*>(s_raw);
  VT* s_val = reinterpret_cast<VT*>(s_raw + xbytes);
  int* s_col = reinterpret_cast<int*>(s_raw + xbytes + sizeof(VT) * 128 * (size_t)(k_hi - k_lo));
  for (int i = threadIdx.x; i < (k_hi - k_lo) * 128; i += blockDim.x) s_val[i] = val[(size_t)k_lo * 128 + i];
  for (int i = threadIdx.x; i < (k_hi - k_lo) * 32; i += blockDim.x) s_col[i] = col[(size_t)k_lo * 32 + i];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int nx = n_slices * 32;
  auto publish = [&](int buf, int t, double2 v) {     // DS: node t of buffer `buf` in every CTA of the cluster
    for (int r = 0; r < C; ++r) cl.map_shared_rank(s_x, r)[buf * nx + t] = v;
  };
  // first sweep from a zero iterate: x = Dw b  (Chebyshev: x = d_0 = Dw b / theta)
  {
    int own = 0;
    for (int s = s0 + warp; s < s1; s += nwarps, ++own) {
      const int t = (s << 5) + lane;
      double2 v = make_double2(0.0, 0.0);
      if (t < n) {
        const double2 bv = reinterpret_cast<const double2*>(b)[t];
        const double2 d01 = reinterpret_cast<const double2*>(dblk)[2 * t], d23 = reinterpret_cast<const double2*>(dblk)[2 * t + 1];
        v = make_double2(d01.x * bv.x + d01.y * bv.y, d23.x * bv.x + d23.y * bv.y);
        if (cheb) { v.x /= theta; v.y /= theta; }
        if (!DS) reinterpret_cast<double2*>(xa)[t] = v;
      }
      if (own < kMaxOwn) dk[own] = v;
      if (DS) publish(0, t, v);       // padding rows of the last slice hold zeros (their columns may be read, their values are zero)
    }
  }
  cl.sync();        // also orders the shared-memory copies (it is a superset of __syncthreads)
  double* src = xa; double* dst = xb;
  for (int sw = 0; sw < nsweeps; ++sw) {
    const bool last = sw == nsweeps - 1;
    if (last) dst = out;
    const double2* x2 = DS ? s_x + (sw & 1) * nx : reinterpret_cast<const double2*>(src);
    const double rho = 1.0 / (2.0 * sigma - rho_old);        // Chebyshev coefficients of this step (the same in every thread)
    const double c_d = rho * rho_old, c_z = 2.0 * rho / delta;
    int own = 0;
    for (int s = s0 + warp; s < s1; s += nwarps, ++own) {
      const int t = (s << 5) + lane;
      const int k0 = slice_ptr[s] - k_lo, k1 = slice_ptr[s + 1] - k_lo;
      double yu = 0.0, yv = 0.0;
      constexpr int U = 8;
      for (int k = k0; k < k1; k += U) {
        double2 xv[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
          if (k + u < k1) xv[u] = DS ? x2[s_col[(k + u) * 32 + lane]] : __ldcg(x2 + s_col[(k + u) * 32 + lane]);
#pragma unroll
        for (int u = 0; u < U; ++u)
          if (k + u < k1) {
            const VT* vp = s_val + (size_t)(k + u) * 128 + lane;
            yu += (double)vp[0] * xv[u].x + (double)vp[32] * xv[u].y;
            yv += (double)vp[64] * xv[u].x + (double)vp[96] * xv[u].y;
          }
      }
      double2 v = make_double2(0.0, 0.0);
      if (t < n) {
        const double2 bv = reinterpret_cast<const double2*>(b)[t];
        const double ru = bv.x - yu, rv = bv.y - yv;
        const double2 d01 = reinterpret_cast<const double2*>(dblk)[2 * t], d23 = reinterpret_cast<const double2*>(dblk)[2 * t + 1];
        const double2 xo = DS ? x2[t] : __ldcg(x2 + t);
        if (cheb) {
          const double2 dprev = dk[own < kMaxOwn ? own : 0];
          const double2 dn = make_double2(c_d * dprev.x + c_z * (d01.x * ru + d01.y * rv), c_d * dprev.y + c_z * (d23.x * ru + d23.y * rv));
          if (own < kMaxOwn) dk[own] = dn;
          v = make_double2(xo.x + dn.x, xo.y + dn.y);
        } else {
          v = make_double2(xo.x + d01.x * ru + d01.y * rv, xo.y + d23.x * ru + d23.y * rv);
        }
        if (!DS || last) reinterpret_cast<double2*>(dst)[t] = v;
      }
      if (DS && !last) publish((sw + 1) & 1, t, v);
    }
    rho_old = rho;
    cl.sync();      // (after the last sweep: no CTA may exit while a peer could still write into its shared memory)
    double* tmp = src; src = dst; dst = (tmp == out) ? xb : tmp;
  }
}
// shared memory a cluster of C CTAs needs for level L (the largest share of slots of any CTA), or 0 if the level has no slices
template <typename VT>
size_t coarsest_cluster_smem(const std::vector<int>& slice_ptr_host, int n, int C) {
  const int n_slices = (n + 31) >> 5, per = (n_slices + C - 1) / C;
  int worst = 0;
  for (int r = 0; r < C; ++r) {
    const int s0 = std::min(n_slices, r * per), s1 = std::min(n_slices, s0 + per);
    worst = std::max(worst, slice_ptr_host[s1] - slice_ptr_host[s0]);
  }
  return (size_t)worst * (128 * sizeof(VT) + 32 * sizeof(int));
}
inline size_t coarsest_cluster_xbytes(int n) { return sizeof(double2) * 2 * (size_t)((n + 31) / 32) * 32; }
template <typename VT, bool DS>
int launch_coarsest_cluster(Ctx* c, int cluster, int threads, size_t smem, const MgLevel& L, const VT* val, const double* b, double* xa, double* xb, double* out, int nsweeps,
                            const double* lmax, double cheb_ratio) {
  static size_t smem_set = 0;
  static bool nonportable = false;
  if (DS) smem += coarsest_cluster_xbytes(L.n);
  if (smem > smem_set) { UFE_CUDA(cudaFuncSetAttribute(k_mg_coarsest_cluster<VT, DS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); smem_set = smem; }
  if (cluster > 8 && !nonportable) { UFE_CUDA(cudaFuncSetAttribute(k_mg_coarsest_cluster<VT, DS>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1)); nonportable = true; }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(cluster); cfg.blockDim = dim3(std::min(kCoarsestBlock, std::max(32, (threads / 32) * 32))); cfg.dynamicSmemBytes = smem; cfg.stream = c->stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  UFE_CUDA(cudaLaunchKernelEx(&cfg, k_mg_coarsest_cluster<VT, DS>, L.n, L.slice_ptr, L.col, val, (const double*)L.dblk.p, b, xa, xb, out, nsweeps, lmax, cheb_ratio));
  return 0;
}

// ---- BiCGStab pieces that differ from the Jacobi path (the preconditioner is no longer fused into the SpMV) ----------

// Bb = M^-1 b, Br = M^-1 (b - A x0) (Br == Bb and x zeroed when there is no initial guess)
__global__ void k_ksmg_init(int m, const double* __restrict__ Bb, const double* __restrict__ Br, int zero_x, double* __restrict__ R,
                            double* __restrict__ RP, double* __restrict__ x_own, double* __restrict__ P_own, double* __restrict__ V, double* scal,
                            double* partials, unsigned* counter, int* flags, double rtol, double abstol) {
  double acc[2] = {0.0, 0.0};
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (long long)gridDim.x * blockDim.x) {
    const double pb = Bb[i], r = Br[i];
    if (zero_x) x_own[i] = 0.0;
    R[i] = r; RP[i] = r; P_own[i] = 0.0; V[i] = 0.0;
    acc[0] += pb * pb;
    acc[1] += r * r;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    flags[KF_DONE] = 0; flags[KF_ITS] = 0; flags[KF_REASON] = 0;
    scal[KS_RTOL] = rtol; scal[KS_ABSTOL] = abstol;
  }
  grid_sum<2>(acc, partials, counter, scal + KS_RR0);
}

__global__ void k_ksmg_sub(int m, const double* __restrict__ b, const double* __restrict__ Ax, double* __restrict__ out) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (long long)gridDim.x * blockDim.x) out[i] = b[i] - Ax[i];
}

// D1 = (v, rp)
__global__ void k_ksmg_dot1(int m, int iter, const double* __restrict__ v, const double* __restrict__ rp, double* scal, double* partials,
                            unsigned* counter, const int* flags) {
  if (flags[KF_DONE]) return;
  double acc[1] = {0.0};
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (long long)gridDim.x * blockDim.x) acc[0] += v[i] * rp[i];
  grid_sum<1>(acc, partials, counter, scal + (iter & 1) * KS_BANK + KS_D1);
}
// ST = (s, t), TT = (t, t)
__global__ void k_ksmg_dot2(int m, int iter, const double* __restrict__ s, const double* __restrict__ t, double* scal, double* partials,
                            unsigned* counter, const int* flags) {
  if (flags[KF_DONE]) return;
  double acc[2] = {0.0, 0.0};
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (long long)gridDim.x * blockDim.x) {
    const double tv = t[i];
    acc[0] += s[i] * tv; acc[1] += tv * tv;
  }
  grid_sum<2>(acc, partials, counter, scal + (iter & 1) * KS_BANK + KS_ST);
}

// Several ranks: every separate reduction is one more ncclAllReduce on the critical path, so the second half of the iteration carries all five
// inner products of S and T at once -- ST = (s, t), TT = (t, t), SS = (s, s), SRP = (s, r~), TRP = (t, r~) (adjacent slots KS_ST..KS_TRP) --
// and the update forms the next RHO = (r, r~) = SRP - omega TRP and RR = (r, r) = SS - 2 omega ST + omega^2 TT from them instead of reducing again:
// two all-reduces per iteration instead of three.  The expansion of RR loses log10( SS / RR) digits to cancellation -- one or two per iteration at
// the contraction rates of this preconditioner -- which is why the single-rank path keeps the direct sums.
__global__ void k_ksmg_dot5(int m, int iter, const double* __restrict__ s, const double* __restrict__ t, const double* __restrict__ rp, double* scal,
                            double* partials, unsigned* counter, const int* flags) {
  if (flags[KF_DONE]) return;
  double acc[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (long long)gridDim.x * blockDim.x) {
    const double sv = s[i], tv = t[i], rv = rp[i];
    acc[0] += sv * tv; acc[1] += tv * tv; acc[2] += sv * sv; acc[3] += sv * rv; acc[4] += tv * rv;
  }
  grid_sum<5>(acc, partials, counter, scal + (iter & 1) * KS_BANK + KS_ST);
}
// x += alpha P + omega S;  R = S - omega T;  RHO, RR of the next iteration from the five inner products (no reduction)
__global__ void k_ksmg_update_xr5(int m, int iter, double* __restrict__ x_own, const double* __restrict__ P_own, const double* __restrict__ S_own,
                                  const double* __restrict__ T, double* __restrict__ R, double* scal, const int* flags) {
  if (flags[KF_DONE]) return;
  const int p = iter & 1;
  const double* cur = scal + p * KS_BANK;
  const double alpha = cur[KS_RHO] / cur[KS_D1];
  const double omega = (cur[KS_TT] == 0.0) ? 0.0 : cur[KS_ST] / cur[KS_TT];
  const bool bad = !(isfinite(alpha) && isfinite(omega));
  if (blockIdx.x == 0 && threadIdx.x == 0) {          // the other bank: nobody reads it during this kernel
    double* nxt = scal + (1 - p) * KS_BANK;
    nxt[KS_RHO] = bad ? nan("") : cur[KS_SRP] - omega * cur[KS_TRP];
    nxt[KS_RR] = bad ? nan("") : fmax(0.0, cur[KS_SS] - 2.0 * omega * cur[KS_ST] + omega * omega * cur[KS_TT]);
  }
  if (bad) return;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (long long)gridDim.x * blockDim.x) {
    const double s = S_own[i];
    x_own[i] += alpha * P_own[i] + omega * s;
    R[i] = s - omega * T[i];
  }
}

// =====================================================================================================================
// host: cycle
// =====================================================================================================================
template <typename VT>
const VT* level_values(const MgLevel& L);
template <>
const double* level_values<double>(const MgLevel& L) { return L.val; }
template <>
const float* level_values<float>(const MgLevel& L) { return L.valf.p; }

// Host side of the cycle operations: launch the kernel, or -- while the coarse-level program is being recorded -- append a step.
template <typename VT, int MODE>
void launch_sweep(Ctx* c, Mg& G, const MgLevel& L, int level, const double* x_ext, const double* b, double* out, double* out2 = nullptr,
                  const int* agg = nullptr, const double* ec = nullptr, double w = 0.0) {
  if (G.rec) {   // the recorded program only knows the unfused operations (mg_cycle does not fuse while recording)
    G.rec->push_back(MgStep{MODE == 0 ? MG_OP_SWEEP : MG_OP_RESID, L.n, L.off, 0, L.slice_ptr, L.col, level_values<VT>(L), x_ext, b, L.dblk.p, out, 0.0});
    return;
  }
  const int grid = grid_for(c, (long long)L.n_slices * 32, kBlock);
  UFE_LAUNCH(c, level == 0 ? TC_MG_FINE : TC_MG_COARSE);
  if (level == 0) k_mg_sweep<VT, MODE, 2><<<grid, kBlock, 0, c->stream>>>(L.n, L.off, L.slice_ptr, L.col, level_values<VT>(L), x_ext, b, L.dblk.p, out, out2, agg, ec, w);
  else k_mg_sweep<VT, MODE, 8><<<grid, kBlock, 0, c->stream>>>(L.n, L.off, L.slice_ptr, L.col, level_values<VT>(L), x_ext, b, L.dblk.p, out, out2, agg, ec, w);
}
inline void op_smooth0(Ctx* c, Mg& G, int level, int n, const double* b, const double* dblk, double* x_own) {
  if (G.rec) { G.rec->push_back(MgStep{MG_OP_SMOOTH0, n, 0, 0, nullptr, nullptr, nullptr, nullptr, b, dblk, x_own, 0.0}); return; }
  UFE_LAUNCH(c, level == 0 ? TC_MG_VEC : TC_MG_COARSE);
  k_mg_smooth0<<<grid_for(c, n, kBlock), kBlock, 0, c->stream>>>(n, b, dblk, x_own);
}
inline void op_restrict(Ctx* c, Mg& G, int level, int nc, const int* mem_ptr, const int* mem, const double* r, double* bc,
                        const double* dblk_c = nullptr, double* xc = nullptr) {
  if (G.rec) { G.rec->push_back(MgStep{MG_OP_RESTRICT, nc, 0, 0, mem_ptr, mem, nullptr, r, nullptr, nullptr, bc, 0.0}); return; }
  UFE_LAUNCH(c, level == 0 ? TC_MG_VEC : TC_MG_COARSE);
  k_mg_restrict<<<grid_for(c, nc, kBlock), kBlock, 0, c->stream>>>(nc, mem_ptr, mem, r, bc, dblk_c, xc);
}
inline void op_prolong(Ctx* c, Mg& G, int level, int n, const int* agg, const double* ec, double w, double* x_own) {
  if (G.rec) { G.rec->push_back(MgStep{MG_OP_PROLONG, n, 0, 0, agg, nullptr, nullptr, ec, nullptr, nullptr, x_own, w}); return; }
  UFE_LAUNCH(c, level == 0 ? TC_MG_VEC : TC_MG_COARSE);
  k_mg_prolong<<<grid_for(c, n, kBlock), kBlock, 0, c->stream>>>(n, agg, ec, w, x_own);
}
inline void op_axpy(Ctx* c, Mg& G, int n2, const double* a, double* y) {
  if (G.rec) { G.rec->push_back(MgStep{MG_OP_AXPY, n2, 0, 0, nullptr, nullptr, nullptr, a, nullptr, nullptr, y, 0.0}); return; }
  UFE_LAUNCH(c, TC_MG_COARSE);
  k_mg_axpy<<<grid_for(c, n2, kBlock), kBlock, 0, c->stream>>>(n2, a, y);
}
inline void op_dense(Ctx* c, Mg& G, const double* b, double* x) {
  if (G.rec) { G.rec->push_back(MgStep{MG_OP_DENSE, G.dense_n, 0, 0, nullptr, nullptr, G.dense_inv.p, nullptr, b, nullptr, x, 0.0}); return; }
  UFE_LAUNCH(c, TC_MG_COARSE);
  k_mg_dense_solve<<<1, std::min(1024, ((G.dense_n + 31) / 32) * 32), sizeof(double) * G.dense_n, c->stream>>>(G.dense_n, G.dense_inv.p, b, x);
}

template <typename VT>
int mg_cycle(Ctx* c, Mg& G, int l, const double* b, double* out, bool x1_ready = false);

// Levels >= 1 are latency-bound, so neighbouring operations are fused there whenever the data flow allows (three launches per level visit
// instead of five): restriction + the coarse level's first sweep, second-visit residual + its first sweep, prolongation + first post-sweep.
// Not while recording the persistent program (it only knows the unfused operations).
inline bool mg_fuse(const Mg& G) { return G.p.fuse != 0 && G.rec == nullptr; }

// coarse-grid correction of level l: lv[l+1].sol ~= A_{l+1}^-1 lv[l+1].b, with a second visit on the first w_levels coarse levels
// (x1_ready: the producer of lv[l+1].b already wrote the first sweep Dw b into lv[l+1].xa)
template <typename VT>
int mg_coarse_correction(Ctx* c, Mg& G, int l, bool x1_ready = false) {
  MgLevel& Cn = G.lv[l + 1];
  const int last = (int)G.lv.size() - 1;
  UFE_TRY(mg_cycle<VT>(c, G, l + 1, Cn.b.p, Cn.sol.p, x1_ready));
  if (l + 1 < last && l + 1 <= G.p.w_levels) {   // e += cycle(b_c - A_c e)
    if (mg_fuse(G)) {   // the residual kernel also does the first sweep of the second visit
      launch_sweep<VT, 2>(c, G, Cn, l + 1, Cn.sol.p, Cn.b.p, Cn.b2.p, Cn.xa.p + 2 * Cn.off);
      UFE_TRY(mg_cycle<VT>(c, G, l + 1, Cn.b2.p, Cn.sol2.p, true));
    } else {
      launch_sweep<VT, 1>(c, G, Cn, l + 1, Cn.sol.p, Cn.b.p, Cn.b2.p);
      UFE_TRY(mg_cycle<VT>(c, G, l + 1, Cn.b2.p, Cn.sol2.p));
    }
    op_axpy(c, G, 2 * Cn.n, Cn.sol2.p, Cn.sol.p);
  }
  return 0;
}

// records the sub-cycle of the tail level for both (right-hand side, solution) pairs its parent uses and uploads the step lists
template <typename VT>
int mg_tail_build(Ctx* c, Mg& G) {
  G.tail_len[0] = G.tail_len[1] = 0;
  const int t = G.tail_level;
  if (t <= 0 || t >= (int)G.lv.size()) return 0;
  MgLevel& L = G.lv[t];
  std::vector<MgStep> all, steps;
  for (int which = 0; which < 2; ++which) {
    steps.clear();
    G.rec = &steps;
    const int rc = mg_cycle<VT>(c, G, t, which ? L.b2.p : L.b.p, which ? L.sol2.p : L.sol.p, false);
    G.rec = nullptr;
    if (rc) return rc;
    G.tail_off[which] = (int)all.size();
    all.insert(all.end(), steps.begin(), steps.end());
  }
  UFE_TRY(G.tail_prog.alloc(all.size() * sizeof(MgStep)));
  UFE_CUDA(cudaMemcpyAsync(G.tail_prog.p, all.data(), all.size() * sizeof(MgStep), cudaMemcpyHostToDevice, c->stream));
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  G.tail_len[0] = G.tail_off[1]; G.tail_len[1] = (int)all.size() - G.tail_off[1];
  return 0;
}

// The same, entered from the fine level: every launch on levels >= 1 is a small, latency-bound kernel, so the whole correction is
// captured once as a CUDA graph and replayed (one timed region of class mg_coarse instead of ~100 individually timed launches).
template <typename VT>
int mg_coarse_from_fine(Ctx* c, Mg& G) {
  cudaStream_t st = c->stream;
  LaunchTimer region(c, TC_MG_COARSE, false);
  c->timing_suppress++;
  int rc = 0;
  if (G.tail_level > 0 && G.tail_len[0] == 0 && (rc = mg_tail_build<VT>(c, G)) != 0) { c->timing_suppress--; return rc; }
  if (!G.p.use_graph) {
    rc = mg_coarse_correction<VT>(c, G, 0);
  } else if (G.p.use_graph == 2) {   // one persistent cooperative kernel walks the recorded step list
    if (G.program_n == 0) {
      std::vector<MgStep> steps;
      G.rec = &steps;
      rc = mg_coarse_correction<VT>(c, G, 0);
      G.rec = nullptr;
      int occ = 0;
      cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_mg_program<VT>, kBlock, 0);
      if (rc == 0 && (e != cudaSuccess || occ < 1)) { set_error("multigrid: persistent coarse kernel cannot be made resident (%s)", cudaGetErrorString(e)); rc = 1; }
      if (rc == 0) {
        int bps = 2;
        if (const char* ev = getenv("UFE_B200_MG_BPS")) bps = atoi(ev);
        G.program_grid = c->sm_count * std::max(1, std::min(occ, bps));
        if (G.program.alloc(steps.size() * sizeof(MgStep)) != 0) rc = 1;
        else if (cudaMemcpyAsync(G.program.p, steps.data(), steps.size() * sizeof(MgStep), cudaMemcpyHostToDevice, st) != cudaSuccess ||
                 cudaStreamSynchronize(st) != cudaSuccess) { set_error("multigrid: upload of the coarse-level program failed"); rc = 1; }
        else G.program_n = (int)steps.size();
      }
    }
    if (rc == 0) {
      const MgStep* d_steps = reinterpret_cast<const MgStep*>(G.program.p);
      int nsteps = G.program_n;
      void* args[] = {(void*)&d_steps, (void*)&nsteps};
      cudaError_t e = cudaLaunchCooperativeKernel((void*)k_mg_program<VT>, dim3(G.program_grid), dim3(kBlock), args, 0, st);
      c->launches += 1;
      if (e != cudaSuccess) { set_error("multigrid: cooperative launch of the coarse-level program failed: %s", cudaGetErrorString(e)); rc = 1; }
    }
  } else {
    if (!G.graph_exec) {
      const long long l0 = c->launches;
      cudaError_t e = cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal);
      if (e == cudaSuccess) {
        rc = mg_coarse_correction<VT>(c, G, 0);
        cudaError_t e2 = cudaStreamEndCapture(st, &G.graph);
        if (rc == 0 && e2 == cudaSuccess) e2 = cudaGraphInstantiate(&G.graph_exec, G.graph, 0);
        if (e2 != cudaSuccess) e = e2;
      }
      G.graph_launches = c->launches - l0;
      c->launches = l0;
      if (e != cudaSuccess || rc != 0) {
        c->timing_suppress--;
        G.drop_graph();
        if (rc == 0) { set_error("multigrid: CUDA graph capture of the coarse-grid correction failed: %s", cudaGetErrorString(e)); rc = 1; }
        return rc;
      }
    }
    cudaError_t e = cudaGraphLaunch(G.graph_exec, st);
    c->launches += G.graph_launches;
    if (e != cudaSuccess) { set_error("multigrid: cudaGraphLaunch failed: %s", cudaGetErrorString(e)); rc = 1; }
  }
  c->timing_suppress--;
  return rc;
}

// solves (approximately) A_l x = b;  the result is written to `out` (node t at out[2t..2t+1])
template <typename VT>
int mg_cycle(Ctx* c, Mg& G, int l, const double* b, double* out, bool x1_ready) {
  MgLevel& L = G.lv[l];
  const MgParams& P = G.p;
  cudaStream_t st = c->stream;
  const int last = (int)G.lv.size() - 1;
  if (l == G.tail_level && l > 0 && !G.rec && G.tail_len[0] > 0) {   // this level and everything below it: one single-CTA kernel
    const int which = (b == L.b.p && out == L.sol.p) ? 0 : (b == L.b2.p && out == L.sol2.p) ? 1 : -1;
    if (which >= 0) {
      UFE_LAUNCH(c, TC_MG_COARSE);
      const MgStep* prog = reinterpret_cast<const MgStep*>(G.tail_prog.p) + G.tail_off[which];
      if (P.tail_cluster > 1) UFE_TRY(launch_tail_cluster<VT>(c, P.tail_cluster, prog, G.tail_len[which]));
      else k_mg_tail<VT><<<1, kTailBlock, 0, st>>>(prog, G.tail_len[which]);
      return 0;
    }
  }
  double* xa = L.xa.p;
  double* xb = L.xb.p;
  if (l == last) {
    if (G.dense_n > 0) {
      op_dense(c, G, b, out);
      return 0;
    }
    if (P.coarsest_cluster > 1 && !G.rec && G.coarsest_smem > 0 && P.coarse_sweeps >= 1 && L.off == 0) {   // all sweeps of this visit in one cluster kernel
      UFE_LAUNCH(c, TC_MG_COARSE);
      // Chebyshev steps need d_k of every own row in registers: at most 4 slices per warp
      const int per = ((L.n_slices + P.coarsest_cluster - 1) / P.coarsest_cluster), warps = std::max(1, std::min(kCoarsestBlock, P.coarsest_threads) / 32);
      const double* lmax = (P.coarsest_cheb && (per + warps - 1) / warps <= 4) ? reinterpret_cast<const double*>(G.lmax_bits.p) : nullptr;
      if (P.coarsest_dsmem && G.coarsest_smem + coarsest_cluster_xbytes(L.n) <= 200 * 1024)
        UFE_TRY((launch_coarsest_cluster<VT, true>(c, P.coarsest_cluster, P.coarsest_threads, G.coarsest_smem, L, level_values<VT>(L), b, xa, xb, out, P.coarse_sweeps, lmax, P.cheb_ratio)));
      else
        UFE_TRY((launch_coarsest_cluster<VT, false>(c, P.coarsest_cluster, P.coarsest_threads, G.coarsest_smem, L, level_values<VT>(L), b, xa, xb, out, P.coarse_sweeps, lmax, P.cheb_ratio)));
      return 0;
    }
    op_smooth0(c, G, l, L.n, b, L.dblk.p, xa + 2 * L.off);
    for (int s = 0; s < P.coarse_sweeps; ++s) {
      double* dst = (s == P.coarse_sweeps - 1) ? out : xb + 2 * L.off;
      launch_sweep<VT, 0>(c, G, L, l, xa, b, dst);
      std::swap(xa, xb);
    }
    return 0;
  }
  MgLevel& Cn = G.lv[l + 1];
  const bool split = l == 0 && G.distributed;   // level 0 is a row block: its iterates need the neighbours' halo before every matrix pass
  // level 0 without pre-smoothing (nu_pre = 0): the iterate is zero and the right-hand side itself is the residual that gets restricted, which
  // saves one of the cycle's two passes over the level-0 matrix
  const bool no_pre = l == 0 && P.nu_pre == 0 && !G.rec;
  const double* res = no_pre ? b : L.r.p;
  if (no_pre) {
    UFE_CUDA(cudaMemsetAsync(xa + 2 * L.off, 0, sizeof(double) * 2 * (size_t)L.n, st));
  } else {
    if (!x1_ready) op_smooth0(c, G, l, L.n, b, L.dblk.p, xa + 2 * L.off);
    for (int s = 1; s < P.nu_pre; ++s) {
      if (split) UFE_TRY(halo_exchange(c, G.rb, xa, 1, 2));
      launch_sweep<VT, 0>(c, G, L, l, xa, b, xb + 2 * L.off);
      std::swap(xa, xb);
    }
    if (split) UFE_TRY(halo_exchange(c, G.rb, xa, 1, 2));
    launch_sweep<VT, 1>(c, G, L, l, xa, b, L.r.p);
  }
  if (split) UFE_CUDA(cudaMemsetAsync(Cn.b.p, 0, sizeof(double) * 2 * (size_t)Cn.n, st));
  // the child's first sweep rides on the restriction unless the child is solved directly or its right-hand side still needs the all-reduce
  const bool child_smooths = !(l + 1 == last && G.dense_n > 0);
  const bool fuse_down = mg_fuse(G) && !split && child_smooths && l >= 1;
  if (fuse_down) op_restrict(c, G, l, L.nc_loc, L.mem_ptr.p, L.mem.p, res, Cn.b.p + 2 * (size_t)L.c_offset, Cn.dblk.p, Cn.xa.p + 2 * Cn.off);
  else op_restrict(c, G, l, L.nc_loc, L.mem_ptr.p, L.mem.p, res, Cn.b.p + 2 * (size_t)L.c_offset);
  if (split) UFE_TRY(allreduce_sum_large(c, Cn.b.p, 2 * (size_t)Cn.n, false));   // every rank now holds the whole level-1 right-hand side
  if (l == 0) UFE_TRY(mg_coarse_from_fine<VT>(c, G));
  else UFE_TRY(mg_coarse_correction<VT>(c, G, l, fuse_down));
  int s0 = 0;
  if (l == 0 && P.adaptive && P.nu_post >= 1 && !G.rec) {   // correction with the residual-minimising step length; doubles as the first post-sweep
    UFE_TRY(G.mr_q.alloc(2 * (size_t)L.n)); UFE_TRY(G.mr_scal.alloc(2)); UFE_TRY(G.mr_partials.alloc(2 * (size_t)c->sm_count * 8));
    if (!G.mr_counter.n) { UFE_TRY(G.mr_counter.alloc(1)); UFE_CUDA(cudaMemsetAsync(G.mr_counter.p, 0, sizeof(unsigned), st)); }
    { UFE_LAUNCH(c, TC_MG_VEC); k_mg_prolong_set<<<grid_for(c, L.n, kBlock), kBlock, 0, st>>>(L.n, L.agg.p, Cn.sol.p, xb + 2 * L.off); }
    if (split) UFE_TRY(halo_exchange(c, G.rb, xb, 1, 2));
    {
      UFE_LAUNCH(c, TC_MG_FINE);
      const int grid = std::min(grid_for(c, (long long)L.n_slices * 32, kBlock), c->sm_count * 8);
      k_mg_aq<VT><<<grid, kBlock, 0, st>>>(L.n, L.slice_ptr, L.col, level_values<VT>(L), xb, res, G.mr_q.p, G.mr_scal.p, G.mr_partials.p, G.mr_counter.p);
    }
    if (split) UFE_TRY(allreduce_sum(c, G.mr_scal.p, 2));
    double* dst = (P.nu_post == 1) ? out : xa + 2 * L.off;
    { UFE_LAUNCH(c, TC_MG_VEC); k_mg_mr_update<<<grid_for(c, L.n, kBlock), kBlock, 0, st>>>(L.n, L.off, xa, xb, res, G.mr_q.p, L.dblk.p, G.mr_scal.p, P.adaptive_max, dst); }
    s0 = 1;
  } else if (mg_fuse(G) && l >= 1 && P.nu_post >= 1) {   // prolongation fused into the first post-sweep (levels >= 1: off = 0, agg covers the whole level)
    double* dst = (P.nu_post == 1) ? out : xb + 2 * L.off;
    launch_sweep<VT, 3>(c, G, L, l, xa, b, dst, nullptr, L.agg.p, Cn.sol.p, P.overcorr);
    std::swap(xa, xb);
    s0 = 1;
  } else {
    op_prolong(c, G, l, L.n, L.agg.p, Cn.sol.p, P.overcorr, xa + 2 * L.off);
  }
  for (int s = s0; s < P.nu_post; ++s) {
    double* dst = (s == P.nu_post - 1) ? out : xb + 2 * L.off;
    if (split) UFE_TRY(halo_exchange(c, G.rb, xa, 1, 2));
    launch_sweep<VT, 0>(c, G, L, l, xa, b, dst);
    std::swap(xa, xb);
  }
  return 0;
}

template <typename VT>
int mg_numeric(Ctx* c, Mg& G) {
  cudaStream_t st = c->stream;
  const int nl = (int)G.lv.size();
  for (int l = 0; l < nl; ++l) {
    MgLevel& L = G.lv[l];
    {
      UFE_LAUNCH(c, TC_MG_SETUP);
      k_mg_diag<VT><<<grid_for(c, L.n, kBlock), kBlock, 0, st>>>(L.n, L.off, L.slice_ptr, L.col, level_values<VT>(L), G.p.omega, L.dblk.p);
    }
    if (l + 1 < nl) {
      MgLevel& Cn = G.lv[l + 1];
      VT* cv = const_cast<VT*>(level_values<VT>(Cn));
      UFE_LAUNCH(c, TC_MG_SETUP);
      k_mg_galerkin<VT, VT><<<grid_for(c, Cn.n_slots, kBlock), kBlock, 0, st>>>(Cn.n_slots, L.gal_ptr.p, L.gal_src.p, level_values<VT>(L), cv);
    }
    if (l == 0 && G.distributed && nl > 1) {   // each rank summed the blocks of its own coarse rows (zeros elsewhere): complete the replicated level 1
      VT* cv = const_cast<VT*>(level_values<VT>(G.lv[1]));
      UFE_TRY(allreduce_sum_large(c, cv, (size_t)G.lv[1].n_slots * 4, sizeof(VT) == 4));
    }
  }
  if (G.p.coarsest_cheb && G.dense_n == 0 && nl > 1) {      // spectral bound of Dw A on the coarsest level for the Chebyshev steps
    MgLevel& L = G.lv[nl - 1];
    UFE_TRY(G.lmax_bits.alloc(1));
    UFE_CUDA(cudaMemsetAsync(G.lmax_bits.p, 0, sizeof(unsigned long long), st));
    UFE_LAUNCH(c, TC_MG_SETUP);
    k_mg_lmax_bound<VT><<<grid_for(c, L.n, kBlock), kBlock, 0, st>>>(L.n, L.slice_ptr, L.col, level_values<VT>(L), L.dblk.p, G.lmax_bits.p);
  }
  if (G.dense_n > 0) {
    MgLevel& L = G.lv[nl - 1];
    const int N = G.dense_n;
    UFE_CUDA(cudaMemsetAsync(G.dense_w.p, 0, sizeof(double) * (size_t)N * 2 * N, st));
    { UFE_LAUNCH(c, TC_MG_SETUP); k_mg_dense_fill<VT><<<grid_for(c, L.n, 128), 128, 0, st>>>(L.n, L.slice_ptr, L.col, level_values<VT>(L), G.dense_w.p); }
    { UFE_LAUNCH(c, TC_MG_SETUP); k_mg_dense_invert<<<1, 1024, sizeof(double) * N, st>>>(N, G.dense_w.p, G.dense_inv.p); }
  }
  UFE_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace

// =====================================================================================================================
// public entry points
// =====================================================================================================================
int mg_build(Ctx* c, Mg& G, const BlkMat& B, const Range& rb) {
  const int off_tri = rb.off(), n_ext_tri = rb.n_ext();
  G.distributed = c->nranks > 1;
  G.rb = rb;
  const auto t_start = std::chrono::steady_clock::now();
  G.valid = false;
  G.drop_graph();
  if (const char* e = getenv("UFE_B200_MG_GRAPH")) G.p.use_graph = atoi(e);   // debugging aid: 0 = plain launches
  if (const char* e = getenv("UFE_B200_MG_DOUBLE")) G.p.double_from = atoi(e);
  if (const char* e = getenv("UFE_B200_MG_FUSE")) G.p.fuse = atoi(e);
  G.lv.clear();
  G.lv.reserve(G.p.max_levels + 1);
  G.gather_entries = 0;
  // level 0 = the Krylov solver's own matrix
  HostLevel H;
  H.n = B.n_tri; H.off = off_tri;
  H.slice_ptr.resize((size_t)B.n_slices + 1);
  H.col.resize((size_t)B.n_slots);
  UFE_CUDA(cudaMemcpyAsync(H.slice_ptr.data(), B.slice_ptr.p, sizeof(int) * H.slice_ptr.size(), cudaMemcpyDeviceToHost, c->stream));
  UFE_CUDA(cudaMemcpyAsync(H.col.data(), B.col.p, sizeof(int) * H.col.size(), cudaMemcpyDeviceToHost, c->stream));
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  {
    G.lv.emplace_back();
    MgLevel& L = G.lv.back();
    L.n = B.n_tri; L.n_ext = n_ext_tri; L.off = off_tri; L.n_slices = B.n_slices; L.n_slots = B.n_slots;
    L.slice_ptr = B.slice_ptr.p; L.col = B.col.p; L.val = B.val.p;
  }
  while (true) {
    MgLevel& L = G.lv.back();
    const bool small = (L.n <= G.p.dense_max || L.n <= G.p.stop_nodes) && !(G.distributed && G.lv.size() == 1);   // a split level 0 is never the coarsest
    if (small || (int)G.lv.size() >= G.p.max_levels) break;
    HostTransfer T;
    const bool split = G.distributed && G.lv.size() == 1;
    if (split) UFE_TRY(build_transfer_distributed(c, H, rb, T));
    else build_transfer(H, T, (G.p.double_from > 0 && (int)G.lv.size() - 1 >= G.p.double_from) ? 2 : 1);
    if (!split && (T.nc >= L.n || T.nc * 10 > L.n * 9)) break;   // coarsening stalled: stop here, smooth on this level
    L.nc = T.nc; L.nc_loc = T.nc_loc; L.c_offset = T.offset;
    UFE_TRY(upload(c, L.agg, T.agg)); UFE_TRY(upload(c, L.mem_ptr, T.mem_ptr)); UFE_TRY(upload(c, L.mem, T.mem));
    UFE_TRY(upload(c, L.gal_ptr, T.gal_ptr)); UFE_TRY(upload(c, L.gal_src, T.gal_src));
    G.gather_entries += (long long)T.gal_src.size();
    UFE_CHECK(T.coarse_slots * 4 < 2147483647LL, "mg_build: coarse level too large for 32-bit value offsets");
    G.lv.emplace_back();
    MgLevel& Cn = G.lv.back();
    Cn.n = T.nc; Cn.n_ext = T.nc; Cn.off = 0; Cn.n_slices = (T.nc + 31) / 32; Cn.n_slots = T.coarse_slots;
    UFE_TRY(upload(c, Cn.slice_ptr_buf, T.coarse.slice_ptr)); UFE_TRY(upload(c, Cn.col_buf, T.coarse.col));
    Cn.slice_ptr = Cn.slice_ptr_buf.p; Cn.col = Cn.col_buf.p;
    UFE_CUDA(cudaStreamSynchronize(c->stream));   // the host vectors of T die at the end of this iteration
    H = std::move(T.coarse);
  }
  // work vectors and value storage
  for (size_t l = 0; l < G.lv.size(); ++l) {
    MgLevel& L = G.lv[l];
    const size_t ne = 2 * (size_t)L.n_ext, no = 2 * (size_t)L.n;
    UFE_TRY(L.dblk.alloc(4 * (size_t)L.n));
    UFE_TRY(L.xa.alloc(ne)); UFE_TRY(L.xb.alloc(ne)); UFE_TRY(L.r.alloc(no));
    UFE_CUDA(cudaMemsetAsync(L.xa.p, 0, sizeof(double) * ne, c->stream));   // halo entries stay zero for ever
    UFE_CUDA(cudaMemsetAsync(L.xb.p, 0, sizeof(double) * ne, c->stream));
    if (l > 0) {
      UFE_TRY(L.b.alloc(no)); UFE_TRY(L.b2.alloc(no)); UFE_TRY(L.sol.alloc(no)); UFE_TRY(L.sol2.alloc(no));
      if (G.p.fp32) {
        UFE_TRY(L.valf.alloc((size_t)L.n_slots * 4));
        UFE_CUDA(cudaMemsetAsync(L.valf.p, 0, sizeof(float) * (size_t)L.n_slots * 4, c->stream));
      } else {
        UFE_TRY(L.val_buf.alloc((size_t)L.n_slots * 4));
        UFE_CUDA(cudaMemsetAsync(L.val_buf.p, 0, sizeof(double) * (size_t)L.n_slots * 4, c->stream));
        L.val = L.val_buf.p;
      }
    } else if (G.p.fp32) {
      UFE_TRY(L.valf.alloc((size_t)L.n_slots * 4));
      UFE_CUDA(cudaMemsetAsync(L.valf.p, 0, sizeof(float) * (size_t)L.n_slots * 4, c->stream));
    }
  }
  // the first coarse level small enough for one CTA (never level 1: its right-hand side comes from the fine level / the all-reduce)
  G.tail_level = 0;
  if (const char* e = getenv("UFE_B200_MG_TAIL")) G.p.tail_max = atoi(e);
  for (int l = 2; l < (int)G.lv.size() && G.p.tail_max > 0; ++l)
    if (G.lv[l].n <= G.p.tail_max) { G.tail_level = l; break; }
  MgLevel& Lc = G.lv.back();
  G.dense_n = 0;
  if (Lc.n <= G.p.dense_max && 2 * (size_t)Lc.n * sizeof(double) <= 40 * 1024) {
    G.dense_n = 2 * Lc.n;
    UFE_TRY(G.dense_w.alloc((size_t)G.dense_n * 2 * G.dense_n));
    UFE_TRY(G.dense_inv.alloc((size_t)G.dense_n * G.dense_n));
  }
  // the sweeps of a coarsest level that is not solved directly: one cluster kernel per visit if every CTA's share of the matrix fits its shared memory
  G.coarsest_smem = 0;
  if (G.p.coarsest_cluster > 1 && G.dense_n == 0 && G.lv.size() > 1 && (int)H.slice_ptr.size() == Lc.n_slices + 1) {
    const size_t need = G.p.fp32 ? coarsest_cluster_smem<float>(H.slice_ptr, Lc.n, G.p.coarsest_cluster) : coarsest_cluster_smem<double>(H.slice_ptr, Lc.n, G.p.coarsest_cluster);
    if (need > 0 && need <= 200 * 1024) G.coarsest_smem = need;
  }
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  G.setup_host_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_start).count();
  G.valid = true;
  return 0;
}

int mg_update_values(Ctx* c, Mg& G, const BlkMat& B, const DevCsr& A) {
  UFE_CHECK(G.valid, "mg_update_values: hierarchy not built");
  if (G.p.fp32) {
    {
      UFE_LAUNCH(c, TC_MG_SETUP);
      k_mg_scatter_f32<<<grid_for(c, A.nnz, 256), 256, 0, c->stream>>>(A.nnz, B.map.p, A.val, G.lv[0].valf.p);
    }
    return mg_numeric<float>(c, G);
  }
  G.lv[0].val = B.val.p;
  return mg_numeric<double>(c, G);
}

int mg_apply(Ctx* c, Mg& G, const double* r, double* z) {
  return G.p.fp32 ? mg_cycle<float>(c, G, 0, r, z) : mg_cycle<double>(c, G, 0, r, z);
}

int bicgstab_solve_mg(Ctx* c, Mg& G, const BlkLauncher& A, const Range& rr, const double* d_b, double* d_x_ext, const KrylovParams& kp, int* its) {
  if (!c->krylov) c->krylov = new Krylov();
  Krylov& K = *c->krylov;
  const int m = rr.n_loc(), off = rr.off();
  UFE_TRY(K.ensure(c, m, rr.n_ext()));
  UFE_TRY(K.W1.alloc(m)); UFE_TRY(K.W2.alloc(m));
  cudaStream_t st = c->stream;
  const int gv = grid_for(c, m, kBlock);
  double* x_own = d_x_ext + off;
  double* P_own = K.P.p + off;
  double* S_own = K.S.p + off;
  if (rr.n_ext() > m) {
    UFE_CUDA(cudaMemsetAsync(K.P.p, 0, sizeof(double) * rr.n_ext(), st));
    UFE_CUDA(cudaMemsetAsync(K.S.p, 0, sizeof(double) * rr.n_ext(), st));
  }
  auto spmv_plain = [&](const double* x_ext, double* y) {
    A.spmv(c, 0, m, off, 0, x_ext, nullptr, y, nullptr, nullptr, nullptr, nullptr, nullptr, TC_SPMV_STIFFNESS);
  };
  // Breakdown handling as in bicgstab_solve_mat (krylov_impl.cuh): restart from the last finite iterate.  In addition the recurrence is
  // restarted from the true residual every `restart_every` iterations: at tight tolerances (rtol <= 1e-9) BiCGStab's recursive residual
  // stagnates for dozens of iterations before it breaks down, and a restart recovers two orders of magnitude within a few iterations
  // (profiles/r01_mg_monitor_tight_tolerance.log).  Solves at the configuration's own rtol 1e-4 finish long before the first restart.
  int* h_flags = reinterpret_cast<int*>(c->h_pinned);
  const bool monitor = getenv("UFE_B200_KSP_MONITOR") != nullptr;   // like -ksp_monitor: preconditioned residual norm per iteration
  const int restart_every = 40;
  // several ranks: merged reductions (k_ksmg_dot5); UFE_B200_KSP_MERGED=0/1 overrides (1 on one rank is how the tests compare the two recurrences)
  const char* merged_env = getenv("UFE_B200_KSP_MERGED");
  const bool merged = merged_env ? atoi(merged_env) != 0 : c->nranks > 1;
  int total_its = 0, use_guess = kp.use_guess, breakdowns = 0;
  for (int attempt = 0;; ++attempt) {
    const int maxit = std::min(kp.maxit - total_its, restart_every);
    // W1 = M^-1 b;  W2 = M^-1 (b - A x0) when an initial guess is given
    UFE_TRY(mg_apply(c, G, d_b, K.W1.p));
    const double* Br = K.W1.p;
    if (use_guess) {
      UFE_TRY(halo_exchange(c, rr, d_x_ext));
      spmv_plain(d_x_ext, K.T.p);
      { UFE_LAUNCH(c, TC_KRYLOV_VEC); k_ksmg_sub<<<gv, kBlock, 0, st>>>(m, d_b, K.T.p, K.V.p); }
      UFE_TRY(mg_apply(c, G, K.V.p, K.W2.p));
      Br = K.W2.p;
    }
    {
      UFE_LAUNCH(c, TC_KRYLOV_VEC);
      k_ksmg_init<<<gv, kBlock, 0, st>>>(m, K.W1.p, Br, use_guess ? 0 : 1, K.R.p, K.RP.p, x_own, P_own, K.V.p, K.scal.p, K.partials.p,
                                         K.counter.p, K.flags.p, kp.rtol, kp.abstol);
    }
    UFE_TRY(allreduce_sum(c, K.scal.p + KS_RR0, 2));
    { UFE_LAUNCH(c, TC_KRYLOV_VEC); k_ks_init_finish<<<1, 1, 0, st>>>(K.scal.p); }
    int iter = 0;
    // one multigrid-preconditioned iteration is milliseconds of device work and its kernels do not test the done flag, so the host
    // looks at the convergence flag after every iteration (the Jacobi path batches 8-32 sub-millisecond iterations instead)
    const int check_every = 1;
    bool done = false;
    while (!done && iter < maxit) {
      const int stop_at = (iter + check_every < maxit) ? iter + check_every : maxit;
      for (; iter < stop_at; ++iter) {
        const int p = iter & 1;
        { UFE_LAUNCH(c, TC_KRYLOV_VEC); k_ks_update_p<<<gv, kBlock, 0, st>>>(m, iter, K.R.p, K.V.p, P_own, K.scal.p, K.flags.p); }
        UFE_TRY(halo_exchange(c, rr, K.P.p));
        spmv_plain(K.P.p, K.W1.p);
        UFE_TRY(mg_apply(c, G, K.W1.p, K.V.p));
        { UFE_LAUNCH(c, TC_KRYLOV_VEC); k_ksmg_dot1<<<gv, kBlock, 0, st>>>(m, iter, K.V.p, K.RP.p, K.scal.p, K.partials.p, K.counter.p + 1, K.flags.p); }
        UFE_TRY(allreduce_sum(c, K.scal.p + p * KS_BANK + KS_D1, 1));
        { UFE_LAUNCH(c, TC_KRYLOV_VEC); k_ks_update_s<<<gv, kBlock, 0, st>>>(m, iter, K.R.p, K.V.p, S_own, K.scal.p, K.flags.p); }
        UFE_TRY(halo_exchange(c, rr, K.S.p));
        spmv_plain(K.S.p, K.W1.p);
        UFE_TRY(mg_apply(c, G, K.W1.p, K.T.p));
        if (merged) {
          { UFE_LAUNCH(c, TC_KRYLOV_VEC); k_ksmg_dot5<<<gv, kBlock, 0, st>>>(m, iter, S_own, K.T.p, K.RP.p, K.scal.p, K.partials.p, K.counter.p + 2, K.flags.p); }
          UFE_TRY(allreduce_sum(c, K.scal.p + p * KS_BANK + KS_ST, 5));
          { UFE_LAUNCH(c, TC_KRYLOV_VEC); k_ksmg_update_xr5<<<gv, kBlock, 0, st>>>(m, iter, x_own, P_own, S_own, K.T.p, K.R.p, K.scal.p, K.flags.p); }
        } else {
          { UFE_LAUNCH(c, TC_KRYLOV_VEC); k_ksmg_dot2<<<gv, kBlock, 0, st>>>(m, iter, S_own, K.T.p, K.scal.p, K.partials.p, K.counter.p + 2, K.flags.p); }
          UFE_TRY(allreduce_sum(c, K.scal.p + p * KS_BANK + KS_ST, 2));
          {
            UFE_LAUNCH(c, TC_KRYLOV_VEC);
            k_ks_update_xr<<<gv, kBlock, 0, st>>>(m, iter, x_own, P_own, S_own, K.T.p, K.R.p, K.RP.p, K.scal.p, K.partials.p, K.counter.p + 3, K.flags.p);
          }
          UFE_TRY(allreduce_sum(c, K.scal.p + (1 - p) * KS_BANK + KS_RHO, 2));
        }
      }
      if (iter >= maxit) {
        UFE_LAUNCH(c, TC_KRYLOV_VEC);
        k_ks_final_check<<<1, 1, 0, st>>>(iter, K.scal.p, K.flags.p);
      } else {
        UFE_LAUNCH(c, TC_KRYLOV_VEC);
        k_ks_update_p<<<1, 32, 0, st>>>(0, iter, K.R.p, K.V.p, P_own, K.scal.p, K.flags.p);
      }
      UFE_CUDA(cudaMemcpyAsync(h_flags, K.flags.p, KF_COUNT * sizeof(int), cudaMemcpyDeviceToHost, st));
      UFE_CUDA(cudaStreamSynchronize(st));
      done = h_flags[KF_DONE] != 0;
      if (monitor) {
        double rr[2];
        UFE_CUDA(cudaMemcpy(rr, K.scal.p + (iter & 1) * KS_BANK + KS_RR, sizeof(double), cudaMemcpyDeviceToHost));
        UFE_CUDA(cudaMemcpy(rr + 1, K.scal.p + KS_RR0, sizeof(double), cudaMemcpyDeviceToHost));
        fprintf(stderr, "[ufe_b200 ksp_monitor] attempt %d it %4d  ||B r|| %.6e  ||B b|| %.6e\n", attempt, total_its + iter, sqrt(rr[0]), sqrt(rr[1]));
      }
    }
    UFE_CUDA(cudaGetLastError());
    total_its += h_flags[KF_ITS];
    if (h_flags[KF_REASON] == 1) ++breakdowns;
    if (h_flags[KF_REASON] == 0 || breakdowns > 3 || total_its >= kp.maxit) break;
    use_guess = 1;
  }
  *its = total_its;
  return 0;
}

}  // namespace ufe

// ---- host-logic hooks (no GPU involved): the pattern phase of one level, for the CPU test-suite ------------------------------------
static thread_local ufe::HostTransfer* g_host_transfer = nullptr;

extern "C" int ufe_b200_mg_host_transfer(int n, int off, const int32_t* slice_ptr, const int32_t* col, int* nc, long long* coarse_slots,
                                         long long* gather_entries) {
  using namespace ufe;
  UFE_CHECK(n > 0 && slice_ptr && col, "ufe_b200_mg_host_transfer: empty pattern");
  HostLevel F;
  F.n = n; F.off = off;
  const int nsl = (n + 31) / 32;
  F.slice_ptr.assign(slice_ptr, slice_ptr + nsl + 1);
  F.col.assign(col, col + (size_t)F.slice_ptr[nsl] * 32);
  delete g_host_transfer;
  g_host_transfer = new HostTransfer();
  build_transfer(F, *g_host_transfer);
  if (nc) *nc = g_host_transfer->nc;
  if (coarse_slots) *coarse_slots = g_host_transfer->coarse_slots;
  if (gather_entries) *gather_entries = (long long)g_host_transfer->gal_src.size();
  return 0;
}

extern "C" int ufe_b200_mg_host_transfer_get(int32_t* agg, int32_t* coarse_slice_ptr, int32_t* coarse_col, int32_t* gal_ptr, int32_t* gal_src) {
  using namespace ufe;
  UFE_CHECK(g_host_transfer != nullptr, "ufe_b200_mg_host_transfer_get: call ufe_b200_mg_host_transfer first");
  const HostTransfer& T = *g_host_transfer;
  auto copy = [](int32_t* dst, const std::vector<int>& v) { if (dst && !v.empty()) std::memcpy(dst, v.data(), sizeof(int) * v.size()); };
  copy(agg, T.agg); copy(coarse_slice_ptr, T.coarse.slice_ptr); copy(coarse_col, T.coarse.col); copy(gal_ptr, T.gal_ptr); copy(gal_src, T.gal_src);
  delete g_host_transfer;
  g_host_transfer = nullptr;
  return 0;
}

// the same pattern phase for a matrix whose rows are split into parts (= ranks), emulated in one process: part p owns the nodes
// off[p] .. off[p] + n[p] - 1 of one global numbering (its extended range is the whole range), its sliced-ELL pattern starts at
// slice_ptr[sp_start[p]] / col[col_start[p]].  Exactly the code path of build_transfer_distributed with the exchanges replaced by copies.
struct HostPartsResult {
  std::vector<int> agg_glob, gal_ptr, gal_src, gal_src_start;
  ufe::HostLevel coarse;
};
static thread_local HostPartsResult* g_host_parts = nullptr;

extern "C" int ufe_b200_mg_host_transfer_parts(int nparts, int n_total, const int32_t* n, const int32_t* off, const int32_t* sp_start,
                                               const int32_t* col_start, const int32_t* slice_ptr, const int32_t* col, int* nc,
                                               long long* coarse_slots, long long* gather_entries) {
  using namespace ufe;
  UFE_CHECK(nparts > 0 && n && off && sp_start && col_start && slice_ptr && col, "ufe_b200_mg_host_transfer_parts: NULL argument");
  std::vector<HostLevel> F(nparts);
  std::vector<LocalAgg> A(nparts);
  std::vector<int> offset(nparts + 1, 0);
  for (int p = 0; p < nparts; ++p) {
    F[p].n = n[p]; F[p].off = off[p];
    const int nsl = (n[p] + 31) / 32;
    F[p].slice_ptr.assign(slice_ptr + sp_start[p], slice_ptr + sp_start[p] + nsl + 1);
    F[p].col.assign(col + col_start[p], col + col_start[p] + (size_t)F[p].slice_ptr[nsl] * 32);
    aggregate_part(F[p], A[p]);
    offset[p + 1] = offset[p] + A[p].nc_loc;
  }
  const int nc_glob = offset[nparts];
  delete g_host_parts;
  g_host_parts = new HostPartsResult();
  HostPartsResult& Rs = *g_host_parts;
  Rs.agg_glob.assign(n_total, -1);
  for (int p = 0; p < nparts; ++p)
    for (int i = 0; i < n[p]; ++i) Rs.agg_glob[off[p] + i] = offset[p] + A[p].agg[i];
  std::vector<LocalRows> R(nparts);
  std::vector<int> all_lens, all_cols, cptr;
  for (int p = 0; p < nparts; ++p) {
    build_local_rows(F[p], A[p], Rs.agg_glob, offset[p], nc_glob, R[p]);
    for (int I = 0; I < A[p].nc_loc; ++I) all_lens.push_back(R[p].cptr[I + 1] - R[p].cptr[I]);
    all_cols.insert(all_cols.end(), R[p].ccol.begin(), R[p].ccol.end());
  }
  rows_from_parts(all_lens, cptr);
  ell_from_rows(nc_glob, cptr, all_cols, Rs.coarse);
  for (int p = 0; p < nparts; ++p) {
    std::vector<int> gp, gs;
    build_gather(F[p], A[p], offset[p], R[p], Rs.coarse, gp, gs);
    Rs.gal_src_start.push_back((int)Rs.gal_src.size());
    Rs.gal_ptr.insert(Rs.gal_ptr.end(), gp.begin(), gp.end());
    Rs.gal_src.insert(Rs.gal_src.end(), gs.begin(), gs.end());
  }
  Rs.gal_src_start.push_back((int)Rs.gal_src.size());
  if (nc) *nc = nc_glob;
  if (coarse_slots) *coarse_slots = (long long)Rs.coarse.col.size();
  if (gather_entries) *gather_entries = (long long)Rs.gal_src.size();
  return 0;
}

extern "C" int ufe_b200_mg_host_transfer_parts_get(int32_t* agg_glob, int32_t* coarse_slice_ptr, int32_t* coarse_col, int32_t* gal_ptr, int32_t* gal_src,
                                                   int32_t* gal_src_start) {
  using namespace ufe;
  UFE_CHECK(g_host_parts != nullptr, "ufe_b200_mg_host_transfer_parts_get: call ufe_b200_mg_host_transfer_parts first");
  const HostPartsResult& T = *g_host_parts;
  auto copy = [](int32_t* dst, const std::vector<int>& v) { if (dst && !v.empty()) std::memcpy(dst, v.data(), sizeof(int) * v.size()); };
  copy(agg_glob, T.agg_glob); copy(coarse_slice_ptr, T.coarse.slice_ptr); copy(coarse_col, T.coarse.col); copy(gal_ptr, T.gal_ptr);
  copy(gal_src, T.gal_src); copy(gal_src_start, T.gal_src_start);
  delete g_host_parts;
  g_host_parts = nullptr;
  return 0;
}
