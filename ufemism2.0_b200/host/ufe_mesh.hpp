// Host-side mesh container and builders (C++17, no CUDA).
//
// Mirrors the subset of UFEMISM's type_mesh that the stress-balance / ice-thickness
// hot path reads (reference: src/UPSY/types/mesh_types.f90:40-131).
//
// Layout contract (identical to what the Fortran host hands over ISO_C_BINDING):
//   * every 2-D array is column-major: A(i,j) of shape (n,m) lives at a[(j-1)*n + (i-1)]
//   * every index stored in an array is 1-based; 0 means "none"
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace ufe {

struct HostMesh {
  // sizes
  int nV = 0, nTri = 0, nE = 0, nC_mem = 32, nz = 12;
  double xmin = 0, xmax = 0, ymin = 0, ymax = 0, tol_dist = 0;

  // primary data (mesh_types.f90:56-68)
  std::vector<double> V;      // (nV,2)
  std::vector<int> nC;        // (nV)
  std::vector<int> C;         // (nV,nC_mem)  ccw, edge-to-edge for border vertices
  std::vector<int> niTri;     // (nV)
  std::vector<int> iTri;      // (nV,nC_mem)  ccw; iTri(vi,k) spans vi,C(vi,k),C(vi,k+1)
  std::vector<int> VBI;       // (nV) 0 free, 1 N, 2 NE, 3 E, 4 SE, 5 S, 6 SW, 7 W, 8 NW
  std::vector<int> Tri;       // (nTri,3) ccw
  std::vector<double> Tricc;  // (nTri,2)
  std::vector<int> TriC;      // (nTri,3) neighbour across from 1st,2nd,3rd vertex

  // secondary data (mesh_types.f90:85-112)
  std::vector<double> A;      // (nV) Voronoi cell area
  std::vector<double> Cw;     // (nV,nC_mem) shared Voronoi boundary length
  std::vector<double> D_x, D_y, D;  // (nV,nC_mem)
  std::vector<int> TriBI;     // (nTri)
  std::vector<double> TriGC;  // (nTri,2)
  std::vector<double> TriA;   // (nTri)
  std::vector<double> E;      // (nE,2)
  std::vector<int> VE;        // (nV,nC_mem)
  std::vector<int> EV;        // (nE,4) [vi,vj,vl,vr]
  std::vector<int> ETri;      // (nE,2) [til,tir]
  std::vector<int> TriE;      // (nTri,3)
  std::vector<int> EBI;       // (nE)
  std::vector<double> zeta;   // (nz)

  std::string last_error;
};

// --- synthetic primary mesh -------------------------------------------------------------
struct SynthParams {
  double xmin, xmax, ymin, ymax;
  double h;            // target edge length [m]
  double jitter;       // jitter amplitude as a fraction of h (0.2 in SURVEY §8d config 3)
  uint64_t seed;       // 20261104 for the bench
  // optional 1-D grading in x (grounding-line refinement stand-in):
  //   local spacing is h * (1 - depth * exp(-((x-x0)/width)^2)),  depth in [0,1)
  double grade_x0 = 0, grade_width = 0, grade_depth = 0;
  int nC_mem = 32;
  int nz = 12;
  int zeta_kind = 0;   // 0 regular, 1 irregular_log (R = zeta_R)
  double zeta_R = 10.0;
};

// Jittered near-equilateral triangulation, Delaunay-repaired by Lawson flips, vertices and
// triangles sorted by x (the ordering UFEMISM relies on, mesh_parallelisation.f90:14-22).
bool make_synthetic_mesh(const SynthParams& p, HostMesh& m);

// The reference's 5-vertex dummy mesh (mesh_dummy_meshes.f90:23-120), used as a hand-checkable fixture.
bool make_dummy_mesh_5(double xmin, double xmax, double ymin, double ymax, HostMesh& m);

// Build nC,C,niTri,iTri,TriC,VBI,Tricc from V + Tri (0-based scratch input).
bool build_connectivity(HostMesh& m);

// --- secondary data: edges, TriBI, Voronoi areas, connection widths/lengths, centres, zeta ---
// (mesh_secondary.f90:31-70; edges/mesh_edges.f90:19-237; mesh_utilities.f90:72-372)
bool calc_secondary(HostMesh& m, int zeta_kind, double zeta_R);

// --- ownership (mpi_distributed_memory.f90:47-73) ---
void partition_list(int ntot, int i, int n, int& i1, int& i2);

struct HaloRanges {  // type_par_arr_info, parallel_array_info_type.f90:9-21 (1-based, i2 < i1 = empty)
  int i1, i2, i1_nih, i2_nih, i1_hle, i2_hle, i1_hli, i2_hli, i1_hre, i2_hre, i1_hri, i2_hri;
};
// determine_halos (mesh_parallelisation.f90:282-408) with one process per node (node == GPU)
void determine_halos(const HostMesh& m, int rank, int nranks, HaloRanges& a, HaloRanges& b, HaloRanges& c);

}  // namespace ufe
