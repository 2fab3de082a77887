// C entry points of libufe_host.so (input generation + ownership; no CUDA, no oracle).
#include <cstring>
#include <map>
#include <string>

#include "ufe_mesh.hpp"

using ufe::HostMesh;

extern "C" {

void* ufe_host_mesh_synth(double xmin, double xmax, double ymin, double ymax, double h, double jitter,
                          unsigned long long seed, double grade_x0, double grade_width, double grade_depth,
                          int nC_mem, int nz, int zeta_kind, double zeta_R) {
  auto* m = new HostMesh();
  ufe::SynthParams p{xmin, xmax, ymin, ymax, h, jitter, seed, grade_x0, grade_width, grade_depth, nC_mem, nz, zeta_kind, zeta_R};
  if (!ufe::make_synthetic_mesh(p, *m)) { std::string e = m->last_error; *m = HostMesh(); m->last_error = e.empty() ? "failed" : e; }
  return m;
}

void* ufe_host_mesh_dummy5(double xmin, double xmax, double ymin, double ymax) {
  auto* m = new HostMesh();
  if (!ufe::make_dummy_mesh_5(xmin, xmax, ymin, ymax, *m)) { std::string e = m->last_error; *m = HostMesh(); m->last_error = e.empty() ? "failed" : e; }
  return m;
}

// A mesh whose primary data (mesh_types.f90:56-68) the caller already holds -- e.g. one of the reference's own NetCDF meshes -- in the Fortran
// layout; the secondary data follow from calc_secondary exactly as for a generated mesh (calc_all_secondary_mesh_data, mesh_secondary.f90:31-70).
void* ufe_host_mesh_from_primary(int nV, int nTri, int nC_mem, int nz, int zeta_kind, double zeta_R, double xmin, double xmax, double ymin,
                                 double ymax, const double* V, const int* nC, const int* C, const int* niTri, const int* iTri,
                                 const int* VBI, const int* Tri, const double* Tricc, const int* TriC) {
  auto* m = new HostMesh();
  m->nV = nV; m->nTri = nTri; m->nC_mem = nC_mem; m->nz = nz;
  m->xmin = xmin; m->xmax = xmax; m->ymin = ymin; m->ymax = ymax;
  m->V.assign(V, V + (size_t)nV * 2); m->nC.assign(nC, nC + nV); m->C.assign(C, C + (size_t)nV * nC_mem);
  m->niTri.assign(niTri, niTri + nV); m->iTri.assign(iTri, iTri + (size_t)nV * nC_mem); m->VBI.assign(VBI, VBI + nV);
  m->Tri.assign(Tri, Tri + (size_t)nTri * 3); m->Tricc.assign(Tricc, Tricc + (size_t)nTri * 2); m->TriC.assign(TriC, TriC + (size_t)nTri * 3);
  if (!ufe::calc_secondary(*m, zeta_kind, zeta_R)) { std::string e = m->last_error; *m = HostMesh(); m->last_error = e.empty() ? "failed" : e; }
  return m;
}

void ufe_host_mesh_free(void* h) { delete static_cast<HostMesh*>(h); }

const char* ufe_host_mesh_error(void* h) { return static_cast<HostMesh*>(h)->last_error.c_str(); }

// out = {nV, nTri, nE, nC_mem, nz}
void ufe_host_mesh_dims(void* h, int* out) {
  auto* m = static_cast<HostMesh*>(h);
  out[0] = m->nV; out[1] = m->nTri; out[2] = m->nE; out[3] = m->nC_mem; out[4] = m->nz;
}

void ufe_host_mesh_domain(void* h, double* out) {
  auto* m = static_cast<HostMesh*>(h);
  out[0] = m->xmin; out[1] = m->xmax; out[2] = m->ymin; out[3] = m->ymax;
}

// Borrow a pointer to a named array. dtype: 0 = int32, 1 = float64. Returns nullptr for unknown names.
const void* ufe_host_mesh_array(void* h, const char* name, int* dtype, long long* count) {
  auto* m = static_cast<HostMesh*>(h);
#define I(n) if (!std::strcmp(name, #n)) { *dtype = 0; *count = (long long)m->n.size(); return m->n.data(); }
#define D(n) if (!std::strcmp(name, #n)) { *dtype = 1; *count = (long long)m->n.size(); return m->n.data(); }
  D(V) I(nC) I(C) I(niTri) I(iTri) I(VBI) I(Tri) D(Tricc) I(TriC)
  D(A) D(Cw) D(D_x) D(D_y) D(D) I(TriBI) D(TriGC) D(TriA) D(E) I(VE) I(EV) I(ETri) I(TriE) I(EBI) D(zeta)
#undef I
#undef D
  return nullptr;
}

void ufe_host_partition_list(int ntot, int i, int n, int* i1, int* i2) { ufe::partition_list(ntot, i, n, *i1, *i2); }

// out[0:12] vertices, out[12:24] triangles, out[24:36] edges; each
// {i1,i2,i1_nih,i2_nih,i1_hle,i2_hle,i1_hli,i2_hli,i1_hre,i2_hre,i1_hri,i2_hri}
void ufe_host_determine_halos(void* h, int rank, int nranks, int* out) {
  ufe::HaloRanges a, b, c;
  ufe::determine_halos(*static_cast<HostMesh*>(h), rank, nranks, a, b, c);
  std::memcpy(out, &a, sizeof(a)); std::memcpy(out + 12, &b, sizeof(b)); std::memcpy(out + 24, &c, sizeof(c));
}

}  // extern "C"
