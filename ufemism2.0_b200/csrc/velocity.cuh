// Device state of the SSA/DIVA velocity solve (type_ice_velocity_solver_DIVA / _SSA,
// reference: src/UFEMISM/types/ice_model_types.f90:149-197).  See velocity.cu.
#pragma once
#include "blk.cuh"
#include "common.cuh"
#include "mg.cuh"

namespace ufe {

constexpr int kMaxNz = 32;

// Every field is stored over the extended range of its grid (owned + halos): a-fields have a.n_ext() elements,
// b-fields b.n_ext(); 3-D fields are [nz][n_ext] (the reference's (i,k) column-major layout).
struct Velocity {
  int nz = 0;
  // --- inputs (type_ice_model / type_bed_roughness_model)
  DBuf<double> Hi, Hs, A_flow, A_flow_vav, fraction_gr_b, till_friction_angle, beta_sq, alpha_sq, effective_pressure, tauc0, tauc;
  DBuf<int> mask_grounded_ice, mask_icefree_land;
  bool have_tfa = false, have_beta_sq = false, have_alpha_sq = false, have_neff = false;
  int grounded_ice_exists = 1;
  // --- state on the b-grid
  DBuf<double> u, v, u_base, v_base, tau_bx, tau_by, eta_3D_b, u_prev, v_prev, tau_dx, tau_dy, u_3D, v_3D;
  DBuf<double> N_b, dN_dx_b, dN_dy_b, beta_b, beta_eff_b, F1_3D_b, F2_3D_b;
  // --- state on the a-grid
  DBuf<double> du_dx, du_dy, dv_dx, dv_dy, du_dz_3D, dv_dz_3D, eta_3D_a, eta_vav, N_a, F1_3D_a, F2_3D_a, beta_a, beta_eff_a;
  // --- layer-interleaved work copies ([i][k] instead of the reference's [k][i]) of the two 3-D fields the kernels GATHER through the mesh operators:
  // a neighbour's column is then one 8 nz-byte segment instead of nz sectors (velocity.cu: k_layers_interleave)
  DBuf<double> rzT_b;          // zeta_k / max( visc_eff_min, eta_3D_b( ti,k)): what k_vertex_diva sums into du_dz / dv_dz
  DBuf<double> etaT_a;         // eta_3D_a( vi,k) for the a->b mapping in k_tri_assemble
  // --- prescribed velocities
  DBuf<int> bc_mask;
  DBuf<double> bc_u, bc_v;
  bool have_bc = false;
  std::vector<int32_t> bc_mask_host;   // the mask the current pattern was built for (velocity.cu: upload_bc)
  // --- linear system: stiffness matrix on a pattern built once per (mesh, BC) (K3 of SURVEY 2e)
  DevCsr A;
  DBuf<int> row_kind;          // per matrix row: 0 free, 1 identity (prescribed / 'zero'), 2 'infinite'
  DBuf<double> bb, uv;         // load vector (owned rows), solution (extended buv range)
  // copy rows of BC 'periodic_ISMIP-HOM' / 'infinite_SSA_icestream' (identity rows whose load vector is the previous solution around a
  // displaced point): compact list built with the pattern -- owned matrix row, BC choice, vertex whose Voronoi cell holds the point
  DBuf<int> copy_row, copy_choice, copy_vi, copy_count;
  int n_copy = 0;              // owned copy rows
  bool copy_any = false;       // the BC choices of the current pattern contain a copy choice (on any rank)
  DBuf<double> uv_glob;        // nranks > 1 only: the previous solution on all triangles (the reference gathers u_b_prev globally too)
  BlkMat blk;                  // 2x2-blocked sliced-ELL copy of A that the Krylov solver streams
  bool blk_filled = false;     // the last assembly wrote blk.val itself (no scatter pass needed)
  bool csr_stale = false;      // ... and skipped the CSR values (re-formed from blk.val on demand)
  Mg mg;                       // multigrid preconditioner on the pattern of blk (pc_type = UFE_PC_MG)
  unsigned long long pattern_key = 0;
  bool pattern_valid = false;
  // --- reductions
  DBuf<double> red, partials;
  DBuf<unsigned> counter;
  bool uploaded = false;
  // what the last solves left on the device (read by ufe_b200_set_ice_velocities)
  bool have_strain_2d = false, have_dudz_3d = false, have_sia = false;
  int last_approximation = -1;   // UFE_APPROX_* of the last velocity solve (SIA: solve_SIA)
  int alloc(Ctx* c);
};

void velocity_free(Velocity* v);
// host -> device copies of the velocity solve's inputs on c->stream (no synchronisation); flags select what is copied
enum { VU_STATE_UV = 1, VU_STATE_REST = 2, VU_SKIP_GEOMETRY = 4, VU_ALL = VU_STATE_UV | VU_STATE_REST };
int velocity_upload(Ctx* c, const ufe_b200_ice_inputs* ice, const ufe_b200_velocity_state* st, int flags = VU_ALL);
// device -> host copies of every non-NULL field of st on c->stream, then (sync) a synchronise of that stream
int velocity_download(Ctx* c, ufe_b200_velocity_state* st, bool sync = true);
int ensure_velocity(Ctx* c);
// solve_DIVA_infinite_slab (approximation = 1) / solve_SSA (0) on device-resident fields
int velocity_solve(Ctx* c, const ufe_b200_config* cfg, int approximation, int* n_visc_its, int* n_Axb_its);

}  // namespace ufe
