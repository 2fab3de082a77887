// Device-resident predictor-corrector time step: the unit "1 time step" = 1 velocity solve + 2 thickness solves.
// Follows iterate_pc_timestep of run_ice_dynamics_model_pc
//   src/UFEMISM/ice_dynamics/time_stepping/predictor_corrector_scheme.f90:118-233 (Robinson et al., 2020, Eqs. 30-31)
// with the three device phases back to back and no field leaving HBM in between.  The host-side stages the
// reference runs between them (alter_ice_thickness, determine_masks, calc_grounded_fractions,
// calc_effective_thickness) run on the device too once ufe_b200_geometry_resident switched them on: masks and grounded
// fractions of the predicted geometry for the velocity solve, fraction_margin of the current geometry for the two thickness solves
// (geometry.cu).  Otherwise masks and sub-grid fractions stay as uploaded and only Hs follows the predicted thickness
// (ice_geometry_basics.f90:33).  alter_ice_thickness and the truncation-error based dt control stay with the host.
#include "geometry.cuh"
#include "mesh.cuh"
#include "thickness.cuh"
#include "velocity.cuh"

namespace ufe {

// The state of the time stepper that lives between calls: one per handle (Ctx::pc), released with the mesh.
struct PcState {
  DBuf<double> dHdt_n, dHdt_nm1;
  DBuf<double> Hi_star;            // pc%Hi_star_np1 of the last step (the truncation error needs it next to Hi_np1)
  DBuf<double> Hi_np1;             // scratch: the corrected thickness before it replaces Hi
  double dt_n = 0.0;
  bool started = false;
  // the reference's loop (ufe_b200_pc_reference_loop; default ON): region%ice%Hi lags behind Hi_prev, Eq. 30 takes the applied rate
  bool reference_loop = true, lag_valid = false;
  double w_lag = 0.0;
  DBuf<double> Hi_lag, Hi_save;
  // alter_ice_thickness inside the step (ufe_b200_alter_resident); off: only ice thinner than C%Hi_min is removed
  bool alter_on = false, alter_have_T = false;
  ufe_b200_alter_params alter;
  double alter_time = 0.0;
  DBuf<double> ref_Hi, ref_Hb, Ti_hom;
};
void pc_state_free(PcState* p) { delete p; }

static PcState& pc_of(Ctx* c) {
  if (!c->pc) c->pc = new PcState();
  return *c->pc;
}

// Start of run_ice_dynamics_model_pc (:88-93): no ice and no thinning where the no-ice mask holds.  Then Eq. 30:
// Hi_star = Hi_prev + dt ((1 + zeta/2) dHdt_n - zeta/2 dHdt_nm1), unmodified (alter_ice_thickness follows, :145)
__global__ void k_pc_predict(int n, int off, double dt, double zeta, const int* __restrict__ mask_noice, double* __restrict__ Hi_prev,
                             const double* __restrict__ dHdt_calc, double* __restrict__ dHdt_n, double* __restrict__ dHdt_nm1,
                             double* __restrict__ Hi_star) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const size_t i = off + r;
    double d = dHdt_calc[i], hp = Hi_prev[i], dm = dHdt_nm1[i];
    if (mask_noice[i]) { d = 0.0; hp = 0.0; dm = 0.0; Hi_prev[i] = 0.0; dHdt_nm1[i] = 0.0; }
    dHdt_n[i] = d;
    Hi_star[i] = hp + dt * ((1.0 + zeta / 2.0) * d - (zeta / 2.0) * dm);
  }
}

// dHi_dt_Hi_n_u_n re-derived from the (possibly altered) Hi_star (:152), geometry of the velocity solve (:156-166)
__global__ void k_pc_rederive(int n, int off, double dt, double zeta, double rho_ratio, const double* __restrict__ Hi_prev,
                              const double* __restrict__ Hi_star, const double* __restrict__ dHdt_nm1, double* __restrict__ dHdt_n,
                              const double* __restrict__ Hb, const double* __restrict__ SL, double* __restrict__ Hi_vel, double* __restrict__ Hs_vel,
                              double* __restrict__ Hi_star_keep) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const size_t i = off + r;
    const double h = Hi_star[i];
    dHdt_n[i] = ((h - Hi_prev[i]) / dt + (zeta / 2.0) * dHdt_nm1[i]) / (1.0 + zeta / 2.0);
    Hi_vel[i] = h;
    Hs_vel[i] = h + fmax(SL[i] - rho_ratio * h, Hb[i]);
    Hi_star_keep[i] = h;
  }
}

// Hi_np1 = Hi_prev + dt/2 (dHdt_n + dHdt_star)  (Eq. 31), unmodified (alter_ice_thickness follows, :233)
__global__ void k_pc_correct(int n, int off, double dt, const int* __restrict__ mask_noice, const double* __restrict__ Hi_prev,
                             const double* __restrict__ dHdt_n, const double* __restrict__ dHdt_star, double* __restrict__ Hi_np1) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const size_t i = off + r;
    double ds = dHdt_star[i];
    if (mask_noice[i]) ds = 0.0;
    Hi_np1[i] = Hi_prev[i] + (dt / 2.0) * (dHdt_n[i] + ds);
  }
}

// the step is accepted: Hi_np1 becomes the model's thickness; the rate handed to the next Eq. 30; the lagged region%ice%Hi
__global__ void k_pc_accept(int n, int off, double dt, double rho_ratio, const double* __restrict__ Hi_np1, const double* __restrict__ dHdt_n,
                            double* __restrict__ dHdt_nm1, const double* __restrict__ Hb, const double* __restrict__ SL, double* __restrict__ Hi,
                            double* __restrict__ Hi_vel, double* __restrict__ Hs_vel, double* __restrict__ Hi_lag, double w_lag) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (long long)gridDim.x * blockDim.x) {
    const size_t i = off + r;
    const double h_old = Hi[i], h = Hi_np1[i];
    Hi[i] = h;
    if (Hi_lag) {   // reference loop: region%ice%Hi as interpolated at the last action time, and the applied rate for the next Eq. 30
      Hi_lag[i] = (1.0 - w_lag) * h_old + w_lag * h;               // ice_dynamics_main.f90:115-121
      dHdt_nm1[i] = (h - h_old) / dt;                              // ice_dynamics_main.f90:140 -> predictor_corrector_scheme.f90:109
    } else {
      dHdt_nm1[i] = dHdt_n[i];                                     // Robinson et al. (2020), Eq. 30 as printed
    }
    Hi_vel[i] = h;
    Hs_vel[i] = h + fmax(SL[i] - rho_ratio * h, Hb[i]);
  }
}

// the three device phases of one predictor-corrector step; `before_velocity` / `after_velocity` let the host-buffer entry
// overlap its transfers with the thickness solves
template <class F1, class F2>
static int pc_step_core(Ctx* c, const ufe_b200_config* cfg, int approximation, double dt, int* n_visc_its, int* n_Axb_its_velocity,
                        int* n_Axb_its_thickness, F1&& before_velocity, F2&& after_velocity) {
  Mesh& M = *c->mesh; Velocity& V = *c->vel; Thickness& T = *c->thick;
  cudaStream_t st = c->stream;
  PcState& P = pc_of(c);
  const size_t na = M.a.n_ext();
  const int n = M.a.n_loc(), grid = grid_for(c, n, 128, 16);
  const double zeta = dt / P.dt_n;
  const double rho_ratio = cfg->ice_density / cfg->seawater_density;
  int its_p = 0, its_c = 0;
  double dt_p = dt, dt_c = dt;
  const bool dynamic = geometry_dynamic(c);
  const bool lagged = P.reference_loop;
  UFE_TRY(P.Hi_star.alloc(na)); UFE_TRY(P.Hi_np1.alloc(na));
  const ufe_b200_alter_params* prm = P.alter_on ? &P.alter : nullptr;
  double* Hi_of_masks = T.Hi.p;
  if (lagged) {   // the predictor works on region%ice%Hi, which the reference leaves at the thickness of the last action time (:129 against :140)
    if (P.Hi_lag.n < na) { UFE_TRY(P.Hi_lag.alloc(na)); UFE_TRY(P.Hi_save.alloc(na)); P.lag_valid = false; }
    if (!P.lag_valid) { UFE_CUDA(cudaMemcpyAsync(P.Hi_lag.p, T.Hi.p, sizeof(double) * na, cudaMemcpyDeviceToDevice, st)); P.lag_valid = true; }
    UFE_CUDA(cudaMemcpyAsync(P.Hi_save.p, T.Hi.p, sizeof(double) * na, cudaMemcpyDeviceToDevice, st));
    UFE_CUDA(cudaMemcpyAsync(T.Hi.p, P.Hi_lag.p, sizeof(double) * na, cudaMemcpyDeviceToDevice, st));
  }
  // effective thickness / margin fraction of region%ice%Hi (:107)
  if (dynamic) UFE_TRY(geometry_for_thickness(c, cfg, T.Hi.p, T.Hb.p, T.SL.p, T.fraction_margin.p));
  // predictor: dHi_dt( Hi, u_n) (:129)
  UFE_TRY(halo_exchange(c, M.b, V.u.p)); UFE_TRY(halo_exchange(c, M.b, V.v.p));
  UFE_TRY(thickness_step(c, cfg, V.u.p, V.v.p, &dt_p, 1, &its_p));
  if (lagged) {   // T.Hi is Hi_prev again; the masks alter_ice_thickness reads are still those of region%ice%Hi (:95)
    UFE_CUDA(cudaMemcpyAsync(T.Hi.p, P.Hi_save.p, sizeof(double) * na, cudaMemcpyDeviceToDevice, st));
    Hi_of_masks = P.Hi_lag.p;
  }
  { UFE_LAUNCH(c, TC_THICKNESS); k_pc_predict<<<grid, 128, 0, st>>>(n, M.a.off(), dt, zeta, T.mask_noice.p, T.Hi.p, T.dHi_dt.p, P.dHdt_n.p, P.dHdt_nm1.p, P.Hi_np1.p); }
  UFE_TRY(geometry_alter(c, cfg, prm, P.alter_time, Hi_of_masks, T.Hi.p, T.Hb.p, T.SL.p, P.ref_Hi.p, P.ref_Hb.p, P.alter_have_T ? P.Ti_hom.p : nullptr, P.Hi_np1.p));   // (:145)
  { UFE_LAUNCH(c, TC_THICKNESS); k_pc_rederive<<<grid, 128, 0, st>>>(n, M.a.off(), dt, zeta, rho_ratio, T.Hi.p, P.Hi_np1.p, P.dHdt_nm1.p, P.dHdt_n.p, T.Hb.p, T.SL.p, V.Hi.p, V.Hs.p, P.Hi_star.p); }
  // velocities for the predicted geometry (:186)
  UFE_TRY(before_velocity());
  if (dynamic) {   // masks and grounded fractions of the predicted geometry (:172-179)
    int any = 0;
    UFE_TRY(geometry_for_velocity(c, cfg, V.Hi.p, T.Hb.p, T.SL.p, V.mask_grounded_ice.p, V.mask_icefree_land.p, V.fraction_gr_b.p, &any));
    if (c->nranks > 1) {   // MPI_ALLREDUCE( ..., MPI_LOR, ...) (DIVA_solver_infinite_slab.f90:88)
      double f = any;
      UFE_CUDA(cudaMemcpyAsync(V.red.p + 4, &f, sizeof(double), cudaMemcpyHostToDevice, st));
      UFE_TRY(allreduce_sum(c, V.red.p + 4, 1));
      UFE_CUDA(cudaMemcpyAsync(&f, V.red.p + 4, sizeof(double), cudaMemcpyDeviceToHost, st));
      UFE_CUDA(cudaStreamSynchronize(st));
      any = f > 0.0;
    }
    V.grounded_ice_exists = any;
  }
  UFE_TRY(velocity_solve(c, cfg, approximation, n_visc_its, n_Axb_its_velocity));
  UFE_TRY(after_velocity());
  // corrector: dHi_dt( Hi_prev, u_np1) (:214); masks and effective thickness are those of Hi_prev again (:198-211)
  if (lagged && dynamic) UFE_TRY(geometry_for_thickness(c, cfg, T.Hi.p, T.Hb.p, T.SL.p, T.fraction_margin.p));
  UFE_TRY(halo_exchange(c, M.b, V.u.p)); UFE_TRY(halo_exchange(c, M.b, V.v.p));
  UFE_TRY(thickness_step(c, cfg, V.u.p, V.v.p, &dt_c, 1, &its_c));
  { UFE_LAUNCH(c, TC_THICKNESS); k_pc_correct<<<grid, 128, 0, st>>>(n, M.a.off(), dt, T.mask_noice.p, T.Hi.p, P.dHdt_n.p, T.dHi_dt.p, P.Hi_np1.p); }
  UFE_TRY(geometry_alter(c, cfg, prm, P.alter_time, T.Hi.p, T.Hi.p, T.Hb.p, T.SL.p, P.ref_Hi.p, P.ref_Hb.p, P.alter_have_T ? P.Ti_hom.p : nullptr, P.Hi_np1.p));   // (:233)
  { UFE_LAUNCH(c, TC_THICKNESS); k_pc_accept<<<grid, 128, 0, st>>>(n, M.a.off(), dt, rho_ratio, P.Hi_np1.p, P.dHdt_n.p, P.dHdt_nm1.p, T.Hb.p, T.SL.p, T.Hi.p, V.Hi.p, V.Hs.p,
                                                             lagged ? P.Hi_lag.p : nullptr, P.w_lag); }
  UFE_CUDA(cudaGetLastError());
  *n_Axb_its_thickness = its_p + its_c;
  return 0;
}

static int pc_state_ensure(Ctx* c, double dt, bool reset) {
  Mesh& M = *c->mesh;
  PcState& P = pc_of(c);
  const size_t na = M.a.n_ext();
  if (reset || !P.started || P.dHdt_n.n < na) {
    UFE_TRY(P.dHdt_n.alloc(na)); UFE_TRY(P.dHdt_nm1.alloc(na));
    UFE_CUDA(cudaMemsetAsync(P.dHdt_n.p, 0, sizeof(double) * na, c->stream)); UFE_CUDA(cudaMemsetAsync(P.dHdt_nm1.p, 0, sizeof(double) * na, c->stream));
    P.dt_n = dt; P.started = true;
  }
  return 0;
}

}  // namespace ufe

using namespace ufe;

extern "C" {

int ufe_b200_thickness_upload(ufe_b200_handle h, const ufe_b200_thickness_io* io) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && io, "NULL argument");
  UFE_CUDA(cudaSetDevice(c->device));
  UFE_TRY(ensure_thickness(c));
  Mesh& M = *c->mesh; Thickness& T = *c->thick;
  UFE_CHECK(io->Hi && io->Hb && io->SL && io->SMB && io->BMB && io->LMB && io->fraction_margin && io->mask_noice && io->dHi_dt_target,
            "ufe_b200_thickness_upload: Hi, Hb, SL, SMB, BMB, LMB, fraction_margin, mask_noice, dHi_dt_target are required");
  auto up = [&](double* d, const double* s) { return cudaMemcpyAsync(d + M.a.off(), s, sizeof(double) * M.a.n_loc(), cudaMemcpyHostToDevice, c->stream); };
  UFE_CUDA(up(T.Hi.p, io->Hi)); UFE_CUDA(up(T.Hb.p, io->Hb)); UFE_CUDA(up(T.SL.p, io->SL)); UFE_CUDA(up(T.SMB.p, io->SMB));
  UFE_CUDA(up(T.BMB.p, io->BMB)); UFE_CUDA(up(T.LMB.p, io->LMB)); UFE_CUDA(up(T.fraction_margin.p, io->fraction_margin));
  UFE_CUDA(up(T.target.p, io->dHi_dt_target));
  UFE_CUDA(cudaMemcpyAsync(T.mask_noice.p + M.a.off(), io->mask_noice, sizeof(int) * M.a.n_loc(), cudaMemcpyHostToDevice, c->stream));
  if (io->BC_prescr_mask && io->BC_prescr_Hi) {
    UFE_CUDA(cudaMemcpyAsync(T.bc_mask.p + M.a.off(), io->BC_prescr_mask, sizeof(int) * M.a.n_loc(), cudaMemcpyHostToDevice, c->stream));
    UFE_CUDA(up(T.bc_Hi.p, io->BC_prescr_Hi));
    T.have_bc = true;
  } else T.have_bc = false;
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  if (c->pc) { c->pc->started = false; c->pc->lag_valid = false; }
  return 0;
}

int ufe_b200_thickness_download(ufe_b200_handle h, double* Hi, double* dHi_dt, double* divQ) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && c->thick, "ufe_b200_thickness_download: nothing on the device");
  UFE_CUDA(cudaSetDevice(c->device));
  Mesh& M = *c->mesh; Thickness& T = *c->thick;
  auto dn = [&](double* d, const double* s) { return d ? cudaMemcpyAsync(d, s + M.a.off(), sizeof(double) * M.a.n_loc(), cudaMemcpyDeviceToHost, c->stream) : cudaSuccess; };
  UFE_CUDA(dn(Hi, T.Hi.p)); UFE_CUDA(dn(dHi_dt, T.dHi_dt.p)); UFE_CUDA(dn(divQ, T.divQ.p));
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

int ufe_b200_pc_step_resident(ufe_b200_handle h, const ufe_b200_config* cfg, int approximation, double dt, int* n_visc_its,
                              int* n_Axb_its_velocity, int* n_Axb_its_thickness) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && cfg && n_visc_its && n_Axb_its_velocity && n_Axb_its_thickness, "NULL argument");
  UFE_CHECK(c->vel && c->vel->uploaded && c->thick, "ufe_b200_pc_step_resident: upload the velocity and thickness inputs first");
  UFE_CUDA(cudaSetDevice(c->device));
  UFE_TRY(pc_state_ensure(c, dt, false));
  auto nop = []() { return 0; };
  UFE_TRY(pc_step_core(c, cfg, approximation, dt, n_visc_its, n_Axb_its_velocity, n_Axb_its_thickness, nop, nop));
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  c->pc->dt_n = dt;
  return 0;
}

// Host-buffer variant.  Transfers that the next phase does not depend on go through a second stream:
//   copy stream : [H2D ice inputs + rest of the velocity state]                [D2H velocity state (3-D fields ...)]
//   main stream : H2D thickness fields, u, v -> predictor -> (wait) velocity solve (record) -> corrector -> D2H thickness results
int ufe_b200_pc_step(ufe_b200_handle h, const ufe_b200_config* cfg, int approximation, double dt, double dt_prev,
                     const ufe_b200_ice_inputs* ice, ufe_b200_velocity_state* state, const ufe_b200_thickness_io* io, double* dHi_dt_nm1,
                     int* n_visc_its, int* n_Axb_its_velocity, int* n_Axb_its_thickness) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && cfg && ice && state && io && dHi_dt_nm1 && n_visc_its && n_Axb_its_velocity && n_Axb_its_thickness, "NULL argument");
  UFE_CHECK(io->Hi && io->Hb && io->SL && io->SMB && io->BMB && io->LMB && io->fraction_margin && io->mask_noice && io->dHi_dt_target,
            "ufe_b200_pc_step: Hi, Hb, SL, SMB, BMB, LMB, fraction_margin, mask_noice, dHi_dt_target are required");
  UFE_CHECK(state->u_vav_b && state->v_vav_b, "ufe_b200_pc_step: the velocity state needs u_vav_b and v_vav_b");
  UFE_CUDA(cudaSetDevice(c->device));
  UFE_TRY(ensure_thickness(c));
  UFE_TRY(ensure_velocity(c));
  Mesh& M = *c->mesh; Thickness& T = *c->thick;
  cudaStream_t main_st = c->stream;
  if (!c->copy_stream) {
    UFE_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    UFE_CUDA(cudaEventCreateWithFlags(&c->ev_copy_a, cudaEventDisableTiming));
    UFE_CUDA(cudaEventCreateWithFlags(&c->ev_copy_b, cudaEventDisableTiming));
  }
  UFE_TRY(pc_state_ensure(c, dt_prev, false));
  PcState& P = pc_of(c);
  P.dt_n = dt_prev > 0.0 ? dt_prev : dt;
  // ufe_b200_set_state_resident: after the first call the solver state (u, v, basal velocities and stresses, eta_3D_b) stays on the device,
  // like the reference's DIVA%... fields stay inside the solver object; `state` then only says which results come back
  const bool keep_state = c->state_resident && c->vel->uploaded;
  const size_t nb_a = sizeof(double) * M.a.n_loc();
  {  // what the predictor needs, on the main stream
    UFE_REGION(c, TC_H2D);
    auto up = [&](double* d, const double* s) { return cudaMemcpyAsync(d + M.a.off(), s, nb_a, cudaMemcpyHostToDevice, main_st); };
    UFE_CUDA(up(T.Hi.p, io->Hi)); UFE_CUDA(up(T.Hb.p, io->Hb)); UFE_CUDA(up(T.SL.p, io->SL)); UFE_CUDA(up(T.SMB.p, io->SMB));
    UFE_CUDA(up(T.BMB.p, io->BMB)); UFE_CUDA(up(T.LMB.p, io->LMB)); UFE_CUDA(up(T.fraction_margin.p, io->fraction_margin));
    UFE_CUDA(up(T.target.p, io->dHi_dt_target)); UFE_CUDA(up(P.dHdt_nm1.p, dHi_dt_nm1));
    UFE_CUDA(cudaMemcpyAsync(T.mask_noice.p + M.a.off(), io->mask_noice, sizeof(int) * M.a.n_loc(), cudaMemcpyHostToDevice, main_st));
    if (io->BC_prescr_mask && io->BC_prescr_Hi) {
      UFE_CUDA(cudaMemcpyAsync(T.bc_mask.p + M.a.off(), io->BC_prescr_mask, sizeof(int) * M.a.n_loc(), cudaMemcpyHostToDevice, main_st));
      UFE_CUDA(up(T.bc_Hi.p, io->BC_prescr_Hi));
      T.have_bc = true;
    } else T.have_bc = false;
    if (!keep_state) UFE_TRY(velocity_upload(c, nullptr, state, VU_STATE_UV));
  }
  {  // everything else the velocity solve needs, on the copy stream, while the predictor runs
    c->stream = c->copy_stream;
    int rc;
    { UFE_REGION(c, TC_H2D); rc = velocity_upload(c, ice, keep_state ? nullptr : state, VU_STATE_REST | VU_SKIP_GEOMETRY); }
    cudaError_t e = cudaEventRecord(c->ev_copy_a, c->copy_stream);
    c->stream = main_st;
    UFE_TRY(rc);
    UFE_CUDA(e);
  }
  auto before_velocity = [&]() -> int { UFE_CUDA(cudaStreamWaitEvent(main_st, c->ev_copy_a, 0)); return 0; };
  auto after_velocity = [&]() -> int {   // the velocity state is final: start bringing it back while the corrector runs
    UFE_CUDA(cudaEventRecord(c->ev_copy_b, main_st));
    UFE_CUDA(cudaStreamWaitEvent(c->copy_stream, c->ev_copy_b, 0));
    c->stream = c->copy_stream;
    int rc;
    { UFE_REGION(c, TC_D2H); rc = velocity_download(c, state, false); }
    c->stream = main_st;
    return rc;
  };
  UFE_TRY(pc_step_core(c, cfg, approximation, dt, n_visc_its, n_Axb_its_velocity, n_Axb_its_thickness, before_velocity, after_velocity));
  {
    UFE_REGION(c, TC_D2H);
    auto dn = [&](double* d, const double* s) { return d ? cudaMemcpyAsync(d, s + M.a.off(), nb_a, cudaMemcpyDeviceToHost, main_st) : cudaSuccess; };
    UFE_CUDA(dn(io->Hi_tplusdt, T.Hi.p)); UFE_CUDA(dn(io->dHi_dt, T.dHi_dt.p)); UFE_CUDA(dn(io->divQ, T.divQ.p)); UFE_CUDA(dn(io->AMB, T.AMB.p));
    UFE_CUDA(dn(dHi_dt_nm1, P.dHdt_nm1.p));
  }
  UFE_CUDA(cudaStreamSynchronize(main_st));
  UFE_CUDA(cudaStreamSynchronize(c->copy_stream));
  P.dt_n = dt;
  return 0;
}

int ufe_b200_pc_reference_loop(ufe_b200_handle h, int on, double wt_next_of_last_action) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c, "NULL handle");
  UFE_CHECK(wt_next_of_last_action >= 0.0 && wt_next_of_last_action <= 1.0, "ufe_b200_pc_reference_loop: the interpolation weight must lie in [0, 1] (got %g)",
            wt_next_of_last_action);
  PcState& P = pc_of(c);
  if (!on) P.lag_valid = false;
  P.reference_loop = on != 0;
  P.w_lag = wt_next_of_last_action;
  return 0;
}

int ufe_b200_pc_set_model_thickness(ufe_b200_handle h, const double* Hi_model) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && c->mesh && Hi_model, "ufe_b200_pc_set_model_thickness: NULL argument or no mesh");
  UFE_CUDA(cudaSetDevice(c->device));
  Mesh& M = *c->mesh; PcState& P = pc_of(c);
  const size_t na = M.a.n_ext();
  if (P.Hi_lag.n < na) { UFE_TRY(P.Hi_lag.alloc(na)); UFE_TRY(P.Hi_save.alloc(na)); UFE_CUDA(cudaMemsetAsync(P.Hi_lag.p, 0, sizeof(double) * na, c->stream)); }
  UFE_CUDA(cudaMemcpyAsync(P.Hi_lag.p + M.a.off(), Hi_model, sizeof(double) * M.a.n_loc(), cudaMemcpyHostToDevice, c->stream));
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  P.lag_valid = true;
  return 0;
}

int ufe_b200_alter_resident(ufe_b200_handle h, const ufe_b200_alter_params* prm, double time, const double* refgeo_Hi, const double* refgeo_Hb,
                            const double* Ti_hom) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && c->mesh, "ufe_b200_alter_resident: no mesh");
  PcState& P = pc_of(c);
  if (!prm) { P.alter_on = false; return 0; }
  UFE_CHECK(prm->modiness_H_style >= UFE_MODINESS_NONE && prm->modiness_H_style <= UFE_MODINESS_NO_THIN_INLAND, "unknown modiness_H_choice %d",
            prm->modiness_H_style);
  const bool need_T = prm->modiness_H_style >= UFE_MODINESS_TI_HOM && prm->modiness_H_style <= UFE_MODINESS_TI_HOM_DOWN;
  UFE_CHECK(!need_T || Ti_hom || P.alter_have_T, "ufe_b200_alter_resident: this modiness style needs Ti_hom");
  UFE_CHECK((refgeo_Hi && refgeo_Hb) || P.ref_Hi.n, "ufe_b200_alter_resident: the first call needs refgeo_Hi and refgeo_Hb");
  UFE_CUDA(cudaSetDevice(c->device));
  Mesh& M = *c->mesh;
  const size_t na = M.a.n_ext(), nb = sizeof(double) * M.a.n_loc();
  if (refgeo_Hi && refgeo_Hb) {
    UFE_TRY(P.ref_Hi.alloc(na)); UFE_TRY(P.ref_Hb.alloc(na));
    UFE_CUDA(cudaMemcpyAsync(P.ref_Hi.p + M.a.off(), refgeo_Hi, nb, cudaMemcpyHostToDevice, c->stream));
    UFE_CUDA(cudaMemcpyAsync(P.ref_Hb.p + M.a.off(), refgeo_Hb, nb, cudaMemcpyHostToDevice, c->stream));
  }
  if (Ti_hom) {
    UFE_TRY(P.Ti_hom.alloc(na));
    UFE_CUDA(cudaMemcpyAsync(P.Ti_hom.p + M.a.off(), Ti_hom, nb, cudaMemcpyHostToDevice, c->stream));
    P.alter_have_T = true;
  }
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  P.alter = *prm; P.alter_time = time; P.alter_on = true;
  return 0;
}

int ufe_b200_pc_download(ufe_b200_handle h, double* Hi_star_np1, double* dHi_dt_Hi_n_u_n, double* dHi_dt_applied, double* Hi_model) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c && c->mesh && c->pc && c->pc->started, "ufe_b200_pc_download: no time step has run");
  UFE_CUDA(cudaSetDevice(c->device));
  Mesh& M = *c->mesh; PcState& P = *c->pc;
  auto dn = [&](double* d, const DBuf<double>& s) { return (d && s.n) ? cudaMemcpyAsync(d, s.p + M.a.off(), sizeof(double) * M.a.n_loc(), cudaMemcpyDeviceToHost, c->stream) : cudaSuccess; };
  UFE_CUDA(dn(Hi_star_np1, P.Hi_star)); UFE_CUDA(dn(dHi_dt_Hi_n_u_n, P.dHdt_n)); UFE_CUDA(dn(dHi_dt_applied, P.dHdt_nm1));
  if (Hi_model) {
    if (P.reference_loop && P.lag_valid) UFE_CUDA(dn(Hi_model, P.Hi_lag));
    else UFE_CUDA(cudaMemcpyAsync(Hi_model, c->thick->Hi.p + M.a.off(), sizeof(double) * M.a.n_loc(), cudaMemcpyDeviceToHost, c->stream));
  }
  UFE_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

int ufe_b200_set_state_resident(ufe_b200_handle h, int on) {
  Ctx* c = reinterpret_cast<Ctx*>(h);
  UFE_CHECK(c, "NULL handle");
  c->state_resident = on ? 1 : 0;
  return 0;
}

}  // extern "C"
