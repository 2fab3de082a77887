// Device state of the ice-thickness step (see thickness.cu).
#pragma once
#include "common.cuh"

namespace ufe {

struct Thickness {
  // a-grid fields over the extended range
  DBuf<double> Hi, Hb, SL, SMB, BMB, LMB, fraction_margin, target, bc_Hi, divQ, dHi_dt, AMB, Hi_new, Hi_ex, Hs, Hs_tot, bb;
  DBuf<int> mask_noice, bc_mask, n_int;
  DBuf<double> u_b, v_b;     // b-grid, extended (only when the caller passes velocities)
  DevCsr M;                  // M_divQ, then AA = 1 + dt f_s M_divQ in place
  DBuf<double> Mval0;        // M_divQ values (kept when AA overwrites M.val)
  DBuf<unsigned long long> dtbits;
  bool have_bc = false, pattern = false;
};
void thickness_free(Thickness* t);
int ensure_thickness(Ctx* c);
// calc_dHi_dt on device-resident inputs (T.Hi etc.); u_b/v_b: extended b-grid vectors with valid halos
int thickness_step(Ctx* c, const ufe_b200_config* cfg, const double* u_b, const double* v_b, double* dt, int finish, int* n_Axb_its);

}  // namespace ufe
