// Device-resident entry points of the geometry / mask / sub-grid stage (geometry.cu) for the predictor-corrector step (timestep.cu).
#pragma once
#include "common.cuh"

namespace ufe {

// true after ufe_b200_geometry_resident( choice >= 0): the resident time step updates masks and sub-grid fractions itself
bool geometry_dynamic(const Ctx* c);
// determine_masks + calc_grounded_fractions for the geometry (Hi, Hb, SL; extended a-range device vectors, owned part valid):
// writes mask_grounded_ice / mask_icefree_land (extended a range, owned part) and fraction_gr_b (extended b range, owned part) of the velocity
// solve, and whether any owned vertex is grounded ice
int geometry_for_velocity(Ctx* c, const ufe_b200_config* cfg, double* Hi, const double* Hb, double* SL, int* mask_grounded_ice, int* mask_icefree_land,
                          double* fraction_gr_b, int* any_grounded);
// calc_effective_thickness for the geometry: writes fraction_margin (extended a range, owned part)
int geometry_for_thickness(Ctx* c, const ufe_b200_config* cfg, double* Hi, const double* Hb, const double* SL, double* fraction_margin);

// alter_ice_thickness on device fields (ice_thickness_safeties.f90:26-288); prm == nullptr: only the removal of ice thinner than C%Hi_min
int geometry_alter(Ctx* c, const ufe_b200_config* cfg, const ufe_b200_alter_params* prm, double time, double* Hi_of_masks, const double* Hi_old,
                   const double* Hb, const double* SL, const double* ref_Hi, const double* ref_Hb, const double* Ti_hom, double* Hi_new);

}  // namespace ufe
