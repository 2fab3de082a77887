#include "thickness.cuh"
namespace ufe { struct Thickness { int dummy; }; void thickness_free(Thickness* t) { delete t; } }
using namespace ufe;
extern "C" {
int ufe_b200_calc_dHi_dt(ufe_b200_handle, const ufe_b200_config*, ufe_b200_thickness_io*, double*, int, int*) { set_error("calc_dHi_dt: not implemented yet"); return 9; }
int ufe_b200_dHi_dt_finish(ufe_b200_handle, const double*, double*, double, double*, double*) { set_error("dHi_dt_finish: not implemented yet"); return 9; }
}
