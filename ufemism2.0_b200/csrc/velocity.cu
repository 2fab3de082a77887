#include "velocity.cuh"
namespace ufe { struct Velocity { int dummy; }; void velocity_free(Velocity* v) { delete v; } }
using namespace ufe;
extern "C" {
#define NI(name) { set_error(name ": not implemented yet"); return 9; }
int ufe_b200_solve_SSA_DIVA_linearised(ufe_b200_handle, const ufe_b200_config*, double*, double*, const double*, const double*, const double*, const double*, const double*, const double*, double*, double*, double, double, int*, const int32_t*, const double*, const double*) NI("solve_SSA_DIVA_linearised")
int ufe_b200_get_stiffness_size(ufe_b200_handle, int*, int*) NI("get_stiffness_size")
int ufe_b200_get_stiffness(ufe_b200_handle, int32_t*, int32_t*, double*, double*) NI("get_stiffness")
int ufe_b200_solve_DIVA(ufe_b200_handle, const ufe_b200_config*, const ufe_b200_ice_inputs*, ufe_b200_velocity_state*, int*, int*, const int32_t*, const double*, const double*) NI("solve_DIVA")
int ufe_b200_solve_SSA(ufe_b200_handle, const ufe_b200_config*, const ufe_b200_ice_inputs*, ufe_b200_velocity_state*, int*, int*, const int32_t*, const double*, const double*) NI("solve_SSA")
int ufe_b200_velocity_upload(ufe_b200_handle, const ufe_b200_ice_inputs*, const ufe_b200_velocity_state*) NI("velocity_upload")
int ufe_b200_velocity_solve_resident(ufe_b200_handle, const ufe_b200_config*, int, int*, int*) NI("velocity_solve_resident")
int ufe_b200_velocity_download(ufe_b200_handle, ufe_b200_velocity_state*) NI("velocity_download")
int ufe_b200_solve_SIA(ufe_b200_handle, const ufe_b200_config*, const double*, const double*, const double*, double*, double*) NI("solve_SIA")
}
