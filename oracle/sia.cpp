// TEST INFRASTRUCTURE ONLY (see oracle.hpp).
// solve_SIA (src/UFEMISM/ice_dynamics/conservation_of_momentum/SIA/SIA_main.f90:65-158)
#include <cmath>

#include "oracle.hpp"

namespace orc {
void solve_SIA(const MeshView&, const Operators&, const Config&, const double*, const double*, const double*, double*, double*) {}
}  // namespace orc
