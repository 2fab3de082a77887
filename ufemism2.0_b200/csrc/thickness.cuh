// Device state of the ice-thickness step (see thickness.cu).
#pragma once
#include "common.cuh"

namespace ufe {
struct Thickness;
void thickness_free(Thickness* t);
}  // namespace ufe
