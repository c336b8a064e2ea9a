// rt::integrator::RayState (src/integrator/ray_state.h:8-13).  On the device the same fields
// live in the structure-of-arrays PathQueue (csrc/rtb_kernels.cuh).
#pragma once
#include "../core/color.h"
#include "../core/ray.h"
namespace rt::integrator {
struct RayState {
  core::Ray r;
  int pixel_index;
  int depth;
  core::Color throughput;
};
}  // namespace rt::integrator
