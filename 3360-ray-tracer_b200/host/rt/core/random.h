// Host RNG used ONLY at scene-build time (random sphere placement, src/main.cc:94-122).
// Same engine and distribution as src/core/random.h:14-25, but seeded explicitly so the
// oracle and the GPU see the same scene; per-sample randomness on the device is Philox.
#pragma once
#include <random>
namespace rt::core {
inline std::mt19937& GetRng() {
  thread_local static std::mt19937 engine(5489u);
  return engine;
}
inline void SeedRng(unsigned int seed) { GetRng().seed(seed); }
inline double RandomDouble() {
  thread_local static std::uniform_real_distribution<double> unit(0.0, 1.0);
  return unit(GetRng());
}
inline double RandomDouble(double lo, double hi) { return lo + (hi - lo) * RandomDouble(); }
inline int RandomInt(int lo, int hi) {
  std::uniform_int_distribution<int> d(lo, hi);
  return d(GetRng());
}
}  // namespace rt::core
