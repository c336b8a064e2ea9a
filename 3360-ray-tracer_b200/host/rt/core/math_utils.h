// Host-side helpers with the reference's names (src/core/math_utils.h:14-123).
// The per-bounce versions of these run on the device (csrc/shade.cuh); the host copies
// exist for scene construction and for user code written against the reference API.
#pragma once
#include <cmath>
#include "constants.h"
#include "random.h"
#include "vec3.h"
namespace rt::core {
inline Vec3 Reflect(const Vec3& v, const Vec3& n) { return v - 2.0 * Dot(v, n) * n; }
inline Vec3 Refract(const Vec3& uv, const Vec3& n, double eta_ratio) {
  const double cos_theta = std::fmin(Dot(-uv, n), 1.0);
  const Vec3 perp = eta_ratio * (uv + cos_theta * n);
  const Vec3 parallel = -std::sqrt(std::fabs(1.0 - perp.length_squared())) * n;
  return perp + parallel;
}
inline Vec3 RandomVec3() { return {RandomDouble(), RandomDouble(), RandomDouble()}; }
inline Vec3 RandomVec3(double lo, double hi) { return {RandomDouble(lo, hi), RandomDouble(lo, hi), RandomDouble(lo, hi)}; }
inline Vec3 Normalize(const Vec3& v) {
  const double len = v.length();
  return len == 0.0 ? Vec3(0, 0, 0) : v / len;
}
inline Vec3 RandomUnitVector() {
  for (;;) {
    const Vec3 p = RandomVec3(-1.0, 1.0);
    const double l2 = p.length_squared();
    if (l2 > 1e-12 && l2 <= 1.0) return p / std::sqrt(l2);
  }
}
inline Vec3 RandomInUnitSphere() {
  for (;;) {
    const Vec3 p = RandomVec3(-1.0, 1.0);
    if (p.length_squared() < 1.0) return p;
  }
}
inline Vec3 RandomOnHemisphere(const Vec3& normal) {
  const Vec3 u = RandomUnitVector();
  return Dot(u, normal) > 0.0 ? u : -u;
}
inline Vec3 RandomInUnitDisk() {
  for (;;) {
    const Vec3 p(RandomDouble(-1, 1), RandomDouble(-1, 1), 0.0);
    if (p.length_squared() < 1.0) return p;
  }
}
}  // namespace rt::core
