// rt::core::Color + PPM pixel formatting (src/core/color.h:10-37).
#pragma once
#include <cmath>
#include <ostream>
#include "interval.h"
#include "vec3.h"
namespace rt::core {
using Color = Vec3;
inline double linear_to_gamma(double x) { return x > 0 ? std::sqrt(x) : 0.0; }
inline int to_byte(double linear) {  // clamp [0,0.999] then int(256 x), color.h:28-30
  static const Interval range(0.000, 0.999);
  return int(256 * range.Clamp(linear_to_gamma(linear)));
}
inline void write_color(std::ostream& out, const Color& c) {
  out << to_byte(c[0]) << ' ' << to_byte(c[1]) << ' ' << to_byte(c[2]) << '\n';
}
inline double luminance(const Color& c) { return 0.2126f * c.x() + 0.7152f * c.y() + 0.0722f * c.z(); }
}  // namespace rt::core
