// rt::core::Ray (src/core/ray.h:7-26): origin + un-normalised direction.
#pragma once
#include "vec3.h"
namespace rt::core {
class Ray {
 public:
  Ray() {}
  Ray(const Point3& origin, const Vec3& direction) : o_(origin), d_(direction) {}
  const Point3& origin() const { return o_; }
  const Vec3& direction() const { return d_; }
  Point3 at(double t) const { return o_ + t * d_; }
 private:
  Point3 o_;
  Vec3 d_;
};
}  // namespace rt::core
