// rt::geom::Aabb (src/geom/aabb.h:10-90).  Host-side only: boxes are built in double exactly
// as the reference builds them and are rounded outward to float by the flattener; the slab
// test itself (aabb.h:92-115) runs on the device (csrc/traverse.cuh).
#pragma once
#include <algorithm>
#include "../core/interval.h"
#include "../core/vec3.h"
namespace rt::geom {
class Aabb {
 public:
  core::Interval x, y, z;
  Aabb() {}  // empty
  Aabb(const core::Interval& ix, const core::Interval& iy, const core::Interval& iz) : x(ix), y(iy), z(iz) {}
  Aabb(const core::Point3& p, const core::Point3& q) {  // any two opposite corners
    x = p[0] <= q[0] ? core::Interval(p[0], q[0]) : core::Interval(q[0], p[0]);
    y = p[1] <= q[1] ? core::Interval(p[1], q[1]) : core::Interval(q[1], p[1]);
    z = p[2] <= q[2] ? core::Interval(p[2], q[2]) : core::Interval(q[2], p[2]);
  }
  Aabb(const Aabb& a, const Aabb& b) : x(a.x, b.x), y(a.y, b.y), z(a.z, b.z) {}  // union
  Aabb(const Aabb& b, const core::Vec3& p)                                         // grow by a point
      : x(std::min(b.x.min_, p.x()), std::max(b.x.max_, p.x())),
        y(std::min(b.y.min_, p.y()), std::max(b.y.max_, p.y())),
        z(std::min(b.z.min_, p.z()), std::max(b.z.max_, p.z())) {}
  const core::Interval& axis_interval(int n) const { return n == 1 ? y : (n == 2 ? z : x); }
  core::Vec3 min() const { return {x.min_, y.min_, z.min_}; }
  core::Vec3 max() const { return {x.max_, y.max_, z.max_}; }
  core::Vec3 center() const { return {0.5 * (x.min_ + x.max_), 0.5 * (y.min_ + y.max_), 0.5 * (z.min_ + z.max_)}; }
  int LongestAxis() const {
    const double ex = x.max_ - x.min_, ey = y.max_ - y.min_, ez = z.max_ - z.min_;
    if (ex >= ey && ex >= ez) return 0;
    return ey >= ez ? 1 : 2;
  }
  double SurfaceArea() const {
    const double ex = x.max_ - x.min_, ey = y.max_ - y.min_, ez = z.max_ - z.min_;
    return 2.0 * (ex * ey + ey * ez + ez * ex);
  }
};
}  // namespace rt::geom
