// rt::geom::xy_rect / xz_rect / yz_rect (src/geom/rect.h:11-144): axis-aligned rectangles,
// boxes 2e-4 thick along the constant axis.
#pragma once
#include <memory>
#include "hittable.h"
namespace rt::geom {
namespace detail {
// axis = constant axis; (a,b) = the remaining axes in increasing order.
class AxisRect : public Hittable {
 public:
  AxisRect(int axis, double a0, double a1, double b0, double b1, double k, std::shared_ptr<material::Material> mat)
      : mat_(std::move(mat)), axis_(axis), a0_(a0), a1_(a1), b0_(b0), b1_(b1), k_(k) {}
  Aabb BoundingBox() const override {
    double lo[3], hi[3];
    const int ia = axis_ == 0 ? 1 : 0, ib = axis_ == 2 ? 1 : 2;
    lo[axis_] = k_ - 0.0001; hi[axis_] = k_ + 0.0001;
    lo[ia] = a0_; hi[ia] = a1_;
    lo[ib] = b0_; hi[ib] = b1_;
    return Aabb(core::Point3(lo[0], lo[1], lo[2]), core::Point3(hi[0], hi[1], hi[2]));
  }
  int TypeId() const override { return HITTABLE_SQUARE; }
  int ObjectIndex() const override { return index_; }
  void set_object_index(int i) override { index_ = i; }
  void Describe(SceneSink& sink) const override { sink.OnRect(this, axis_, k_, a0_, a1_, b0_, b1_, mat_); }
 private:
  std::shared_ptr<material::Material> mat_;
  int axis_;
  double a0_, a1_, b0_, b1_, k_;
  int index_ = -1;
};
}  // namespace detail
class xy_rect : public detail::AxisRect {
 public:
  xy_rect(double x0, double x1, double y0, double y1, double k, std::shared_ptr<material::Material> mat)
      : AxisRect(2, x0, x1, y0, y1, k, std::move(mat)) {}
};
class xz_rect : public detail::AxisRect {
 public:
  xz_rect(double x0, double x1, double z0, double z1, double k, std::shared_ptr<material::Material> mat)
      : AxisRect(1, x0, x1, z0, z1, k, std::move(mat)) {}
};
class yz_rect : public detail::AxisRect {
 public:
  yz_rect(double y0, double y1, double z0, double z1, double k, std::shared_ptr<material::Material> mat)
      : AxisRect(0, y0, y1, z0, z1, k, std::move(mat)) {}
};
}  // namespace rt::geom
