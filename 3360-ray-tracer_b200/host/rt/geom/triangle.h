// rt::geom::Triangle (src/geom/triangle.h:14-109).  Box = vertex min/max padded by the
// float literal 1e-6f widened to double (triangle.h:33-36).
#pragma once
#include <cmath>
#include <memory>
#include "hittable.h"
namespace rt::geom {
class Triangle : public Hittable {
 public:
  int gpu_index = -1;
  Triangle(const core::Point3& a, const core::Point3& b, const core::Point3& c,
           std::shared_ptr<material::Material> mat)
      : a_(a), b_(b), c_(c), mat_(std::move(mat)) {
    core::Point3 lo(std::fmin(a.x(), std::fmin(b.x(), c.x())), std::fmin(a.y(), std::fmin(b.y(), c.y())),
                    std::fmin(a.z(), std::fmin(b.z(), c.z())));
    core::Point3 hi(std::fmax(a.x(), std::fmax(b.x(), c.x())), std::fmax(a.y(), std::fmax(b.y(), c.y())),
                    std::fmax(a.z(), std::fmax(b.z(), c.z())));
    const double pad = 1e-6f;
    lo += -core::Vec3(pad, pad, pad);
    hi += core::Vec3(pad, pad, pad);
    bbox_ = Aabb(lo, hi);
  }
  Aabb BoundingBox() const override { return bbox_; }
  int TypeId() const override { return HITTABLE_TRIANGLE; }
  int ObjectIndex() const override { return gpu_index; }
  void set_object_index(int i) override { gpu_index = i; }
  void Describe(SceneSink& sink) const override { sink.OnTriangle(this, a_, b_, c_, mat_); }
 private:
  core::Point3 a_, b_, c_;
  std::shared_ptr<material::Material> mat_;
  Aabb bbox_;
};
}  // namespace rt::geom
