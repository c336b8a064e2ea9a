// rt::geom::Sphere (src/geom/sphere.h:13-86).  Box = centre +/- the *unclamped* radius
// argument (sphere.h:17-21); the stored radius is max(0, radius).
#pragma once
#include <cmath>
#include <memory>
#include "../core/constants.h"
#include "hittable.h"
namespace rt::geom {
class Sphere : public Hittable {
 public:
  int gpu_index = -1;
  Sphere(const core::Point3& center, double radius, std::shared_ptr<material::Material> mat)
      : center_(center), radius_(std::fmax(0, radius)), mat_(std::move(mat)) {
    const core::Vec3 ext(radius, radius, radius);
    bbox_ = Aabb(core::Point3(center + ext), core::Point3(center - ext));
  }
  Aabb BoundingBox() const override { return bbox_; }
  int TypeId() const override { return HITTABLE_SPHERE; }
  int ObjectIndex() const override { return gpu_index; }
  void set_object_index(int i) override { gpu_index = i; }
  void Describe(SceneSink& sink) const override { sink.OnSphere(this, center_, radius_, mat_); }
  // (u,v) of a point on the unit sphere (sphere.h:73-79); the device evaluates the same in shade.
  static void get_sphere_uv(const core::Point3& p, double& u, double& v) {
    const double theta = std::acos(-p.y());
    const double phi = std::atan2(-p.z(), p.x()) + core::kPi;
    u = phi / (2 * core::kPi);
    v = theta / core::kPi;
  }
 private:
  core::Point3 center_;
  double radius_;
  std::shared_ptr<material::Material> mat_;
  Aabb bbox_;
};
}  // namespace rt::geom
