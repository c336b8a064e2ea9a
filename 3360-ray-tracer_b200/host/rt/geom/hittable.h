// rt::geom::Hittable / HitRecord / HittableType (src/geom/hittable.h:18-62).
//
// Differences from the reference, both forced by the B200 design:
//  * there is no per-object CPU `Hit`: closest-hit queries run on the device
//    (rtb_intersect_batch / the extend kernel); a CPU fallback is out of scope by contract;
//  * every object gains `Describe(SceneSink&)`, the export hook the reference lacks (its
//    primitive and material parameters are private with no getters, sphere.h:81-85,
//    triangle.h:105-108, rect.h:51-54).
#pragma once
#include <memory>
#include <vector>
#include "../core/interval.h"
#include "../core/ray.h"
#include "../core/vec3.h"
#include "aabb.h"

namespace rt::material { class Material; }

namespace rt::geom {

class Bvh;
class Hittable;

class HitRecord {
 public:
  bool hit = false;
  core::Point3 p;
  core::Vec3 normal;
  std::shared_ptr<material::Material> mat;
  double t = 0.0;
  bool front_face = false;
  double u = 0.0, v = 0.0;
  int prim_id = -1;  // extension: global primitive id (the reference record carries none)
  void set_face_normal(const core::Ray& r, const core::Vec3& outward_normal) {
    front_face = core::Dot(r.direction(), outward_normal) < 0;
    normal = front_face ? outward_normal : -outward_normal;
  }
};

enum HittableType { HITTABLE_SPHERE = 0, HITTABLE_TRIANGLE = 1, HITTABLE_SQUARE = 2 };

// Receiver of a scene walk.  `self` identifies the object (used for de-duplication / ids).
class SceneSink {
 public:
  virtual ~SceneSink() = default;
  virtual void OnSphere(const Hittable* self, const core::Point3& center, double radius,
                        const std::shared_ptr<material::Material>& mat) = 0;
  virtual void OnTriangle(const Hittable* self, const core::Point3& a, const core::Point3& b, const core::Point3& c,
                          const std::shared_ptr<material::Material>& mat) = 0;
  // axis = constant axis (2: xy_rect, 1: xz_rect, 0: yz_rect); (a,b) = remaining axes ascending
  virtual void OnRect(const Hittable* self, int axis, double k, double a0, double a1, double b0, double b1,
                      const std::shared_ptr<material::Material>& mat) = 0;
  virtual void OnBvh(const Bvh& bvh) = 0;
  virtual void OnList(const std::vector<std::shared_ptr<Hittable>>& objects) = 0;
};

class Hittable {
 public:
  virtual ~Hittable() = default;
  virtual Aabb BoundingBox() const = 0;
  virtual int TypeId() const = 0;
  virtual int ObjectIndex() const = 0;
  virtual void set_object_index(int i) = 0;
  virtual void Describe(SceneSink& sink) const = 0;
};

}  // namespace rt::geom
