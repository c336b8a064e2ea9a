// rt::integrator::RayIntegrator + the B200 implementation of the batch-intersection seam
// (src/integrator/ray_integrator.h:30-37; CPU counterpart src/integrator/cpu_ray_integrator.h:18-46).
#pragma once
#include <rtb.h>

#include <cmath>
#include <limits>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../b200/flatten.h"
#include "../core/ray.h"
#include "../geom/hittable.h"
#include "../geom/sphere.h"
#include "../scene/scene.h"

namespace rt::integrator {

class RayIntegrator {
 public:
  virtual ~RayIntegrator() = default;
  virtual void IntersectBatch(const std::vector<core::Ray>& rays, std::vector<geom::HitRecord>& hits) const = 0;
};

// Drop-in for CPURayIntegrator: same constructor argument, same IntersectBatch contract (hits is
// resized; rec.hit set; interval (0.001f, +inf) as cpu_ray_integrator.h:21-22).  Rays are
// narrowed to float for the device; t and the primitive are the reference's for that float ray.
// p / normal / front_face / u / v / mat are filled on the host in double with the reference's
// formulas (sphere.h:44-51, triangle.h:78-84, rect.h:26-38).
class B200RayIntegrator : public RayIntegrator {
 public:
  explicit B200RayIntegrator(const scene::Scene* world, int device = 0) {
    if (rtb_init(device) != RTB_OK) throw std::runtime_error(std::string("B200RayIntegrator: ") + rtb_last_error());
    b200::Flattener fl(1024, /*keep_prim_info=*/true);
    flat_ = fl.Run(*world);
    const RtbSceneDesc d = flat_.desc();
    if (rtb_scene_create(&d, &scene_) != RTB_OK) throw std::runtime_error(std::string("B200RayIntegrator: ") + rtb_last_error());
  }
  ~B200RayIntegrator() override { rtb_scene_destroy(scene_); }
  B200RayIntegrator(const B200RayIntegrator&) = delete;
  B200RayIntegrator& operator=(const B200RayIntegrator&) = delete;

  RtbScene* device_scene() const { return scene_; }

  void IntersectBatch(const std::vector<core::Ray>& rays, std::vector<geom::HitRecord>& hits) const override {
    hits.resize(rays.size());
    if (rays.empty()) return;
    std::vector<RtbRay> in(rays.size());
    for (size_t i = 0; i < rays.size(); ++i) {
      const core::Ray& r = rays[i];
      for (int c = 0; c < 3; ++c) {
        in[i].o[c] = static_cast<float>(r.origin()[c]);
        in[i].d[c] = static_cast<float>(r.direction()[c]);
      }
      in[i].tmin = 0.001f;
      in[i].tmax = std::numeric_limits<float>::infinity();
    }
    std::vector<RtbHit> out(rays.size());
    if (rtb_intersect_batch(scene_, in.data(), in.size(), out.data()) != RTB_OK)
      throw std::runtime_error(std::string("IntersectBatch: ") + rtb_last_error());
    for (size_t i = 0; i < rays.size(); ++i) {
      geom::HitRecord& rec = hits[i];
      rec = geom::HitRecord();
      if (out[i].prim_id < 0) continue;
      const core::Ray fr(core::Point3(in[i].o[0], in[i].o[1], in[i].o[2]), core::Vec3(in[i].d[0], in[i].d[1], in[i].d[2]));
      const b200::PrimInfo& pi = flat_.prim_info[static_cast<size_t>(out[i].prim_id)];
      rec.hit = true;
      rec.t = out[i].t;
      rec.p = fr.at(rec.t);
      rec.mat = pi.mat;
      rec.prim_id = out[i].prim_id;
      if (pi.type == RTB_PRIM_SPHERE) {
        const core::Vec3 n = (rec.p - core::Point3(pi.p[0], pi.p[1], pi.p[2])) / pi.p[3];
        rec.set_face_normal(fr, n);
        geom::Sphere::get_sphere_uv(n, rec.u, rec.v);
      } else if (pi.type == RTB_PRIM_TRIANGLE) {
        const core::Vec3 a(pi.p[0], pi.p[1], pi.p[2]), b(pi.p[3], pi.p[4], pi.p[5]), c(pi.p[6], pi.p[7], pi.p[8]);
        rec.set_face_normal(fr, core::Normalize(core::Cross(b - a, c - a)));
      } else {
        const int axis = static_cast<int>(pi.p[0]);
        const int ia = axis == 0 ? 1 : 0, ib = axis == 2 ? 1 : 2;
        const double pa = fr.origin()[ia] + rec.t * fr.direction()[ia], pb = fr.origin()[ib] + rec.t * fr.direction()[ib];
        rec.u = (pa - pi.p[2]) / (pi.p[3] - pi.p[2]);
        rec.v = (pb - pi.p[4]) / (pi.p[5] - pi.p[4]);
        rec.set_face_normal(fr, core::Vec3(axis == 0, axis == 1, axis == 2));
      }
    }
  }

 private:
  b200::FlatScene flat_;
  RtbScene* scene_ = nullptr;
};

}  // namespace rt::integrator
