// rt::geom::Mesh (src/geom/mesh.h:13-38): indexed triangle list -> a Scene of Triangles.
// Concrete here (the reference's class misses three overrides and cannot be instantiated).
#pragma once
#include <array>
#include <memory>
#include <vector>
#include "../scene/scene.h"
#include "hittable.h"
#include "triangle.h"
namespace rt::geom {
class Mesh : public Hittable {
 public:
  scene::Scene tris;
  Mesh(const std::vector<core::Point3>& vertices, const std::vector<std::array<int, 3>>& indices,
       std::shared_ptr<material::Material> mat) {
    tris.objects_.reserve(indices.size());
    for (const auto& f : indices) tris.Add(std::make_shared<Triangle>(vertices[f[0]], vertices[f[1]], vertices[f[2]], mat));
  }
  Aabb BoundingBox() const override { return tris.BoundingBox(); }
  int TypeId() const override { return -1; }
  int ObjectIndex() const override { return -1; }
  void set_object_index(int) override {}
  void Describe(SceneSink& sink) const override { tris.Describe(sink); }
};
}  // namespace rt::geom
