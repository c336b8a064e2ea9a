// rt::scene::Scene (src/scene/scene.h:15-74): an ordered list of hittables.  The reference's
// Scene::Hit walks the list linearly (scene.h:47-61); here the list is handed to the flattener
// in the same order, which keeps the reference's tie-break order.
#pragma once
#include <memory>
#include <vector>
#include "../geom/hittable.h"
namespace rt::scene {
class Scene : public geom::Hittable {
 public:
  std::vector<std::shared_ptr<geom::Hittable>> objects_;
  Scene() = default;
  explicit Scene(std::shared_ptr<geom::Hittable> object) { Add(object); }
  void Add(const std::shared_ptr<geom::Hittable>& object) {
    if (!object) return;
    objects_.push_back(object);
    bbox_ = geom::Aabb(bbox_, object->BoundingBox());
  }
  void Clear() {
    objects_.clear();
    bbox_ = geom::Aabb();
  }
  const std::vector<std::shared_ptr<geom::Hittable>>& Objects() const { return objects_; }
  geom::Aabb BoundingBox() const override { return bbox_; }
  int TypeId() const override { return -1; }
  int ObjectIndex() const override { return -1; }
  void set_object_index(int) override {}
  void Describe(geom::SceneSink& sink) const override { sink.OnList(objects_); }
 private:
  geom::Aabb bbox_;
};
}  // namespace rt::scene
