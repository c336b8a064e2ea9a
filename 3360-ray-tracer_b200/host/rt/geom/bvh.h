// rt::geom::Bvh — host SAH builder (src/geom/bvh.h:28-368).
//
// Same algorithm and constants as the reference (16 centroid bins on the longest centroid
// axis, leaf at <= 4 primitives or when SAH refuses, cost 1 + (A_L/A) N_L + (A_R/A) N_R,
// std::partition on the bin index), evaluated in double in the same operation order, so
// nodes() and prim_indices() come out identical to the reference's: pre-order, left child at
// idx+1, leaves holding [first, count) into prim_indices().  That identity is what lets the
// device break exact-t ties the way Bvh::Hit's slot order does (bvh.h:95-106).
// Unlike the reference the tree is emitted directly in pre-order during the recursion instead
// of through a pointer tree plus a flatten pass; Hit() lives on the device.
#pragma once
#include <algorithm>
#include <cstdint>
#include <limits>
#include <memory>
#include <vector>
#include "../scene/scene.h"
#include "aabb.h"
#include "hittable.h"
#include "mesh.h"

namespace rt::geom {

// Same fields as the reference's BvhNodeGPU (bvh.h:20-25); the 32-byte float node the
// kernels read is RtbNode (include/rtb.h), produced from these by b200::Flattener.
struct BvhNodeGPU {
  Aabb bbox;
  uint32_t left_pIdx;   // internal: left child; leaf: first slot in prim_indices()
  uint32_t right_pCnt;  // internal: right child; leaf: primitive count
  uint32_t isLeaf;
};

class Bvh : public Hittable {
 public:
  Bvh(scene::Scene& scene) : Bvh(scene.objects_) {}
  Bvh(Mesh& mesh) : Bvh(mesh.tris) {}
  // While set (per thread), constructors only record their primitives and bounding box: for scenes whose traversal
  // tree is built on the device (b200::Flattener::RunSoup, RTB_SCENE_DEVICE_BUILD) the SAH build is wasted work.
  static bool& DeferBuild() {
    thread_local bool defer = false;
    return defer;
  }

  Bvh(std::vector<std::shared_ptr<Hittable>>& objects) {
    if (objects.empty()) return;
    primitives_ = objects;
    const int n = static_cast<int>(primitives_.size());
    if (DeferBuild()) {
      deferred_box_ = primitives_[0]->BoundingBox();
      for (int i = 1; i < n; ++i) deferred_box_ = Aabb(deferred_box_, primitives_[i]->BoundingBox());
      return;
    }
    prim_indices_.resize(n);
    bounds_.resize(n);
    centroids_.resize(n);
    for (int i = 0; i < n; ++i) {
      prim_indices_[i] = i;
      bounds_[i] = primitives_[i]->BoundingBox();
      centroids_[i] = bounds_[i].center();
    }
    nodes_.reserve(2 * static_cast<size_t>(n));
    root_index_ = Build(0, n);
    bounds_.clear();
    bounds_.shrink_to_fit();
    centroids_.clear();
    centroids_.shrink_to_fit();
  }

  Aabb BoundingBox() const override { return root_index_ < 0 ? deferred_box_ : nodes_[root_index_].bbox; }
  int TypeId() const override { return -1; }
  int ObjectIndex() const override { return -1; }
  void set_object_index(int) override {}
  void Describe(SceneSink& sink) const override { sink.OnBvh(*this); }

  const std::vector<BvhNodeGPU>& nodes() const { return nodes_; }
  const std::vector<int>& prim_indices() const { return prim_indices_; }
  const std::vector<std::shared_ptr<Hittable>>& primitives() const { return primitives_; }
  int root_index() const { return root_index_; }

 private:
  static constexpr int kMaxLeaf = 4;
  static constexpr int kBins = 16;

  int BinOf(int prim, int axis, double lo, double inv_extent) const {
    int b = static_cast<int>((centroids_[prim][axis] - lo) * inv_extent * kBins);
    return b < 0 ? 0 : (b >= kBins ? kBins - 1 : b);
  }

  int MakeLeaf(int idx, int first, int count) {
    nodes_[idx].isLeaf = 1;
    nodes_[idx].left_pIdx = static_cast<uint32_t>(first);
    nodes_[idx].right_pCnt = static_cast<uint32_t>(count);
    return idx;
  }

  // Emits the subtree over prim_indices_[start,end) in pre-order; returns its node index.
  int Build(int start, int end) {
    const int idx = static_cast<int>(nodes_.size());
    nodes_.push_back({});
    const int count = end - start;

    Aabb box = bounds_[prim_indices_[start]];
    for (int i = start + 1; i < end; ++i) box = Aabb(box, bounds_[prim_indices_[i]]);
    nodes_[idx].bbox = box;
    if (count <= kMaxLeaf) return MakeLeaf(idx, start, count);

    const core::Vec3& c0 = centroids_[prim_indices_[start]];
    Aabb cbox(c0, c0);
    for (int i = start + 1; i < end; ++i) cbox = Aabb(cbox, centroids_[prim_indices_[i]]);
    const int axis = cbox.LongestAxis();
    const double lo = cbox.axis_interval(axis).min_;
    const double extent = cbox.axis_interval(axis).max_ - lo;
    if (extent <= 0.0) return MakeLeaf(idx, start, count);
    const double inv_extent = 1.0 / extent;

    int bin_count[kBins] = {};
    Aabb bin_box[kBins];  // empty boxes: union with empty is the identity
    for (int i = start; i < end; ++i) {
      const int p = prim_indices_[i];
      const int b = BinOf(p, axis, lo, inv_extent);
      bin_box[b] = bin_count[b] ? Aabb(bin_box[b], bounds_[p]) : bounds_[p];
      ++bin_count[b];
    }
    Aabb lbox[kBins], rbox[kBins];
    int lcnt[kBins], rcnt[kBins];
    {
      Aabb acc;
      int n = 0;
      for (int i = 0; i < kBins; ++i) {
        if (bin_count[i]) { acc = n ? Aabb(acc, bin_box[i]) : bin_box[i]; n += bin_count[i]; }
        lbox[i] = acc; lcnt[i] = n;
      }
      n = 0;
      for (int i = kBins - 1; i >= 0; --i) {
        if (bin_count[i]) { acc = n ? Aabb(acc, bin_box[i]) : bin_box[i]; n += bin_count[i]; }
        rbox[i] = acc; rcnt[i] = n;
      }
    }
    double best_cost = std::numeric_limits<double>::infinity();
    int best = -1;
    const double area = box.SurfaceArea();
    for (int i = 0; i + 1 < kBins; ++i) {
      if (lcnt[i] == 0 || rcnt[i + 1] == 0) continue;
      const double cost = 1.0 + (lbox[i].SurfaceArea() / area) * lcnt[i] + (rbox[i + 1].SurfaceArea() / area) * rcnt[i + 1];
      if (cost < best_cost) { best_cost = cost; best = i; }
    }
    // leaf cost is int * float in the reference (bvh.h:301), i.e. rounded to float first
    const double leaf_cost = static_cast<float>(count) * 1.0f;
    if (best < 0 || best_cost >= leaf_cost) return MakeLeaf(idx, start, count);

    auto mid_it = std::partition(prim_indices_.begin() + start, prim_indices_.begin() + end,
                                 [&](int p) { return BinOf(p, axis, lo, inv_extent) <= best; });
    const int mid = static_cast<int>(mid_it - prim_indices_.begin());
    if (mid == start || mid == end) return MakeLeaf(idx, start, count);

    const int left = Build(start, mid);
    const int right = Build(mid, end);
    nodes_[idx].isLeaf = 0;
    nodes_[idx].left_pIdx = static_cast<uint32_t>(left);
    nodes_[idx].right_pCnt = static_cast<uint32_t>(right);
    return idx;
  }

  std::vector<std::shared_ptr<Hittable>> primitives_;
  std::vector<int> prim_indices_;
  std::vector<Aabb> bounds_;
  std::vector<core::Vec3> centroids_;
  std::vector<BvhNodeGPU> nodes_;
  Aabb deferred_box_;
  int root_index_ = -1;
};

}  // namespace rt::geom
