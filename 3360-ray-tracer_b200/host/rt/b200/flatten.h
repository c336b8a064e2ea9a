// rt::b200::Flattener — host scene graph -> RtbSceneDesc (include/rtb.h).
//
// Replaces the role of geom::Bvh::flatten (src/geom/bvh.h:347-367) and of the GPU-facing
// accessors (bvh.h:134-136).  The reference tree (pre-order, 64-byte double nodes, nested Bvh
// objects as leaf primitives, a linear scene::Scene list on top) is re-laid-out for the device:
//   * one node array for all nesting levels; a nested Bvh becomes an RTB_PRIM_SUBTREE leaf entry;
//   * 32-byte float nodes, boxes rounded outward so float culling never drops a reference hit;
//   * siblings adjacent (one 64-byte fetch per internal node); the first n_top_nodes are in
//     breadth-first order (the prefix the extend kernel stages in shared memory), the rest in
//     depth-first order;
//   * primitive tables permuted into leaf order; gid[] keeps the creation-order id and rank[]
//     the reference's traversal slot order (bvh.h:95-106) for exact-t tie breaks;
//   * leaves larger than RTB_MAX_LEAF (SAH-refused leaves can be any size) are split.
// It also serialises the scene as flat double records for the test oracle (Export*()).
#pragma once
#include <rtb.h>

#include <cmath>
#include <cstdint>
#include <deque>
#include <map>
#include <unordered_map>
#include <stdexcept>
#include <string>
#include <vector>

#include "../geom/bvh.h"
#include "../geom/hittable.h"
#include "../material/material.h"
#include "../scene/scene.h"

namespace rt::b200 {

// Record kinds of the oracle export stream (16 doubles per record: kind, material, 12 params).
enum ExportKind { kExSphere = 0, kExTriangle = 1, kExRect = 2, kExBvhBegin = 10, kExBvhEnd = 11, kExListBegin = 12, kExListEnd = 13 };
constexpr int kExportRecord = 16;

// Per-primitive parameters in creation (gid) order; kept only on request (B200RayIntegrator uses
// them to fill HitRecord on the host).  sphere: c,r; triangle: a,b,c; rect: axis,k,a0,a1,b0,b1.
struct PrimInfo {
  uint32_t type;
  double p[9];
  std::shared_ptr<material::Material> mat;
};

struct FlatScene {
  std::vector<PrimInfo> prim_info;
  std::vector<RtbNode> nodes;
  std::vector<uint32_t> prim_refs;
  std::vector<RtbSphere> spheres;
  std::vector<RtbTriangle> triangles;
  std::vector<RtbRect> rects;
  std::vector<uint32_t> sphere_mat, sphere_gid, sphere_rank, triangle_gid, triangle_rank, rect_gid, rect_rank;
  std::vector<RtbMaterial> materials;
  std::vector<RtbTexture> textures;
  std::vector<RtbImage> images;
  std::vector<std::vector<uint8_t>> image_bytes;
  std::vector<std::string> image_paths;
  uint32_t n_top_nodes = 0, max_depth = 0, n_prims = 0;
  // oracle export
  std::vector<double> prim_records, material_records, texture_records;
  // gid -> RTB_REF(type, table index)
  std::vector<uint32_t> gid_to_ref;

  bool device_build = false;  // RunSoup(): primitive tables only, RTB_SCENE_DEVICE_BUILD

  RtbSceneDesc desc() const {
    RtbSceneDesc d{};
    d.abi_version = RTB_ABI_VERSION;
    d.n_nodes = static_cast<uint32_t>(nodes.size());
    d.nodes = nodes.data();
    d.n_prim_refs = static_cast<uint32_t>(prim_refs.size());
    d.prim_refs = prim_refs.data();
    d.n_spheres = static_cast<uint32_t>(spheres.size());
    d.spheres = spheres.data();
    d.sphere_mat = sphere_mat.data();
    d.n_triangles = static_cast<uint32_t>(triangles.size());
    d.triangles = triangles.data();
    d.n_rects = static_cast<uint32_t>(rects.size());
    d.rects = rects.data();
    d.sphere_gid = sphere_gid.data();
    d.triangle_gid = triangle_gid.data();
    d.rect_gid = rect_gid.data();
    d.sphere_rank = sphere_rank.data();
    d.triangle_rank = triangle_rank.data();
    d.rect_rank = rect_rank.data();
    d.n_materials = static_cast<uint32_t>(materials.size());
    d.materials = materials.data();
    d.n_textures = static_cast<uint32_t>(textures.size());
    d.textures = textures.data();
    d.n_images = static_cast<uint32_t>(images.size());
    d.images = images.data();
    d.n_top_nodes = n_top_nodes;
    d.max_depth = max_depth;
    d.compiled = nullptr;  // rtb_tree_compile's result, set by whoever owns it (rth_capi.cc, B200Renderer)
    d.flags = device_build ? RTB_SCENE_DEVICE_BUILD : 0u;
    if (device_build) {  // no host tree: the device builds it, creation order breaks exact-t ties
      d.nodes = nullptr;
      d.prim_refs = nullptr;
      d.sphere_rank = d.triangle_rank = d.rect_rank = nullptr;
    }
    return d;
  }
};

class Flattener : public geom::SceneSink, public material::MaterialSink, public material::TextureSink {
 public:
  // top_budget: how many nodes go into the breadth-first prefix (shared-memory stage).
  explicit Flattener(uint32_t top_budget = 1024, bool keep_prim_info = false) : top_budget_(top_budget), keep_info_(keep_prim_info) {}

  FlatScene Run(const scene::Scene& world) {
    out_ = FlatScene();
    fnodes_.clear();
    Item root = Convert(world);
    if (root.type != RTB_PRIM_SUBTREE) throw std::runtime_error("flatten: empty scene");
    // a world holding exactly one BVH / list: use that tree as the root directly
    while (fnodes_[root.index].left < 0 && fnodes_[root.index].items.size() == 1 &&
           fnodes_[root.index].items[0].type == RTB_PRIM_SUBTREE)
      root = fnodes_[root.index].items[0];
    AssignRanks(root.index);
    Layout(root.index);
    return std::move(out_);
  }

  // Primitive tables, materials and textures only, in creation order: the traversal tree is built on the device
  // (RTB_SCENE_DEVICE_BUILD).  Host geom::Bvh objects met on the way contribute their primitives, never their nodes,
  // so a scene assembled under geom::Bvh::DeferBuild never runs the SAH build at all.
  FlatScene RunSoup(const scene::Scene& world) {
    out_ = FlatScene();
    fnodes_.clear();
    soup_ = true;
    Convert(world);
    soup_ = false;
    if (out_.n_prims == 0) throw std::runtime_error("flatten: empty scene");
    for (size_t i = 0; i < out_.images.size(); ++i) out_.images[i].rgb = out_.image_bytes[i].empty() ? nullptr : out_.image_bytes[i].data();
    out_.device_build = true;
    return std::move(out_);
  }

  // ---- geom::SceneSink ----
  void OnSphere(const geom::Hittable*, const core::Point3& c, double r,
                const std::shared_ptr<material::Material>& mat) override {
    const uint32_t m = MaterialId(mat);
    const uint32_t gid = NextGid(RTB_PRIM_SPHERE, static_cast<uint32_t>(out_.spheres.size()));
    out_.spheres.push_back(RtbSphere{{c[0], c[1], c[2]}, r});
    out_.sphere_mat.push_back(m);
    out_.sphere_gid.push_back(gid);
    Export(kExSphere, m, {c[0], c[1], c[2], r});
    if (keep_info_) out_.prim_info.push_back(PrimInfo{RTB_PRIM_SPHERE, {c[0], c[1], c[2], r}, mat});
    last_ = Item{RTB_PRIM_SPHERE, static_cast<uint32_t>(out_.spheres.size() - 1)};
  }
  void OnTriangle(const geom::Hittable*, const core::Point3& a, const core::Point3& b, const core::Point3& c,
                  const std::shared_ptr<material::Material>& mat) override {
    const uint32_t m = MaterialId(mat);
    const uint32_t gid = NextGid(RTB_PRIM_TRIANGLE, static_cast<uint32_t>(out_.triangles.size()));
    const core::Vec3 e1 = b - a, e2 = c - a;  // the doubles Triangle::Hit forms per call (triangle.h:45-46)
    out_.triangles.push_back(RtbTriangle{{a[0], a[1], a[2]}, {e1[0], e1[1], e1[2]}, {e2[0], e2[1], e2[2]}, m, 0});
    out_.triangle_gid.push_back(gid);
    Export(kExTriangle, m, {a[0], a[1], a[2], b[0], b[1], b[2], c[0], c[1], c[2]});
    if (keep_info_) out_.prim_info.push_back(PrimInfo{RTB_PRIM_TRIANGLE, {a[0], a[1], a[2], b[0], b[1], b[2], c[0], c[1], c[2]}, mat});
    last_ = Item{RTB_PRIM_TRIANGLE, static_cast<uint32_t>(out_.triangles.size() - 1)};
  }
  void OnRect(const geom::Hittable*, int axis, double k, double a0, double a1, double b0, double b1,
              const std::shared_ptr<material::Material>& mat) override {
    const uint32_t m = MaterialId(mat);
    const uint32_t gid = NextGid(RTB_PRIM_RECT, static_cast<uint32_t>(out_.rects.size()));
    out_.rects.push_back(RtbRect{k, a0, a1, b0, b1, axis, m});
    out_.rect_gid.push_back(gid);
    Export(kExRect, m, {static_cast<double>(axis), k, a0, a1, b0, b1});
    if (keep_info_) out_.prim_info.push_back(PrimInfo{RTB_PRIM_RECT, {static_cast<double>(axis), k, a0, a1, b0, b1}, mat});
    last_ = Item{RTB_PRIM_RECT, static_cast<uint32_t>(out_.rects.size() - 1)};
  }
  void OnBvh(const geom::Bvh& bvh) override {
    Export(kExBvhBegin, 0, {});
    std::vector<Item> by_prim;
    by_prim.reserve(bvh.primitives().size());
    for (const auto& p : bvh.primitives()) by_prim.push_back(Convert(*p));
    Export(kExBvhEnd, 0, {});
    if (soup_ || bvh.nodes().empty()) {
      last_ = Item{kNone, 0};
      return;
    }
    last_ = Item{RTB_PRIM_SUBTREE, static_cast<uint32_t>(FromBvh(bvh, bvh.root_index(), by_prim))};
  }
  void OnList(const std::vector<std::shared_ptr<geom::Hittable>>& objects) override {
    Export(kExListBegin, 0, {});
    std::vector<Item> items;
    geom::Aabb box;
    for (const auto& o : objects) {
      Item it = Convert(*o);
      if (it.type == kNone) continue;
      items.push_back(it);
      box = geom::Aabb(box, o->BoundingBox());
    }
    Export(kExListEnd, 0, {});
    if (soup_ || items.empty()) {
      last_ = Item{kNone, 0};
      return;
    }
    last_ = Item{RTB_PRIM_SUBTREE, static_cast<uint32_t>(MakeLeaf(box, items.data(), items.size()))};
  }

  // ---- material::MaterialSink ----
  void OnLambertian(const std::shared_ptr<material::Texture>& tex) override {
    mat_tmp_ = RtbMaterial{RTB_MAT_LAMBERTIAN, static_cast<int32_t>(TextureId(tex)), {0, 0, 0}, 0, 0, 0};
  }
  void OnMetal(const core::Color& albedo, double fuzz) override {
    mat_tmp_ = RtbMaterial{RTB_MAT_METAL, -1, {float(albedo[0]), float(albedo[1]), float(albedo[2])}, float(fuzz), 0, 0};
    mat_d_[0] = albedo[0]; mat_d_[1] = albedo[1]; mat_d_[2] = albedo[2]; mat_d_[3] = fuzz;
  }
  void OnDielectric(double ref_idx) override {
    mat_tmp_ = RtbMaterial{RTB_MAT_DIELECTRIC, -1, {0, 0, 0}, 0, float(ref_idx), 0};
    mat_d_[4] = ref_idx;
  }
  void OnDiffuseLight(const std::shared_ptr<material::Texture>& tex) override {
    mat_tmp_ = RtbMaterial{RTB_MAT_DIFFUSE_LIGHT, static_cast<int32_t>(TextureId(tex)), {0, 0, 0}, 0, 0, 0};
  }

  // ---- material::TextureSink ----
  void OnSolid(const core::Color& c) override {
    tex_tmp_ = RtbTexture{RTB_TEX_SOLID, -1, -1, -1, {float(c[0]), float(c[1]), float(c[2])}, 0};
    tex_d_[0] = c[0]; tex_d_[1] = c[1]; tex_d_[2] = c[2];
  }
  void OnChecker(double scale, double inv_scale, const std::shared_ptr<material::Texture>& even,
                 const std::shared_ptr<material::Texture>& odd) override {
    const int32_t e = static_cast<int32_t>(TextureId(even)), o = static_cast<int32_t>(TextureId(odd));
    tex_tmp_ = RtbTexture{RTB_TEX_CHECKER, -1, e, o, {0, 0, 0}, float(inv_scale)};
    tex_d_[0] = tex_d_[1] = tex_d_[2] = 0;
    tex_d_[3] = scale;
  }
  void OnImage(const scene::Image& img) override {
    const int32_t id = static_cast<int32_t>(out_.image_bytes.size());
    out_.image_bytes.emplace_back(img.Bytes().begin(), img.Bytes().end());
    out_.image_paths.push_back(img.Path());
    out_.images.push_back(RtbImage{img.Width(), img.Height(), nullptr});  // pointers fixed in Layout()
    tex_tmp_ = RtbTexture{RTB_TEX_IMAGE, id, -1, -1, {0, 0, 0}, 0};
    tex_d_[0] = tex_d_[1] = tex_d_[2] = tex_d_[3] = 0;
  }

 private:
  static constexpr uint32_t kNone = 0xffffffffu;
  struct Item { uint32_t type; uint32_t index; };
  struct FNode {
    geom::Aabb box;
    int left = -1, right = -1;   // FNode indices; -1/-1 = leaf
    std::vector<Item> items;     // leaf entries in the reference's slot order
    uint32_t dev = 0;            // device node index
  };

  Item Convert(const geom::Hittable& h) {
    h.Describe(*this);
    return last_;
  }

  uint32_t NextGid(uint32_t type, uint32_t index) {
    out_.gid_to_ref.push_back(RTB_REF(type, index));
    return out_.n_prims++;
  }

  void Export(int kind, uint32_t mat, std::initializer_list<double> params) {
    if (soup_) return;  // the oracle's export stream is not produced for device-built scenes (640 MB at 5 M primitives)
    const size_t base = out_.prim_records.size();
    out_.prim_records.resize(base + kExportRecord, 0.0);
    out_.prim_records[base] = kind;
    out_.prim_records[base + 1] = mat;
    size_t i = 2;
    for (double p : params) out_.prim_records[base + i++] = p;
  }

  uint32_t MaterialId(const std::shared_ptr<material::Material>& m) {
    if (!m) throw std::runtime_error("flatten: primitive without material");
    if (m.get() == last_mat_) return last_mat_id_;  // a mesh's triangles share one material
    auto it = mat_ids_.find(m.get());
    if (it != mat_ids_.end()) {
      last_mat_ = m.get();
      last_mat_id_ = it->second;
      return it->second;
    }
    for (double& d : mat_d_) d = 0;
    m->Describe(*this);
    const uint32_t id = static_cast<uint32_t>(out_.materials.size());
    out_.materials.push_back(mat_tmp_);
    const double rec[8] = {double(mat_tmp_.type), double(mat_tmp_.tex), mat_d_[0], mat_d_[1], mat_d_[2], mat_d_[3], mat_d_[4], 0};
    out_.material_records.insert(out_.material_records.end(), rec, rec + 8);
    mat_ids_[m.get()] = id;
    return id;
  }

  uint32_t TextureId(const std::shared_ptr<material::Texture>& t) {
    if (!t) throw std::runtime_error("flatten: material without texture");
    auto it = tex_ids_.find(t.get());
    if (it != tex_ids_.end()) return it->second;
    t->Describe(*this);  // children (checker) are registered first, so they get smaller ids
    const RtbTexture tex = tex_tmp_;
    const double rec[8] = {double(tex.type), double(tex.image), double(tex.even), double(tex.odd), tex_d_[0], tex_d_[1], tex_d_[2], tex_d_[3]};
    const uint32_t id = static_cast<uint32_t>(out_.textures.size());
    out_.textures.push_back(tex);
    out_.texture_records.insert(out_.texture_records.end(), rec, rec + 8);
    tex_ids_[t.get()] = id;
    return id;
  }

  int MakeLeaf(const geom::Aabb& box, const Item* items, size_t n) {
    const int idx = static_cast<int>(fnodes_.size());
    fnodes_.push_back(FNode{});
    fnodes_[idx].box = box;
    if (n <= RTB_MAX_LEAF) {
      fnodes_[idx].items.assign(items, items + n);
    } else {  // oversize leaf: halve in slot order under the same box
      const size_t half = n / 2;
      const int l = MakeLeaf(box, items, half);
      const int r = MakeLeaf(box, items + half, n - half);
      fnodes_[idx].left = l;
      fnodes_[idx].right = r;
    }
    return idx;
  }

  int FromBvh(const geom::Bvh& bvh, int node, const std::vector<Item>& by_prim) {
    const geom::BvhNodeGPU& n = bvh.nodes()[node];
    if (n.isLeaf) {
      std::vector<Item> items;
      for (uint32_t i = 0; i < n.right_pCnt; ++i) {
        const Item it = by_prim[bvh.prim_indices()[n.left_pIdx + i]];
        if (it.type != kNone) items.push_back(it);
      }
      return MakeLeaf(n.bbox, items.data(), items.size());
    }
    const int idx = static_cast<int>(fnodes_.size());
    fnodes_.push_back(FNode{});
    fnodes_[idx].box = n.bbox;
    const int l = FromBvh(bvh, static_cast<int>(n.left_pIdx), by_prim);
    const int r = FromBvh(bvh, static_cast<int>(n.right_pCnt), by_prim);
    fnodes_[idx].left = l;
    fnodes_[idx].right = r;
    return idx;
  }

  // Reference visiting order: pre-order, left before right, leaf slots ascending (bvh.h:84-116).
  void AssignRanks(int root) {
    out_.sphere_rank.assign(out_.spheres.size(), 0);
    out_.triangle_rank.assign(out_.triangles.size(), 0);
    out_.rect_rank.assign(out_.rects.size(), 0);
    uint32_t next = 0;
    std::vector<std::pair<int, size_t>> stack;  // (fnode, next item) / internal nodes expand once
    std::vector<int> todo{root};
    while (!todo.empty()) {
      const int n = todo.back();
      todo.pop_back();
      if (n < 0) {  // marker: resume a leaf after a nested subtree
        auto [leaf, pos] = stack.back();
        stack.pop_back();
        WalkLeaf(leaf, pos, next, todo, stack);
        continue;
      }
      const FNode& f = fnodes_[n];
      if (f.left >= 0) {
        todo.push_back(f.right);
        todo.push_back(f.left);
      } else {
        WalkLeaf(n, 0, next, todo, stack);
      }
    }
  }
  void WalkLeaf(int leaf, size_t pos, uint32_t& next, std::vector<int>& todo, std::vector<std::pair<int, size_t>>& stack) {
    const FNode& f = fnodes_[leaf];
    for (size_t i = pos; i < f.items.size(); ++i) {
      const Item it = f.items[i];
      if (it.type == RTB_PRIM_SUBTREE) {
        stack.emplace_back(leaf, i + 1);
        todo.push_back(-1);
        todo.push_back(static_cast<int>(it.index));
        return;
      }
      (it.type == RTB_PRIM_SPHERE ? out_.sphere_rank : it.type == RTB_PRIM_TRIANGLE ? out_.triangle_rank : out_.rect_rank)[it.index] = next++;
    }
  }

  static float RoundDown(double v) {
    float f = static_cast<float>(v);
    if (static_cast<double>(f) > v) f = std::nextafterf(f, -INFINITY);
    return f;
  }
  static float RoundUp(double v) {
    float f = static_cast<float>(v);
    if (static_cast<double>(f) < v) f = std::nextafterf(f, INFINITY);
    return f;
  }

  // A unit is a single node (root or nested-subtree root, padded to two slots) or a sibling pair.
  struct Unit { int a, b; uint32_t depth; };

  void PushChildren(const Unit& u, std::vector<Unit>& out) const {
    for (int n : {u.a, u.b}) {
      if (n < 0) continue;
      const FNode& f = fnodes_[n];
      if (f.left >= 0) {
        out.push_back(Unit{f.left, f.right, u.depth + 1});
      } else {
        for (const Item& it : f.items)
          if (it.type == RTB_PRIM_SUBTREE) out.push_back(Unit{static_cast<int>(it.index), -1, u.depth + 1});
      }
    }
  }

  void Layout(int root) {
    // 1. order the units: breadth-first until the budget is used, then depth-first per pending unit
    std::vector<Unit> order;
    std::deque<Unit> bfs{Unit{root, -1, 1}};
    uint32_t assigned = 0, depth_max = 1;
    std::vector<Unit> kids;
    while (!bfs.empty() && assigned + 2 <= top_budget_) {
      const Unit u = bfs.front();
      bfs.pop_front();
      order.push_back(u);
      assigned += 2;
      kids.clear();
      PushChildren(u, kids);
      for (const Unit& k : kids) bfs.push_back(k);
    }
    const uint32_t top_nodes = assigned;
    std::vector<Unit> dfs(bfs.rbegin(), bfs.rend());
    while (!dfs.empty()) {
      const Unit u = dfs.back();
      dfs.pop_back();
      order.push_back(u);
      kids.clear();
      PushChildren(u, kids);
      for (auto it = kids.rbegin(); it != kids.rend(); ++it) dfs.push_back(*it);
    }
    // 2. device indices
    uint32_t next = 0;
    for (const Unit& u : order) {
      fnodes_[u.a].dev = next;
      if (u.b >= 0) fnodes_[u.b].dev = next + 1;
      next += 2;
      depth_max = std::max(depth_max, u.depth);
    }
    size_t sub_max = 0;
    for (const FNode& f : fnodes_) {
      size_t subs = 0;
      for (const Item& it : f.items) subs += it.type == RTB_PRIM_SUBTREE;
      sub_max = std::max(sub_max, subs);
    }
    if (depth_max + sub_max > 60) throw std::runtime_error("flatten: BVH deeper than the 64-entry traversal stack (bvh.h:80)");
    // 3. emit nodes, leaf refs and the permuted primitive tables
    // padding slots keep an inverted box: the slab test can never enter them
    out_.nodes.assign(next, RtbNode{{INFINITY, INFINITY, INFINITY}, 0, {-INFINITY, -INFINITY, -INFINITY}, 0});
    std::vector<uint32_t> new_s(out_.spheres.size(), kNone), new_t(out_.triangles.size(), kNone), new_r(out_.rects.size(), kNone);
    uint32_t ns = 0, nt = 0, nr = 0;
    for (const Unit& u : order) {
      for (int n : {u.a, u.b}) {
        if (n < 0) continue;
        const FNode& f = fnodes_[n];
        RtbNode& d = out_.nodes[f.dev];
        d.lo[0] = RoundDown(f.box.x.min_); d.lo[1] = RoundDown(f.box.y.min_); d.lo[2] = RoundDown(f.box.z.min_);
        d.hi[0] = RoundUp(f.box.x.max_);   d.hi[1] = RoundUp(f.box.y.max_);   d.hi[2] = RoundUp(f.box.z.max_);
        if (f.left >= 0) {
          d.a = fnodes_[f.left].dev;
          d.b = 0;
          continue;
        }
        d.a = static_cast<uint32_t>(out_.prim_refs.size());
        d.b = static_cast<uint32_t>(f.items.size());
        if (d.b == 0) {  // empty leaf (all entries were empty BVHs): an inverted box never hits
          d.lo[0] = d.lo[1] = d.lo[2] = INFINITY;
          d.hi[0] = d.hi[1] = d.hi[2] = -INFINITY;
          d.b = 1;
          out_.prim_refs.push_back(RTB_REF(RTB_PRIM_SUBTREE, f.dev));
          continue;
        }
        for (const Item& it : f.items) {
          uint32_t idx = 0;
          switch (it.type) {
            case RTB_PRIM_SPHERE: idx = new_s[it.index] = ns++; break;
            case RTB_PRIM_TRIANGLE: idx = new_t[it.index] = nt++; break;
            case RTB_PRIM_RECT: idx = new_r[it.index] = nr++; break;
            default: idx = fnodes_[it.index].dev; break;
          }
          out_.prim_refs.push_back(RTB_REF(it.type, idx));
        }
      }
    }
    Permute(out_.spheres, new_s); Permute(out_.sphere_mat, new_s); Permute(out_.sphere_gid, new_s); Permute(out_.sphere_rank, new_s);
    Permute(out_.triangles, new_t); Permute(out_.triangle_gid, new_t); Permute(out_.triangle_rank, new_t);
    Permute(out_.rects, new_r); Permute(out_.rect_gid, new_r); Permute(out_.rect_rank, new_r);
    for (uint32_t& ref : out_.gid_to_ref) {
      const uint32_t t = RTB_REF_TYPE(ref), i = RTB_REF_INDEX(ref);
      ref = RTB_REF(t, (t == RTB_PRIM_SPHERE ? new_s : t == RTB_PRIM_TRIANGLE ? new_t : new_r)[i]);
    }
    for (size_t i = 0; i < out_.images.size(); ++i) out_.images[i].rgb = out_.image_bytes[i].empty() ? nullptr : out_.image_bytes[i].data();
    out_.n_top_nodes = top_nodes;
    out_.max_depth = depth_max;
  }

  template <class T>
  static void Permute(std::vector<T>& v, const std::vector<uint32_t>& new_index) {
    std::vector<T> tmp(v.size());
    for (size_t i = 0; i < v.size(); ++i) {
      if (new_index[i] == kNone) throw std::runtime_error("flatten: primitive not referenced by any leaf");
      tmp[new_index[i]] = v[i];
    }
    v.swap(tmp);
  }

  uint32_t top_budget_;
  bool keep_info_ = false;
  bool soup_ = false;
  FlatScene out_;
  std::vector<FNode> fnodes_;
  Item last_{kNone, 0};
  RtbMaterial mat_tmp_{};
  RtbTexture tex_tmp_{};
  double mat_d_[5] = {0, 0, 0, 0, 0};
  double tex_d_[4] = {0, 0, 0, 0};
  std::unordered_map<const material::Material*, uint32_t> mat_ids_;
  std::unordered_map<const material::Texture*, uint32_t> tex_ids_;
  const material::Material* last_mat_ = nullptr;
  uint32_t last_mat_id_ = 0;
};

}  // namespace rt::b200
