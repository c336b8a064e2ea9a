// rtb_compile.h — host side of rtb_scene_create: RtbSceneDesc (the interchange layout of
// include/rtb.h) -> the device-native traversal layout.
//
// The closest hit the kernels return is a pure function of (exact primitive tests, tie rule)
// (rtb_device.cuh), so the device tree only has to be CONSERVATIVE, not equal to the reference's
// geom::Bvh (src/geom/bvh.h:166-367).  That freedom is used here:
//   * sibling PAIRS are the unit: one 64-byte record {box0 box1 | desc0 desc1 | pad}, box = {lo.x hi.x lo.y hi.y lo.z hi.z}, per
//     internal node, i.e. one 128-bit x4 fetch and both slab tests per traversal step;
//   * a child descriptor says everything needed to continue (no second fetch): internal -> pair
//     index; leaf -> primitive type, count and first index into that type's leaf-ordered table.
//     The prim_refs indirection of the interchange layout disappears;
//   * leaves are homogeneous: a reference leaf holding several primitive types (Cornell: 6 rects +
//     3 spheres in one SAH-refused leaf) or nested geom::Bvh objects (RTB_PRIM_SUBTREE) is expanded
//     into a small subtree, with boxes recomputed from the primitives, rounded outward;
//   * pair 0 holds the root's two children (traversal starts there);
//   * the first `top_pairs` records are in breadth-first order (the prefix the traversal kernels
//     stage in shared memory), the rest depth-first, near the order rays walk them.
#pragma once
#include <rtb.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <deque>
#include <stdexcept>
#include <vector>

namespace rtbc {

constexpr uint32_t kLeafBit = 0x80000000u;
constexpr uint32_t kMaxLeafCount = 32;      // 5-bit count-1 field
constexpr uint32_t kMaxFirst = 1u << 24;    // 24-bit first-index field

inline uint32_t LeafDesc(uint32_t type, uint32_t first, uint32_t count) {
  return kLeafBit | (type << 29) | ((count - 1u) << 24) | first;
}

struct Box {
  float lo[3], hi[3];
  void Grow(const Box& o) {
    for (int k = 0; k < 3; ++k) {
      lo[k] = std::min(lo[k], o.lo[k]);
      hi[k] = std::max(hi[k], o.hi[k]);
    }
  }
  static Box Empty() { return Box{{INFINITY, INFINITY, INFINITY}, {-INFINITY, -INFINITY, -INFINITY}}; }
};

// one float towards -inf / +inf (finite, non-NaN input; bit arithmetic instead of nextafterf: this runs per primitive)
inline float StepDown(float f) {
  uint32_t u;
  std::memcpy(&u, &f, 4);
  if (f > 0.f) --u;
  else if (f < 0.f) ++u;
  else u = 0x80000001u;  // below +-0: the smallest negative denormal
  std::memcpy(&f, &u, 4);
  return f;
}
inline float StepUp(float f) { return -StepDown(-f); }
inline float Down(double v) {
  float f = static_cast<float>(v);
  if (static_cast<double>(f) > v && std::isfinite(f)) f = StepDown(f);
  return f;
}
inline float Up(double v) {
  float f = static_cast<float>(v);
  if (static_cast<double>(f) < v && std::isfinite(f)) f = StepUp(f);
  return f;
}

struct Pair {  // 64 bytes, the device record: two 256-bit loads per traversal step
  float box0[6], box1[6];  // each {lo.x hi.x lo.y hi.y lo.z hi.z}: (lo, hi) of an axis is one aligned 64-bit pair (FADD2 / FMUL2 operand)
  uint32_t d0, d1;
  uint32_t pad[2];
};
static_assert(sizeof(Pair) == 64, "pair record is one 64-byte fetch");

struct Compiled {
  std::vector<Pair> pairs;  // pairs[0] = the root's children; empty for an empty scene
  uint32_t top_pairs = 0;
  uint32_t depth = 0;       // deepest chain of pairs (stack bound)
  // 4-wide form of the same tree (Run(.., wide = true); the round-2 experiment of DESIGN.md 3.9): quad q = records
  // quads[2q], quads[2q + 1] (128 bytes, four child slots); an internal slot's descriptor is a quad index
  std::vector<Pair> quads;
  uint32_t quad_stack = 0;  // most entries a traversal of the quads can have on its stack
};
constexpr uint32_t kEmptyQuadSlot = 0xfffffffdu;  // descriptor of an unused quad slot (never a leaf: type bits 3; never a quad index)

class Compiler {
 public:
  // leaf_max: most primitives a device leaf may hold (1..32); prim_cost: cost of one sphere / triangle / rect test in
  // pair steps (nullptr = the measured defaults: gpurun_out r1t sweep)
  explicit Compiler(const RtbSceneDesc& d, uint32_t leaf_max = 4, const float* prim_cost = nullptr)
      : d_(d), leaf_max_(std::min(std::max(leaf_max, 1u), kMaxLeafCount)) {
    const float def[3] = {2.f, 8.f, 4.f};
    for (int k = 0; k < 3; ++k) prim_cost_[k] = prim_cost ? prim_cost[k] : def[k];
  }

  // rebuild = true: ignore the interchange topology and build the device tree from the primitive tables alone with a
  // surface-area heuristic over all three axes (32 bins; every split position for small sets) and the device's costs.
  // The reference's builder bins the longest centroid axis only (bvh.h:232-250) and stops at 4 primitives.
  Compiled Run(bool rebuild = false, bool wide = false) {
    int root = rebuild ? Rebuild() : FromNode(0, 0);
    Compiled out;
    if (root < 0) return out;  // nothing to hit: no pairs, the kernels report a miss for every ray
    // Traversal starts at pair 0 = the root's two children (the root's own box is never tested: a
    // ray that misses it misses both children).  A scene that is a single leaf is split in two.
    if (nodes_[root].left < 0) {
      const uint32_t d = nodes_[root].desc, ty = (d >> 29) & 3u, first = d & 0x00ffffffu, count = ((d >> 24) & 31u) + 1u;
      const uint32_t half = count > 1 ? count / 2 : 1;  // one primitive: both children hold it (a self-tie never replaces)
      const int l = NewLeaf(RunBox(ty, first, half), ty, first, half);
      const int r = count > 1 ? NewLeaf(RunBox(ty, first + half, count - half), ty, first + half, count - half) : l;
      root = NewInternal(l, r);
    }
    Layout(root, out);
    if (wide) LayoutWide(root, out);
    return out;
  }

 private:
  struct CNode {
    Box box;
    int left = -1, right = -1;  // CNode indices, -1/-1 = leaf
    uint32_t desc = 0;          // leaf descriptor
    uint32_t pair = 0;          // device pair index of an internal node's children
  };

  int NewLeaf(const Box& b, uint32_t type, uint32_t first, uint32_t count) {
    if (first + count > kMaxFirst) throw std::runtime_error("more than 2^24 primitives of one type");
    if (count > 1) {
      // The reference stops at <= 4 primitives per leaf (bvh.h:204).  On the device a primitive test is an order of
      // magnitude dearer per lane than a box test (it runs at a handful of active lanes, ncu profiles/r01_path_*),
      // so leaves are split further wherever the surface-area heuristic with the device's costs says it pays.
      return SplitRun(type, first, count);
    }
    CNode n;
    n.box = b;
    n.desc = LeafDesc(type, first, count);
    nodes_.push_back(n);
    return static_cast<int>(nodes_.size()) - 1;
  }
  int NewInternal(int l, int r) {
    CNode n;
    n.box = nodes_[l].box;
    n.box.Grow(nodes_[r].box);
    n.left = l;
    n.right = r;
    nodes_.push_back(n);
    return static_cast<int>(nodes_.size()) - 1;
  }

  // Conservative float boxes from the exact primitive parameters (only used where a reference
  // leaf had to be regrouped; the reference's own boxes are geom/sphere.h:17-21, triangle.h:18-39
  // (1e-6 pad), rect.h:11-17 (1e-4 pad) — these are supersets of them up to rounding).
  Box PrimBox(uint32_t type, uint32_t i) const {
    Box b = Box::Empty();
    if (type == RTB_PRIM_SPHERE) {
      const RtbSphere& s = d_.spheres[i];
      const double r = std::fabs(s.r);
      for (int k = 0; k < 3; ++k) {
        b.lo[k] = Down(s.c[k] - r);
        b.hi[k] = Up(s.c[k] + r);
      }
    } else if (type == RTB_PRIM_TRIANGLE) {
      const RtbTriangle& t = d_.triangles[i];
      for (int k = 0; k < 3; ++k) {
        const double v0 = t.a[k], v1 = t.a[k] + t.e1[k], v2 = t.a[k] + t.e2[k];
        const double lo = std::min(v0, std::min(v1, v2)), hi = std::max(v0, std::max(v1, v2));
        const double pad = 2e-6 + 1e-12 * std::max(std::fabs(lo), std::fabs(hi));
        b.lo[k] = Down(lo - pad);
        b.hi[k] = Up(hi + pad);
      }
    } else {
      const RtbRect& r = d_.rects[i];
      const int ka = r.axis == 0 ? 1 : 0, kb = r.axis == 2 ? 1 : 2;
      b.lo[r.axis] = Down(r.k - 2e-4);
      b.hi[r.axis] = Up(r.k + 2e-4);
      b.lo[ka] = Down(r.a0); b.hi[ka] = Up(r.a1);
      b.lo[kb] = Down(r.b0); b.hi[kb] = Up(r.b1);
    }
    return b;
  }
  Box RunBox(uint32_t type, uint32_t first, uint32_t count) const {
    Box b = Box::Empty();
    for (uint32_t i = 0; i < count; ++i) b.Grow(PrimBox(type, first + i));
    return b;
  }

  // Tree over the primitives items[lo, hi) of one type, chosen by a surface-area heuristic with the DEVICE's
  // costs: one pair step = 1, one primitive test = prim_cost_[type] (a primitive test runs at a handful of active lanes
  // and is several times dearer per lane than a box test).  Small sets try every axis and split position; a set
  // stays one leaf when no split is cheaper (e.g. large overlapping spheres).  Returns the expected cost.
  static float Area(const Box& b) {
    const float dx = std::max(b.hi[0] - b.lo[0], 0.f), dy = std::max(b.hi[1] - b.lo[1], 0.f), dz = std::max(b.hi[2] - b.lo[2], 0.f);
    return 2.f * (dx * dy + dy * dz + dz * dx);
  }
  struct Item {  // one primitive of a run being split: its index in the type's table, box and centroid
    uint32_t idx;
    Box box;
    float c[3];
  };
  static Box ItemsBox(const Item* it, size_t lo, size_t hi) {
    Box b = Box::Empty();
    for (size_t i = lo; i < hi; ++i) b.Grow(it[i].box);
    return b;
  }
  template <class Less>
  static void SmallSort(Item* it, size_t lo, size_t hi, Less less) {  // insertion sort: runs are short (<= 4 almost always)
    if (hi - lo > 32) {
      std::sort(it + lo, it + hi, less);
      return;
    }
    for (size_t i = lo + 1; i < hi; ++i) {
      Item v = it[i];
      size_t k = i;
      for (; k > lo && less(v, it[k - 1]); --k) it[k] = it[k - 1];
      it[k] = v;
    }
  }
  int SplitRun(uint32_t type, uint32_t first, uint32_t count) {
    Item small[kMaxLeafCount];
    std::vector<Item> big;
    Item* items = small;
    if (count > kMaxLeafCount) {
      big.resize(count);
      items = big.data();
    }
    for (uint32_t i = 0; i < count; ++i) {
      Item& t = items[i];
      t.idx = first + i;
      t.box = PrimBox(type, t.idx);
      for (int k = 0; k < 3; ++k) t.c[k] = 0.5f * (t.box.lo[k] + t.box.hi[k]);
    }
    return SplitItems(type, items, 0, count, nullptr);
  }
  int SplitItems(uint32_t type, Item* items, size_t lo, size_t hi, float* cost_out) {
    const size_t n = hi - lo;
    auto by_index = [](const Item& a, const Item& b) { return a.idx < b.idx; };
    SmallSort(items, lo, hi, by_index);
    bool consecutive = true;
    for (size_t i = lo + 1; i < hi; ++i) consecutive = consecutive && items[i].idx == items[i - 1].idx + 1;
    const Box box = ItemsBox(items, lo, hi);
    const bool may_leaf = n <= leaf_max_ && consecutive;
    const float leaf_cost = prim_cost_[type] * static_cast<float>(n);
    auto make_leaf = [&]() {
      CNode c;
      c.box = box;
      c.desc = LeafDesc(type, items[lo].idx, static_cast<uint32_t>(n));
      nodes_.push_back(c);
      if (cost_out) *cost_out = leaf_cost;
      return static_cast<int>(nodes_.size()) - 1;
    };
    if (n == 1) return make_leaf();
    // candidate split: (axis, position) minimising area_L * n_L + area_R * n_R; all of them for small sets,
    // the median of the widest axis for large ones
    int best_axis = 0;
    size_t best_mid = lo + n / 2;
    auto by_axis = [](int axis) {
      return [axis](const Item& a, const Item& b) { return a.c[axis] < b.c[axis] || (a.c[axis] == b.c[axis] && a.idx < b.idx); };
    };
    if (n <= 16) {
      float best = INFINITY;
      for (int axis = 0; axis < 3; ++axis) {
        SmallSort(items, lo, hi, by_axis(axis));
        for (size_t k = 1; k < n; ++k) {
          const float c = Area(ItemsBox(items, lo, lo + k)) * static_cast<float>(k) + Area(ItemsBox(items, lo + k, hi)) * static_cast<float>(n - k);
          if (c < best) {
            best = c;
            best_axis = axis;
            best_mid = lo + k;
          }
        }
      }
    } else {
      float cmin[3] = {INFINITY, INFINITY, INFINITY}, cmax[3] = {-INFINITY, -INFINITY, -INFINITY};
      for (size_t i = lo; i < hi; ++i)
        for (int k = 0; k < 3; ++k) {
          cmin[k] = std::min(cmin[k], items[i].c[k]);
          cmax[k] = std::max(cmax[k], items[i].c[k]);
        }
      for (int k = 1; k < 3; ++k)
        if (cmax[k] - cmin[k] > cmax[best_axis] - cmin[best_axis]) best_axis = k;
    }
    SmallSort(items, lo, hi, by_axis(best_axis));
    const size_t mark = nodes_.size();
    float cl = 0.f, cr = 0.f;
    const int l = SplitItems(type, items, lo, best_mid, &cl), r = SplitItems(type, items, best_mid, hi, &cr);
    const float a = std::max(Area(box), 1e-30f);
    const float split_cost = 1.f + (Area(nodes_[l].box) * cl + Area(nodes_[r].box) * cr) / a;
    if (may_leaf && leaf_cost <= split_cost) {
      nodes_.resize(mark);  // drop the trial subtree
      SmallSort(items, lo, hi, by_index);
      return make_leaf();
    }
    if (cost_out) *cost_out = split_cost;
    return NewInternal(l, r);
  }

  // ---- full rebuild -------------------------------------------------------------------------------------------
  struct RItem {
    uint32_t type, idx;
    Box box;
    float c[3];
  };
  static Box RBox(const RItem* it, size_t lo, size_t hi) {
    Box b = Box::Empty();
    for (size_t i = lo; i < hi; ++i) b.Grow(it[i].box);
    return b;
  }
  int Rebuild() {
    std::vector<RItem> items;
    items.reserve((size_t)d_.n_spheres + d_.n_triangles + d_.n_rects);
    const uint32_t counts[3] = {d_.n_spheres, d_.n_triangles, d_.n_rects};
    for (uint32_t ty = 0; ty < 3; ++ty)
      for (uint32_t i = 0; i < counts[ty]; ++i) {
        RItem t;
        t.type = ty;
        t.idx = i;
        t.box = PrimBox(ty, i);
        for (int k = 0; k < 3; ++k) t.c[k] = 0.5f * t.box.lo[k] + 0.5f * t.box.hi[k];
        items.push_back(t);
      }
    if (items.empty()) return -1;
    if (items.size() > kMaxFirst) throw std::runtime_error("more than 2^24 primitives");
    return BuildSah(items.data(), 0, items.size(), nullptr);
  }
  float RLeafCost(const RItem* it, size_t lo, size_t hi) const {
    float c = 0.f;
    for (size_t i = lo; i < hi; ++i) c += prim_cost_[it[i].type];
    return c;
  }
  int BuildSah(RItem* it, size_t lo, size_t hi, float* cost_out) {
    const size_t n = hi - lo;
    const Box box = RBox(it, lo, hi);
    auto by_index = [](const RItem& a, const RItem& b) { return a.type < b.type || (a.type == b.type && a.idx < b.idx); };
    const float leaf_cost = RLeafCost(it, lo, hi);
    auto make_leaf = [&]() {
      CNode c;
      c.box = box;
      c.desc = LeafDesc(it[lo].type, it[lo].idx, static_cast<uint32_t>(n));
      nodes_.push_back(c);
      if (cost_out) *cost_out = leaf_cost;
      return static_cast<int>(nodes_.size()) - 1;
    };
    if (n == 1) return make_leaf();
    bool may_leaf = n <= leaf_max_;
    if (may_leaf) {
      std::sort(it + lo, it + hi, by_index);
      for (size_t i = lo + 1; i < hi; ++i) may_leaf = may_leaf && it[i].type == it[lo].type && it[i].idx == it[i - 1].idx + 1;
    }
    size_t best_mid = lo + n / 2;
    int best_axis = -1;
    auto by_axis = [](int axis) {
      return [axis](const RItem& a, const RItem& b) { return a.c[axis] < b.c[axis] || (a.c[axis] == b.c[axis] && (a.type < b.type || (a.type == b.type && a.idx < b.idx))); };
    };
    if (n <= 12) {  // every axis, every position
      float best = INFINITY;
      for (int axis = 0; axis < 3; ++axis) {
        std::sort(it + lo, it + hi, by_axis(axis));
        for (size_t k = 1; k < n; ++k) {
          const float c = Area(RBox(it, lo, lo + k)) * RLeafCost(it, lo, lo + k) + Area(RBox(it, lo + k, hi)) * RLeafCost(it, lo + k, hi);
          if (c < best) {
            best = c;
            best_axis = axis;
            best_mid = lo + k;
          }
        }
      }
      std::sort(it + lo, it + hi, by_axis(best_axis));
    } else {  // 32 centroid bins per axis
      constexpr int kBins = 32;
      float cmin[3] = {INFINITY, INFINITY, INFINITY}, cmax[3] = {-INFINITY, -INFINITY, -INFINITY};
      for (size_t i = lo; i < hi; ++i)
        for (int k = 0; k < 3; ++k) {
          cmin[k] = std::min(cmin[k], it[i].c[k]);
          cmax[k] = std::max(cmax[k], it[i].c[k]);
        }
      float best = INFINITY;
      int best_bin = -1;
      for (int axis = 0; axis < 3; ++axis) {
        const float ext = cmax[axis] - cmin[axis];
        if (!(ext > 0.f)) continue;
        const float scale = kBins / ext;
        Box bb[kBins];
        float bc[kBins];
        for (int b = 0; b < kBins; ++b) { bb[b] = Box::Empty(); bc[b] = 0.f; }
        for (size_t i = lo; i < hi; ++i) {
          const int b = std::min(kBins - 1, std::max(0, static_cast<int>((it[i].c[axis] - cmin[axis]) * scale)));
          bb[b].Grow(it[i].box);
          bc[b] += prim_cost_[it[i].type];
        }
        float la[kBins], lc[kBins];
        Box acc = Box::Empty();
        float cc = 0.f;
        for (int b = 0; b < kBins; ++b) { acc.Grow(bb[b]); cc += bc[b]; la[b] = cc > 0.f ? Area(acc) : 0.f; lc[b] = cc; }
        acc = Box::Empty();
        cc = 0.f;
        for (int b = kBins - 1; b >= 1; --b) {
          acc.Grow(bb[b]);
          cc += bc[b];
          if (cc <= 0.f || lc[b - 1] <= 0.f) continue;
          const float c = la[b - 1] * lc[b - 1] + Area(acc) * cc;
          if (c < best) { best = c; best_axis = axis; best_bin = b; }
        }
      }
      if (best_axis >= 0) {
        const float scale = kBins / (cmax[best_axis] - cmin[best_axis]), c0 = cmin[best_axis];
        const int axis = best_axis, bin = best_bin;
        RItem* mid = std::partition(it + lo, it + hi, [&](const RItem& t) {
          return std::min(kBins - 1, std::max(0, static_cast<int>((t.c[axis] - c0) * scale))) < bin;
        });
        best_mid = static_cast<size_t>(mid - it);
      }
      if (best_axis < 0 || best_mid == lo || best_mid == hi) {  // coincident centroids: halve in table order
        std::sort(it + lo, it + hi, by_index);
        best_mid = lo + n / 2;
      }
    }
    const size_t mark = nodes_.size();
    float cl = 0.f, cr = 0.f;
    const int l = BuildSah(it, lo, best_mid, &cl), r = BuildSah(it, best_mid, hi, &cr);
    const float a = std::max(Area(box), 1e-30f);
    const float split_cost = 1.f + (Area(nodes_[l].box) * cl + Area(nodes_[r].box) * cr) / a;
    if (may_leaf && leaf_cost <= split_cost) {
      nodes_.resize(mark);  // drop the trial subtree
      std::sort(it + lo, it + hi, by_index);
      return make_leaf();
    }
    if (cost_out) *cost_out = split_cost;
    return NewInternal(l, r);
  }

  // Balanced tree over a short list of already-built children.
  int Combine(std::vector<int>& kids, size_t lo, size_t hi) {
    if (hi - lo == 1) return kids[lo];
    const size_t mid = lo + (hi - lo) / 2;
    const int l = Combine(kids, lo, mid), r = Combine(kids, mid, hi);
    return NewInternal(l, r);
  }

  // Interchange node -> CNode index (-1 when the subtree holds nothing).
  int FromNode(uint32_t i, int guard) {
    if (guard > 256) throw std::runtime_error("scene nesting too deep (cycle in RTB_PRIM_SUBTREE refs?)");
    const RtbNode& n = d_.nodes[i];
    if (!(n.lo[0] <= n.hi[0]) && n.b == 0) return -1;  // padding slot
    Box nb;
    for (int k = 0; k < 3; ++k) {
      nb.lo[k] = n.lo[k];
      nb.hi[k] = n.hi[k];
    }
    if (n.b == 0) {
      const int l = FromNode(n.a, guard + 1), r = FromNode(n.a + 1, guard + 1);
      if (l < 0 || r < 0) return l < 0 ? r : l;
      const int id = NewInternal(l, r);
      nodes_[id].box = nb;  // the reference's (outward-rounded) box is already the tight union
      return id;
    }
    // leaf: group entries by type (each type's table indices are consecutive inside one leaf)
    uint32_t first[3] = {0, 0, 0}, count[3] = {0, 0, 0};
    std::vector<int> kids;
    for (uint32_t k = 0; k < n.b; ++k) {
      const uint32_t ref = d_.prim_refs[n.a + k], ty = RTB_REF_TYPE(ref), idx = RTB_REF_INDEX(ref);
      if (ty == RTB_PRIM_SUBTREE) {
        if (idx == i) continue;  // the flattener's "empty leaf" marker points at itself
        const int sub = FromNode(idx, guard + 1);
        if (sub >= 0) kids.push_back(sub);
        continue;
      }
      if (count[ty] == 0) first[ty] = idx;
      if (idx != first[ty] + count[ty]) throw std::runtime_error("leaf entries of one type must be consecutive in that type's table");
      ++count[ty];
    }
    const int n_types = (count[0] > 0) + (count[1] > 0) + (count[2] > 0);
    if (n_types == 1 && kids.empty()) {
      const uint32_t ty = count[0] ? 0u : count[1] ? 1u : 2u;
      return NewLeaf(nb, ty, first[ty], count[ty]);
    }
    for (uint32_t ty = 0; ty < 3; ++ty)
      if (count[ty]) kids.push_back(NewLeaf(RunBox(ty, first[ty], count[ty]), ty, first[ty], count[ty]));
    if (kids.empty()) return -1;
    return Combine(kids, 0, kids.size());
  }

  static void SetChild(Pair& p, int side, const Box& b, uint32_t desc) {
    float* box = side ? p.box1 : p.box0;
    for (int k = 0; k < 3; ++k) {
      box[2 * k] = b.lo[k];
      box[2 * k + 1] = b.hi[k];
    }
    (side ? p.d1 : p.d0) = desc;
    p.pad[0] = p.pad[1] = 0;
  }

  void Layout(int root, Compiled& out) {
    // pair 0 holds the root itself next to an empty sibling; every internal node owns one pair
    struct Unit { int a, b; uint32_t depth; };  // the two sibling CNodes stored in this pair
    const uint32_t budget = std::max<uint32_t>(1u, d_.n_top_nodes / 2);
    std::vector<Unit> order;
    std::deque<Unit> bfs{Unit{nodes_[root].left, nodes_[root].right, 1}};
    auto children = [&](const Unit& u, std::vector<Unit>& kids) {
      kids.clear();
      for (int n : {u.a, u.b})
        if (nodes_[n].left >= 0) kids.push_back(Unit{nodes_[n].left, nodes_[n].right, u.depth + 1});
    };
    std::vector<Unit> kids;
    while (!bfs.empty() && order.size() < budget) {
      const Unit u = bfs.front();
      bfs.pop_front();
      order.push_back(u);
      children(u, kids);
      for (const Unit& k : kids) bfs.push_back(k);
    }
    out.top_pairs = static_cast<uint32_t>(order.size());
    std::vector<Unit> dfs(bfs.rbegin(), bfs.rend());
    while (!dfs.empty()) {
      const Unit u = dfs.back();
      dfs.pop_back();
      order.push_back(u);
      children(u, kids);
      for (auto it = kids.rbegin(); it != kids.rend(); ++it) dfs.push_back(*it);
    }
    if (order.size() >= kLeafBit) throw std::runtime_error("too many BVH nodes");
    // a node's children live in the pair whose .a is its left child
    for (size_t p = 0; p < order.size(); ++p) {
      out.depth = std::max(out.depth, order[p].depth);
      pair_of_left_[order[p].a] = static_cast<uint32_t>(p);
    }
    out.pairs.resize(order.size());
    for (size_t p = 0; p < order.size(); ++p) {
      const Unit& u = order[p];
      int side = 0;
      for (int n : {u.a, u.b}) {
        const CNode& c = nodes_[n];
        SetChild(out.pairs[p], side++, c.box, c.left >= 0 ? pair_of_left_.at(c.left) : c.desc);
      }
    }
  }

  // The binary tree collapsed to 4-wide nodes: a quad holds a node's two children, the child with the largest box replaced by
  // ITS two children while slots are free (so a quad's slots are 2-4 descendants that partition the node's primitives).
  // Quads in depth-first order, slot order irrelevant (the traversal sorts the slots a ray enters by entry distance).
  void LayoutWide(int root, Compiled& out) {
    struct Quad { int slot[4]; int n; };
    auto collapse = [&](int node) {
      Quad q{{nodes_[node].left, nodes_[node].right, -1, -1}, 2};
      while (q.n < 4) {
        int best = -1;
        float best_area = -1.f;
        for (int k = 0; k < q.n; ++k)
          if (nodes_[q.slot[k]].left >= 0 && Area(nodes_[q.slot[k]].box) > best_area) {
            best = k;
            best_area = Area(nodes_[q.slot[k]].box);
          }
        if (best < 0) break;
        const int opened = q.slot[best];
        q.slot[best] = nodes_[opened].left;
        q.slot[q.n++] = nodes_[opened].right;
      }
      return q;
    };
    std::vector<Quad> order;
    std::vector<int> todo{root};
    IntMap quad_of;
    while (!todo.empty()) {
      const int node = todo.back();
      todo.pop_back();
      quad_of[node] = static_cast<uint32_t>(order.size());
      order.push_back(collapse(node));
      const Quad q = order.back();
      for (int k = q.n - 1; k >= 0; --k)
        if (nodes_[q.slot[k]].left >= 0) todo.push_back(q.slot[k]);
    }
    if (order.size() >= (1u << 28)) throw std::runtime_error("too many BVH nodes");
    out.quads.assign(2 * order.size(), Pair{});
    std::vector<uint32_t> need(order.size(), 0);  // stack entries a traversal below quad q can hold: children come after parents
    for (size_t p = order.size(); p-- > 0;) {
      const Quad& q = order[p];
      uint32_t below = 0;
      for (int k = 0; k < 4; ++k) {
        Pair& rec = out.quads[2 * p + k / 2];
        if (k >= q.n) {
          SetChild(rec, k & 1, Box{{0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}}, kEmptyQuadSlot);
          continue;
        }
        const CNode& c = nodes_[q.slot[k]];
        if (c.left >= 0) below = std::max(below, need[quad_of.at(q.slot[k])]);
        SetChild(rec, k & 1, c.box, c.left >= 0 ? quad_of.at(q.slot[k]) : c.desc);
      }
      need[p] = static_cast<uint32_t>(q.n - 1) + below;
    }
    out.quad_stack = need.empty() ? 0 : need[0];
  }

  const RtbSceneDesc& d_;
  uint32_t leaf_max_;
  float prim_cost_[3];
  std::vector<CNode> nodes_;
  struct IntMap {  // CNode index -> pair index (dense)
    std::vector<uint32_t> v;
    uint32_t& operator[](int i) {
      if (static_cast<size_t>(i) >= v.size()) v.resize(i + 1, 0);
      return v[i];
    }
    uint32_t at(int i) const { return v.at(i); }
  } pair_of_left_;
};

}  // namespace rtbc
