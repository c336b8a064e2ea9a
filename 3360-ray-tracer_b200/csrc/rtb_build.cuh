// rtb_build.cuh — the traversal tree built ON THE DEVICE from the primitive tables.
//
// The reference builds its tree on one host thread (geom::Bvh, src/geom/bvh.h:166-344: 16-bin SAH, ~5 s for the
// 5 M-primitive stress scene, plus this library's own flatten + compile passes).  The closest hit is a pure
// function of the exact primitive tests and the tie rule (rtb_device.cuh), so any conservative tree gives the same
// hits; this one is a linear BVH (Karras 2012):
//   prim_bounds   conservative float box + leaf descriptor per primitive, scene centroid bounds
//   morton        63-bit Morton key of the box centre (cubic cells over the largest extent)
//   radix sort    (key, primitive) pairs - cub::DeviceRadixSort, the one library call
//   hierarchy     one thread per internal node: range and split from the longest common key prefix
//   emit          one thread per leaf walks up; the second arrival at a node unions its two children and carries
//                 on; each child writes its own half of the parent's 64-byte PAIR record, the layout trace_loop
//                 walks (rtb_compile.h), so no separate node array is ever materialised.
// Leaves hold one primitive (a primitive test costs several pair steps per lane), the descriptor
// points into the type's table in CREATION order, so the tables need no permutation.
#pragma once
#include <cub/device/device_radix_sort.cuh>

#include "rtb_device.cuh"

namespace rtb {

struct BuildBox {
  float lo[3], hi[3];
};

// Float shadows of the exact double tables (what the FP32 filters and shading read: rtb_device.cuh DevScene), computed on the
// device for scenes that are not served from a pinned arena image: the stress scene's 4 M triangles would otherwise cost a
// host pass over 320 MB and another 192 MB of H2D traffic per rtb_scene_create.  Bit-identical to the host loop in
// rtb_scene_create: round to nearest for the values, round UP for the max |component| bounds of the triangle filter.
__global__ void __launch_bounds__(256) shadow_spheres_kernel(const double2* __restrict__ spheres, uint32_t n, float4* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double2 q0 = spheres[2 * (size_t)i], q1 = spheres[2 * (size_t)i + 1];
  out[i] = make_float4((float)q0.x, (float)q0.y, (float)q1.x, (float)q1.y);
}
__global__ void __launch_bounds__(256) shadow_triangles_kernel(const double2* __restrict__ tris, uint32_t n, float4* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double2* p = tris + 5 * (size_t)i;
  const double2 q0 = p[0], q1 = p[1], q2 = p[2], q3 = p[3], q4 = p[4];
  const float mat = __uint_as_float((uint32_t)(__double_as_longlong(q4.y) & 0xffffffffll));
  const double m1 = fmax(fabs(q1.y), fmax(fabs(q2.x), fabs(q2.y))), m2 = fmax(fabs(q3.x), fmax(fabs(q3.y), fabs(q4.x)));
  out[3 * (size_t)i] = make_float4((float)q0.x, (float)q0.y, (float)q1.x, mat);
  out[3 * (size_t)i + 1] = make_float4((float)q1.y, (float)q2.x, (float)q2.y, __double2float_ru(m1));
  out[3 * (size_t)i + 2] = make_float4((float)q3.x, (float)q3.y, (float)q4.x, __double2float_ru(m2));
}
__global__ void __launch_bounds__(256) shadow_rects_kernel(const double2* __restrict__ rects, uint32_t n, float4* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double2* p = rects + 3 * (size_t)i;
  const double2 q0 = p[0], q1 = p[1], q2 = p[2];  // (k,a0) (a1,b0) (b1, axis|mat)
  const long long bits = __double_as_longlong(q2.y);
  out[2 * (size_t)i] = make_float4((float)q0.x, (float)q0.y, (float)q1.x, (float)q1.y);
  out[2 * (size_t)i + 1] = make_float4((float)q2.x, __uint_as_float((uint32_t)(bits & 0xffffffffll)), __uint_as_float((uint32_t)((unsigned long long)bits >> 32)), 0.f);
}

// ValidateDesc's material-index checks for large scenes, on the device (a host pass over the stress scene's 5 M primitives is
// 22 ms of every rtb_scene_create): bad[0] counts primitives whose material index is out of range, rects also their axis.
__global__ void __launch_bounds__(256) validate_materials_kernel(DevScene s, uint32_t n_spheres, uint32_t n_triangles, uint32_t n_rects, uint32_t n_materials,
                                                                 unsigned* bad) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  bool ok = true;
  if (i < n_spheres) {
    ok = s.sphere_mat[i] < n_materials;
  } else if (i < n_spheres + n_triangles) {
    const double2 q4 = s.triangles[5 * (size_t)(i - n_spheres) + 4];
    ok = (uint32_t)(__double_as_longlong(q4.y) & 0xffffffffll) < n_materials;
  } else if (i < n_spheres + n_triangles + n_rects) {
    const long long bits = __double_as_longlong(s.rects[3 * (size_t)(i - n_spheres - n_triangles) + 2].y);
    const int axis = (int)(bits & 0xffffffffll);
    ok = (uint32_t)((unsigned long long)bits >> 32) < n_materials && axis >= 0 && axis <= 2;
  }
  if (!ok) atomicAdd(bad, 1u);
}

// float <-> unsigned with the same order (for atomicMin / atomicMax on floats of any sign)
__device__ __forceinline__ unsigned ordered_bits(float f) {
  const unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float ordered_float(unsigned o) {
  const unsigned u = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
#ifdef __CUDA_ARCH__
  return __uint_as_float(u);
#else
  float f;
  memcpy(&f, &u, 4);
  return f;
#endif
}

// rtb_compile.h PrimBox on the device: supersets of the reference's own boxes (sphere.h:17-21, triangle.h:18-39
// with its 1e-6 pad, rect.h:11-17 with its 1e-4 pad), rounded outward to float.
__device__ __forceinline__ BuildBox prim_box(const DevScene& s, uint32_t type, uint32_t i) {
  BuildBox b;
  if (type == RTB_PRIM_SPHERE) {
    const double2 q0 = s.spheres[2 * (size_t)i], q1 = s.spheres[2 * (size_t)i + 1];
    const double c[3] = {q0.x, q0.y, q1.x}, r = fabs(q1.y);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      b.lo[k] = __double2float_rd(c[k] - r);
      b.hi[k] = __double2float_ru(c[k] + r);
    }
  } else if (type == RTB_PRIM_TRIANGLE) {
    const double2* p = s.triangles + 5 * (size_t)i;
    const double2 q0 = p[0], q1 = p[1], q2 = p[2], q3 = p[3], q4 = p[4];
    const double a[3] = {q0.x, q0.y, q1.x}, e1[3] = {q1.y, q2.x, q2.y}, e2[3] = {q3.x, q3.y, q4.x};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const double v0 = a[k], v1 = a[k] + e1[k], v2 = a[k] + e2[k];
      const double lo = fmin(v0, fmin(v1, v2)), hi = fmax(v0, fmax(v1, v2));
      const double pad = 2e-6 + 1e-12 * fmax(fabs(lo), fabs(hi));
      b.lo[k] = __double2float_rd(lo - pad);
      b.hi[k] = __double2float_ru(hi + pad);
    }
  } else {
    const double2* p = s.rects + 3 * (size_t)i;
    const double2 q0 = p[0], q1 = p[1], q2 = p[2];  // (k,a0) (a1,b0) (b1, axis|mat)
    const int axis = (int)(__double_as_longlong(q2.y) & 0xffffffffll);
    const int ka = axis == 0 ? 1 : 0, kb = axis == 2 ? 1 : 2;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      double lo, hi;
      if (k == axis) { lo = q0.x - 2e-4; hi = q0.x + 2e-4; }
      else if (k == ka) { lo = q0.y; hi = q1.x; }
      else { (void)kb; lo = q1.y; hi = q2.x; }
      b.lo[k] = __double2float_rd(lo);
      b.hi[k] = __double2float_ru(hi);
    }
  }
  return b;
}

// Primitive i of the concatenation [spheres | triangles | rects] -> (type, index in its table).
__device__ __forceinline__ void split_index(uint32_t i, uint32_t n_s, uint32_t n_t, uint32_t& type, uint32_t& idx) {
  if (i < n_s) { type = RTB_PRIM_SPHERE; idx = i; }
  else if (i < n_s + n_t) { type = RTB_PRIM_TRIANGLE; idx = i - n_s; }
  else { type = RTB_PRIM_RECT; idx = i - n_s - n_t; }
}

// bounds[0..2] = min, bounds[3..5] = max of the box CENTRES, as ordered bits (initialised to ~0 / 0 by the host)
__global__ void __launch_bounds__(256) build_bounds_kernel(DevScene s, uint32_t n_s, uint32_t n_t, uint32_t n, BuildBox* __restrict__ boxes,
                                                           unsigned* bounds) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  float c[3] = {INFINITY, INFINITY, INFINITY};
  if (i < n) {
    uint32_t type, idx;
    split_index(i, n_s, n_t, type, idx);
    const BuildBox b = prim_box(s, type, idx);
    boxes[i] = b;
#pragma unroll
    for (int k = 0; k < 3; ++k) c[k] = 0.5f * b.lo[k] + 0.5f * b.hi[k];
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const float v = i < n ? c[k] : INFINITY, w = i < n ? c[k] : -INFINITY;
    const unsigned mn = __reduce_min_sync(0xffffffffu, ordered_bits(v)), mx = __reduce_max_sync(0xffffffffu, ordered_bits(w));
    if ((threadIdx.x & 31u) == 0u) {
      atomicMin(&bounds[k], mn);
      atomicMax(&bounds[3 + k], mx);
    }
  }
}

__device__ __forceinline__ unsigned long long expand21(unsigned long long x) {
  x &= 0x1fffffull;
  x = (x | x << 32) & 0x1f00000000ffffull;
  x = (x | x << 16) & 0x1f0000ff0000ffull;
  x = (x | x << 8) & 0x100f00f00f00f00full;
  x = (x | x << 4) & 0x10c30c30c30c30c3ull;
  x = (x | x << 2) & 0x1249249249249249ull;
  return x;
}

__global__ void __launch_bounds__(256) build_morton_kernel(const BuildBox* __restrict__ boxes, uint32_t n, const unsigned* __restrict__ bounds,
                                                           unsigned long long* __restrict__ keys, uint32_t* __restrict__ order) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float lo[3] = {ordered_float(bounds[0]), ordered_float(bounds[1]), ordered_float(bounds[2])};
  const float ext = fmaxf(fmaxf(ordered_float(bounds[3]) - lo[0], ordered_float(bounds[4]) - lo[1]), ordered_float(bounds[5]) - lo[2]);
  const float scale = ext > 0.0f ? 2097152.0f / ext : 0.0f;  // 2^21 cubic cells along the largest extent
  const BuildBox b = boxes[i];
  unsigned long long key = 0;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const float c = 0.5f * b.lo[k] + 0.5f * b.hi[k];
    const float q = fminf(fmaxf((c - lo[k]) * scale, 0.0f), 2097151.0f);
    key |= expand21((unsigned long long)q) << (2 - k);
  }
  keys[i] = key;
  order[i] = i;
}

// longest common prefix of the keys at sorted positions i and j (ties: the positions themselves), -1 out of range
__device__ __forceinline__ int build_delta(const unsigned long long* __restrict__ keys, int n, int i, int j) {
  if (j < 0 || j >= n) return -1;
  const unsigned long long a = keys[i], b = keys[j];
  return a == b ? 64 + __clz(i ^ j) : __clzll((long long)(a ^ b));
}

// Karras 2012, one thread per internal node i in [0, n-1).  links[2i], links[2i+1] = left / right child: an internal
// index, or kLeafBit | sorted leaf position.  parent_* = 2 * parent + side.
__global__ void __launch_bounds__(256) build_hierarchy_kernel(const unsigned long long* __restrict__ keys, int n, uint32_t* __restrict__ parent_internal,
                                                              uint32_t* __restrict__ parent_leaf) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1) return;
  const int d = build_delta(keys, n, i, i + 1) - build_delta(keys, n, i, i - 1) >= 0 ? 1 : -1;
  const int dmin = build_delta(keys, n, i, i - d);
  int lmax = 2;
  while (build_delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
  int l = 0;
  for (int t = lmax >> 1; t >= 1; t >>= 1)
    if (build_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
  const int j = i + l * d;
  const int dnode = build_delta(keys, n, i, j);
  int sft = 0, t = l;
  do {
    t = (t + 1) >> 1;
    if (build_delta(keys, n, i, i + (sft + t) * d) > dnode) sft += t;
  } while (t > 1);
  const int gamma = i + sft * d + min(d, 0);
  const int lo = min(i, j), hi = max(i, j);
  if (lo == gamma) parent_leaf[gamma] = 2u * (uint32_t)i;
  else parent_internal[gamma] = 2u * (uint32_t)i;
  if (hi == gamma + 1) parent_leaf[gamma + 1] = 2u * (uint32_t)i + 1u;
  else parent_internal[gamma + 1] = 2u * (uint32_t)i + 1u;
}

__device__ __forceinline__ void store_half(float* pair, uint32_t side, const BuildBox& b, uint32_t desc) {
  float* p = pair + 6 * side;
  p[0] = b.lo[0]; p[1] = b.hi[0];  // {lo.x hi.x lo.y hi.y lo.z hi.z}, rtb_compile.h Pair
  p[2] = b.lo[1]; p[3] = b.hi[1];
  p[4] = b.lo[2]; p[5] = b.hi[2];
  reinterpret_cast<uint32_t*>(pair)[12 + side] = desc;
}

// One thread per leaf (sorted position k).  pairs: (n-1) x 16 floats, flags: n-1 zeros, depth: one zero.
// Depth: every arrival leaves the HEIGHT of its subtree in its half's pad word (14 + side) of the pair record; the
// second arrival carries max(own, sibling) + 1 upward, so the value that reaches the root is the height of the whole
// tree (the length of one leaf-to-root chain, which is what the first version counted, is the depth of whichever
// leaf happened to finish last).  The traversal never reads the pad words.
__global__ void __launch_bounds__(256) build_emit_kernel(const BuildBox* __restrict__ boxes, const uint32_t* __restrict__ order, uint32_t n, uint32_t n_s,
                                                         uint32_t n_t, const uint32_t* __restrict__ parent_internal,
                                                         const uint32_t* __restrict__ parent_leaf, float* pairs, unsigned* flags, unsigned* depth) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const uint32_t prim = order[k];
  uint32_t type, idx;
  split_index(prim, n_s, n_t, type, idx);
  BuildBox box = boxes[prim];
  uint32_t desc = kLeafBit | (type << 29) | idx;  // count - 1 = 0
  uint32_t link = parent_leaf[k];
  unsigned levels = 0;  // height of the subtree this thread carries (a leaf: 0)
  for (;;) {
    const uint32_t p = link >> 1, side = link & 1u;
    float* pair = pairs + 16 * (size_t)p;
    store_half(pair, side, box, desc);
    reinterpret_cast<unsigned*>(pair)[14 + side] = levels;
    __threadfence();
    if (atomicAdd(&flags[p], 1u) == 0u) return;  // the sibling subtree is not finished: its last thread carries on
    // second arrival: the other half was written (and fenced) before its atomicAdd
    const float* o = pair + 6 * (side ^ 1u);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      box.lo[c] = fminf(box.lo[c], __ldcg(o + 2 * c));  // {lo.x hi.x lo.y hi.y lo.z hi.z}
      box.hi[c] = fmaxf(box.hi[c], __ldcg(o + 2 * c + 1));
    }
    desc = p;
    levels = max(levels, __ldcg(reinterpret_cast<const unsigned*>(pair) + 14 + (side ^ 1u))) + 1u;
    if (p == 0u) break;
    link = parent_internal[p];
  }
  atomicMax(depth, levels);
}

}  // namespace rtb
