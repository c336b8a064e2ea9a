// rtb_device.cuh — device-side scene view, RNG, exact primitive tests and the BVH traversal.
//
// Numerical contract (DESIGN.md "Exactness"): the closest hit returned is the one
// geom::Bvh::Hit (reference src/geom/bvh.h:71-119) returns for the same float ray widened to
// double.  Box culling runs in FP32 and is only ever CONSERVATIVE (outward-rounded boxes plus an
// error slack), while every primitive decision is an FP64/FP32 replay of the reference's own
// arithmetic, operation by operation, with contraction disabled (__dmul_rn/__dadd_rn are never
// fused).  Because the accepted set and the t values are then exactly the reference's, the
// winner is a pure function of (t, tie rule) and does not depend on traversal order, which
// frees the kernel to walk near-child-first.
#pragma once
#include <cuda_runtime.h>
#include <rtb.h>
#include <stdint.h>

namespace rtb {

// ---------------------------------------------------------------------------------------------
struct DevScene {
  const float4* nodes;        // 2 x float4 per RtbNode
  const uint32_t* prim_refs;
  const double2* spheres;     // RtbSphere  = 2 x double2
  const double2* triangles;   // RtbTriangle = 5 x double2 (last .y holds mat/pad bits)
  const double2* rects;       // RtbRect = 3 x double2
  const float4* sphere_f;     // (cx,cy,cz,r) float, shading + pre-filter
  const float4* triangle_f;   // 3 x float4: (a, mat bits) (e1, 0) (e2, 0)
  const uint32_t* sphere_mat;
  const uint32_t* sphere_gid;
  const uint32_t* triangle_gid;
  const uint32_t* rect_gid;
  const uint32_t* sphere_rank;
  const uint32_t* triangle_rank;
  const uint32_t* rect_rank;
  const RtbMaterial* materials;
  const RtbTexture* textures;
  const cudaTextureObject_t* image_tex;  // one per image (0 when the image failed to load)
  const int2* image_dim;
  uint32_t n_nodes, n_top_nodes;
};

constexpr uint32_t kNoHit = 0xffffffffu;
constexpr uint32_t kLeafBit = 0x80000000u;
constexpr int kStackSize = 64;  // the reference's own bound (bvh.h:80); the flattener enforces it

struct HitCounters {  // only touched when RTB_COUNTERS is defined
  unsigned long long node_pairs, sphere_tests, triangle_tests, rect_tests, rays;
};

// ---------------------------------------------------------------------------------------------
// Philox4x32-10, counter = (pixel, sample, bounce, stream), key = seed.
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    const uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0;
    key.y += W1;
  }
  return ctr;
}
__device__ __forceinline__ float u01(uint32_t x) { return (x >> 8) * (1.0f / 16777216.0f); }  // [0,1)

// ---------------------------------------------------------------------------------------------
// Ray prepared for traversal.
struct TravRay {
  float ox, oy, oz, dx, dy, dz;
  float tmin, tmax;
  // FP32 slab constants: t_near_i = fma(plane_near_i, inv_i, bn_i), t_far_i = fma(plane_far_i, inv_i, bf_i)
  float ix, iy, iz;
  float bnx, bny, bnz, bfx, bfy, bfz;
  // FP64 replay constants
  double a;  // |d|^2 exactly as Vec3::length_squared sums it (vec3.h:47-49)
};

__device__ __forceinline__ float safe_dir(float d) {
  // Box test only: |d| below 1e-20 behaves like an axis-parallel ray; keeps 1/d and o/d finite.
  return fabsf(d) < 1e-20f ? copysignf(1e-20f, d) : d;
}

__device__ __forceinline__ void prepare_ray(TravRay& r, float4 o, float4 d) {
  r.ox = o.x; r.oy = o.y; r.oz = o.z; r.tmin = o.w;
  r.dx = d.x; r.dy = d.y; r.dz = d.z; r.tmax = d.w;
  r.ix = __frcp_rn(safe_dir(d.x));
  r.iy = __frcp_rn(safe_dir(d.y));
  r.iz = __frcp_rn(safe_dir(d.z));
  // computed t = fma(p, inv, -o*inv); |error| <= ~2^-23 (|o*inv| + |t|).  The |o*inv| part is
  // folded into the per-axis constants here, the relative part is applied in box_hit().
  const float kAbs = 4.0f * 1.1920929e-7f;
  const float ox = o.x * r.ix, oy = o.y * r.iy, oz = o.z * r.iz;
  const float ex = kAbs * fabsf(ox), ey = kAbs * fabsf(oy), ez = kAbs * fabsf(oz);
  r.bnx = -ox - ex; r.bny = -oy - ey; r.bnz = -oz - ez;
  r.bfx = -ox + ex; r.bfy = -oy + ey; r.bfz = -oz + ez;
  const double ddx = d.x, ddy = d.y, ddz = d.z;
  r.a = __dadd_rn(__dadd_rn(__dmul_rn(ddx, ddx), __dmul_rn(ddy, ddy)), __dmul_rn(ddz, ddz));
}

// Conservative slab test against [tmin, tbest].  Returns entry distance in tn.
__device__ __forceinline__ bool box_hit(const TravRay& r, float4 lo, float4 hi, float tbest, float& tn) {
  const float nx = r.ix >= 0.f ? lo.x : hi.x, fx = r.ix >= 0.f ? hi.x : lo.x;
  const float ny = r.iy >= 0.f ? lo.y : hi.y, fy = r.iy >= 0.f ? hi.y : lo.y;
  const float nz = r.iz >= 0.f ? lo.z : hi.z, fz = r.iz >= 0.f ? hi.z : lo.z;
  const float t0x = fmaf(nx, r.ix, r.bnx), t0y = fmaf(ny, r.iy, r.bny), t0z = fmaf(nz, r.iz, r.bnz);
  const float t1x = fmaf(fx, r.ix, r.bfx), t1y = fmaf(fy, r.iy, r.bfy), t1z = fmaf(fz, r.iz, r.bfz);
  tn = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, r.tmin));
  const float tf = fminf(fminf(t1x, t1y), fminf(t1z, tbest));
  // relative slack (~8 ulp each side); fminf/fmaxf drop a NaN operand, which is the conservative side
  return fmaf(fabsf(tn), -1e-6f, tn) <= fmaf(fabsf(tf), 1e-6f, tf);
}

// ---------------------------------------------------------------------------------------------
// Exact replays.  Each returns true with the reference's HitRecord::t in `t` when the
// reference's Hit() would return true for the interval (tmin, tmax) of the ORIGINAL ray.

// geom::Sphere::Hit, src/geom/sphere.h:23-55
__device__ __forceinline__ bool sphere_exact(const DevScene& s, uint32_t idx, const TravRay& r, double& t) {
  const double2 q0 = __ldg(&s.spheres[2 * idx]), q1 = __ldg(&s.spheres[2 * idx + 1]);
  const double ocx = __dsub_rn(q0.x, (double)r.ox), ocy = __dsub_rn(q0.y, (double)r.oy), ocz = __dsub_rn(q1.x, (double)r.oz);
  const double dx = r.dx, dy = r.dy, dz = r.dz;
  const double h = __dadd_rn(__dadd_rn(__dmul_rn(dx, ocx), __dmul_rn(dy, ocy)), __dmul_rn(dz, ocz));
  const double oc2 = __dadd_rn(__dadd_rn(__dmul_rn(ocx, ocx), __dmul_rn(ocy, ocy)), __dmul_rn(ocz, ocz));
  const double c = __dsub_rn(oc2, __dmul_rn(q1.y, q1.y));
  const double disc = __dsub_rn(__dmul_rn(h, h), __dmul_rn(r.a, c));
  if (disc < 0) return false;
  const double sq = __dsqrt_rn(disc);
  const double tmin = r.tmin, tmax = r.tmax;
  double root = __ddiv_rn(__dsub_rn(h, sq), r.a);
  if (!(tmin < root && root < tmax)) {
    root = __ddiv_rn(__dadd_rn(h, sq), r.a);
    if (!(tmin < root && root < tmax)) return false;
  }
  t = root;
  return true;
}

// geom::Triangle::Hit, src/geom/triangle.h:41-87 (double products, float det/inv_det/u/v/t)
__device__ __forceinline__ bool triangle_exact(const DevScene& s, uint32_t idx, const TravRay& r, double& t) {
  const double2* p = s.triangles + 5 * (size_t)idx;
  const double2 q0 = __ldg(p), q1 = __ldg(p + 1), q2 = __ldg(p + 2), q3 = __ldg(p + 3), q4 = __ldg(p + 4);
  const double ax = q0.x, ay = q0.y, az = q1.x, e1x = q1.y, e1y = q2.x, e1z = q2.y, e2x = q3.x, e2y = q3.y, e2z = q4.x;
  const double dx = r.dx, dy = r.dy, dz = r.dz;
  // pvec = Cross(dir, edge2)
  const double px = __dsub_rn(__dmul_rn(dy, e2z), __dmul_rn(dz, e2y));
  const double py = __dsub_rn(__dmul_rn(dz, e2x), __dmul_rn(dx, e2z));
  const double pz = __dsub_rn(__dmul_rn(dx, e2y), __dmul_rn(dy, e2x));
  const float det = __double2float_rn(__dadd_rn(__dadd_rn(__dmul_rn(e1x, px), __dmul_rn(e1y, py)), __dmul_rn(e1z, pz)));
  if (fabsf(det) < 1e-6f) return false;
  const float inv_det = __fdiv_rn(1.0f, det);
  const double inv = inv_det;
  const double tx = __dsub_rn((double)r.ox, ax), ty = __dsub_rn((double)r.oy, ay), tz = __dsub_rn((double)r.oz, az);
  const float u = __double2float_rn(__dmul_rn(__dadd_rn(__dadd_rn(__dmul_rn(tx, px), __dmul_rn(ty, py)), __dmul_rn(tz, pz)), inv));
  if (u < 0.0f || u > 1.0f) return false;
  // qvec = Cross(tvec, edge1)
  const double qx = __dsub_rn(__dmul_rn(ty, e1z), __dmul_rn(tz, e1y));
  const double qy = __dsub_rn(__dmul_rn(tz, e1x), __dmul_rn(tx, e1z));
  const double qz = __dsub_rn(__dmul_rn(tx, e1y), __dmul_rn(ty, e1x));
  const float v = __double2float_rn(__dmul_rn(__dadd_rn(__dadd_rn(__dmul_rn(dx, qx), __dmul_rn(dy, qy)), __dmul_rn(dz, qz)), inv));
  if (v < 0.0f || __fadd_rn(u, v) > 1.0f) return false;
  const float tf = __double2float_rn(__dmul_rn(__dadd_rn(__dadd_rn(__dmul_rn(e2x, qx), __dmul_rn(e2y, qy)), __dmul_rn(e2z, qz)), inv));
  const double td = tf;
  if (td < (double)r.tmin || td > (double)r.tmax) return false;
  t = td;
  return true;
}

// geom::xy_rect / xz_rect / yz_rect::Hit, src/geom/rect.h:19-40,65-85,109-130
__device__ __forceinline__ bool rect_exact(const DevScene& s, uint32_t idx, const TravRay& r, double& t) {
  const double2* p = s.rects + 3 * (size_t)idx;
  const double2 q0 = __ldg(p), q1 = __ldg(p + 1), q2 = __ldg(p + 2);  // (k,a0) (a1,b0) (b1, axis|mat)
  const int axis = (int)(__double_as_longlong(q2.y) & 0xffffffffll);
  // constant axis k, in-plane axes (a,b) = the other two in increasing order
  const double ok = axis == 0 ? r.ox : axis == 1 ? r.oy : r.oz, dk = axis == 0 ? r.dx : axis == 1 ? r.dy : r.dz;
  const double oa = axis == 0 ? r.oy : r.ox, da = axis == 0 ? r.dy : r.dx;
  const double ob = axis == 2 ? r.oy : r.oz, db = axis == 2 ? r.dy : r.dz;
  const double tt = __ddiv_rn(__dsub_rn(q0.x, ok), dk);
  if (!((double)r.tmin < tt && tt < (double)r.tmax)) return false;
  const double a = __dadd_rn(oa, __dmul_rn(tt, da));
  const double b = __dadd_rn(ob, __dmul_rn(tt, db));
  if (a < q0.y || a > q1.x || b < q1.y || b > q2.x) return false;
  t = tt;
  return true;
}

__device__ __forceinline__ uint32_t rank_of(const DevScene& s, uint32_t ref) {
  const uint32_t ty = RTB_REF_TYPE(ref), i = RTB_REF_INDEX(ref);
  return ty == RTB_PRIM_SPHERE ? __ldg(&s.sphere_rank[i]) : ty == RTB_PRIM_TRIANGLE ? __ldg(&s.triangle_rank[i]) : __ldg(&s.rect_rank[i]);
}

// Does candidate (t, ref) replace (best_t, best_ref)?  Strictly smaller t wins.  On an exact
// tie the reference's slot order decides (bvh.h:95-106 visits slots in rank order with the
// interval (tmin, closest)): Triangle::Hit accepts t == closest (triangle.h:74), Sphere/rect
// use the strict Interval::Surrounds (sphere.h:39-41, rect.h:22) and so never replace.
__device__ __forceinline__ bool replaces(const DevScene& s, double t, uint32_t ref, double best_t, uint32_t best_ref) {
  if (t > best_t) return false;
  if (t < best_t || best_ref == kNoHit) return true;  // t == tmax with nothing hit yet: the prim test already applied its own bound
  const bool tri_new = RTB_REF_TYPE(ref) == RTB_PRIM_TRIANGLE, tri_old = RTB_REF_TYPE(best_ref) == RTB_PRIM_TRIANGLE;
  if (tri_new != tri_old) return tri_new;
  const uint32_t rn = rank_of(s, ref), ro = rank_of(s, best_ref);
  return tri_new ? rn > ro : rn < ro;
}

// ---------------------------------------------------------------------------------------------
template <bool kUseSmem>
__device__ __forceinline__ void load_pair(const DevScene& s, const float4* smem_nodes, uint32_t pair, float4& llo, float4& lhi,
                                          float4& rlo, float4& rhi) {
  if (kUseSmem && pair < s.n_top_nodes) {
    const float4* p = smem_nodes + 2 * (size_t)pair;
    llo = p[0]; lhi = p[1]; rlo = p[2]; rhi = p[3];
  } else {
    const float4* p = s.nodes + 2 * (size_t)pair;
    llo = __ldg(p); lhi = __ldg(p + 1); rlo = __ldg(p + 2); rhi = __ldg(p + 3);
  }
}

__device__ __forceinline__ uint32_t node_desc(float4 lo, float4 hi) {
  const uint32_t a = __float_as_uint(lo.w), b = __float_as_uint(hi.w);
  return b == 0 ? a : (kLeafBit | (b << 25) | a);
}

// Closest hit.  Returns the primitive ref (kNoHit on a miss) and the reference's t.
template <bool kUseSmem>
__device__ __forceinline__ uint32_t closest_hit(const DevScene& s, const float4* smem_nodes, const TravRay& r, double& t_out,
                                                HitCounters* cnt) {
  uint32_t stack[kStackSize];
  int sp = 0;
  double best_t = r.tmax;
  float best_f = r.tmax;  // float upper bound of best_t for the box test
  uint32_t best_ref = kNoHit;
  uint32_t cur = 0;  // pair 0 = {root, padding}
#ifdef RTB_COUNTERS
  unsigned n_pairs = 0, n_sph = 0, n_tri = 0, n_rect = 0;
#endif
  for (;;) {
    // ---- descend through internal nodes
    while (!(cur & kLeafBit)) {
      float4 llo, lhi, rlo, rhi;
      load_pair<kUseSmem>(s, smem_nodes, cur, llo, lhi, rlo, rhi);
#ifdef RTB_COUNTERS
      ++n_pairs;
#endif
      float tl, tr;
      const bool hl = box_hit(r, llo, lhi, best_f, tl);
      const bool hr = box_hit(r, rlo, rhi, best_f, tr);
      const uint32_t dl = node_desc(llo, lhi), dr = node_desc(rlo, rhi);
      if (hl && hr) {
        const bool left_first = tl <= tr;
        stack[sp++] = left_first ? dr : dl;
        cur = left_first ? dl : dr;
      } else if (hl || hr) {
        cur = hl ? dl : dr;
      } else {
        if (sp == 0) goto done;
        cur = stack[--sp];
      }
    }
    // ---- leaf
    {
      const uint32_t first = cur & 0x01ffffffu, count = (cur >> 25) & 63u;
      for (uint32_t i = 0; i < count; ++i) {
        const uint32_t ref = __ldg(&s.prim_refs[first + i]);
        const uint32_t ty = RTB_REF_TYPE(ref), idx = RTB_REF_INDEX(ref);
        if (ty == RTB_PRIM_SUBTREE) {
          stack[sp++] = idx;  // pair {subtree root, padding}
          continue;
        }
#ifdef RTB_COUNTERS
        if (ty == RTB_PRIM_TRIANGLE) ++n_tri; else if (ty == RTB_PRIM_SPHERE) ++n_sph; else ++n_rect;
#endif
        double t;
        bool ok;
        if (ty == RTB_PRIM_TRIANGLE) ok = triangle_exact(s, idx, r, t);
        else if (ty == RTB_PRIM_SPHERE) ok = sphere_exact(s, idx, r, t);
        else ok = rect_exact(s, idx, r, t);
        if (ok && replaces(s, t, ref, best_t, best_ref)) {
          best_t = t;
          best_ref = ref;
          best_f = __double2float_ru(t);
        }
      }
      if (sp == 0) goto done;
      cur = stack[--sp];
    }
  }
done:
#ifdef RTB_COUNTERS
  if (cnt) {
    atomicAdd(&cnt->node_pairs, (unsigned long long)n_pairs);
    atomicAdd(&cnt->sphere_tests, (unsigned long long)n_sph);
    atomicAdd(&cnt->triangle_tests, (unsigned long long)n_tri);
    atomicAdd(&cnt->rect_tests, (unsigned long long)n_rect);
    atomicAdd(&cnt->rays, 1ull);
  }
#else
  (void)cnt;
#endif
  t_out = best_t;
  return best_ref;
}

}  // namespace rtb
