// rtb_device.cuh — device-side scene view, RNG, primitive tests (FP32 filters + exact FP64
// replays) and the BVH traversal loop.
//
// Numerical contract (DESIGN.md "Exactness"): the closest hit returned is the one
// geom::Bvh::Hit (reference src/geom/bvh.h:71-119) returns for the same float ray widened to
// double.  Three layers make that cheap:
//   1. box culling runs in FP32 and is only ever CONSERVATIVE (outward-rounded boxes, slack on
//      the slab comparison);
//   2. every primitive is first classified by an FP32 (spheres: mixed FP64/FP32) FILTER that
//      carries a rigorous running error bound: MISS (the reference certainly rejects), HIT with an
//      interval [t-e, t+e] that certainly contains the reference's t, or UNSURE;
//   3. only UNSURE candidates, and HITs whose interval overlaps the current best's, are decided
//      by an EXACT replay of the reference's own arithmetic, operation by operation, with
//      contraction disabled (__dmul_rn/__dadd_rn are never fused).
// Because the accepted set and the order of the t values are then exactly the reference's, the
// winner is a pure function of (t, tie rule) and does not depend on traversal order or on the
// shape of the device tree, which frees the kernel to walk near-child-first over its own layout.
#pragma once
#include <cuda_runtime.h>
#include <rtb.h>
#include <stdint.h>

namespace rtb {

// ---------------------------------------------------------------------------------------------
struct DevScene {
  const float4* pairs;        // 4 x float4 per sibling pair (rtb_compile.h)
  const double2* spheres;     // RtbSphere  = 2 x double2
  const double2* triangles;   // RtbTriangle = 5 x double2 (last .y holds mat/pad bits)
  const double2* rects;       // RtbRect = 3 x double2
  const float4* sphere_f;     // (cx,cy,cz,r) float, shading frame
  const float4* triangle_f;   // 3 x float4: (a, mat bits) (e1, max|e1|) (e2, max|e2|)
  const float4* rect_f;       // 2 x float4: (k, a0, a1, b0) (b1, axis bits, mat bits, 0)
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
  uint32_t n_pairs, n_top_pairs;
  uint32_t n_spheres, n_triangles, n_rects, n_materials, n_textures;  // table sizes: only read by the RTB_CHECKED build
  uint32_t n_quads;
  const float4* quads;        // the 4-wide form of the tree, 8 x float4 per quad (RTB_WIDE4 at compile time; nullptr otherwise)
};

// RTB_CHECKED build (make VARIANT=checked EXTRA=-DRTB_CHECKED): every index the kernels form is compared with the size of what
// it indexes and violations are COUNTED per class in g_check_fail (read by rtb_debug_check_failures), the access itself is
// clamped.  This is the library's own bounds checker: compute-sanitizer is not available on the B200 pool this was
// developed on (profiles/r02_checked_build.txt).
enum { kChkStack, kChkPair, kChkPrim, kChkPixel, kChkMaterial, kChkChunk, kChkPoolSlot, kChkClasses = 8 };
#ifdef RTB_CHECKED
__device__ unsigned long long g_check_fail[kChkClasses];
#define RTB_CHECK(klass, cond)                          \
  do {                                                  \
    if (!(cond)) atomicAdd(&rtb::g_check_fail[klass], 1ull); \
  } while (0)
#else
#define RTB_CHECK(klass, cond) \
  do {                         \
  } while (0)
#endif

constexpr uint32_t kNoHit = 0xffffffffu;
constexpr uint32_t kLeafBit = 0x80000000u;
constexpr uint32_t kEmptyQuadSlot = 0xfffffffdu;  // unused slot of a quad (rtb_compile.h)
constexpr int kStackSize = 64;  // the reference's own bound (bvh.h:80); rtb_scene_create enforces it
constexpr float kU = 5.9604645e-8f;  // 2^-24, FP32 unit roundoff

// leaf descriptor (rtb_compile.h): bit31 | type<<29 | (count-1)<<24 | first
__device__ __forceinline__ uint32_t desc_type(uint32_t d) { return (d >> 29) & 3u; }
__device__ __forceinline__ uint32_t desc_first(uint32_t d) { return d & 0x00ffffffu; }
__device__ __forceinline__ uint32_t desc_more(uint32_t d) { return d & 0x1f000000u; }  // count-1 != 0

struct HitCounters {  // only touched when RTB_COUNTERS is defined
  unsigned long long node_pairs, sphere_tests, triangle_tests, rect_tests, rays, exact_evals, filter_violations;
  unsigned long long max_steps, slow_rays;  // longest traversal (steps of one ray), rays above 256 steps
  float slow_ray[8];                        // one of them (o, tmin, d, tmax)
  unsigned long long step_hist[16];         // rays by floor(log2(steps + 1))
};

// ---------------------------------------------------------------------------------------------
// Philox4x32-10, counter = (pixel, sample, bounce, stream), key = seed.
#ifndef RTB_PHILOX_ROUNDS
#define RTB_PHILOX_ROUNDS 10
#endif
constexpr uint32_t kPhiloxW0 = 0x9E3779B9u, kPhiloxW1 = 0xBB67AE85u;
// The key schedule (key + r * W) depends on the seed only: the host writes it once per render into the kernel's
// parameter block (RenderConsts::round_keys), where every round reads it as a constant-bank operand — 2 integer adds
// per round and call less than stepping the key in registers.
struct PhiloxKeys {
  uint32_t k[2 * RTB_PHILOX_ROUNDS];
};
inline __host__ __device__ PhiloxKeys philox_key_schedule(uint32_t k0, uint32_t k1) {
  PhiloxKeys ks;
  for (int r = 0; r < RTB_PHILOX_ROUNDS; ++r) {
    ks.k[2 * r] = k0 + (uint32_t)r * kPhiloxW0;
    ks.k[2 * r + 1] = k1 + (uint32_t)r * kPhiloxW1;
  }
  return ks;
}
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, const PhiloxKeys& ks) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
#pragma unroll
  for (int r = 0; r < RTB_PHILOX_ROUNDS; ++r) {
    const uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    const uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ ks.k[2 * r], lo1, hi0 ^ ctr.w ^ ks.k[2 * r + 1], lo0);
  }
  return ctr;
}
__device__ __forceinline__ float u01(uint32_t x) { return (x >> 8) * (1.0f / 16777216.0f); }  // [0,1)

// ---------------------------------------------------------------------------------------------
#ifndef RTB_SLAB_FORM
#define RTB_SLAB_FORM 2  // the forms are listed in front of slab() below
#endif
// Per-lane ray state of the traversal loop.
struct TravRay {
  float ox, oy, oz, dx, dy, dz;
  float tmin, tmax;
  // FP32 slab constants: t_plane = fma(plane, inv, b) with b = -o*inv.  The only ABSOLUTE error of that form is
  // the rounding of b (<= u |o*inv|; it can be large: |o| ~ 500 and |d_k| ~ 1e-3 give ~0.03 in t), so each axis
  // carries two copies of b, pushed outward by its own bound: bl* is added to the box's lo plane, bh* to its hi
  // plane, ordered by the sign of inv so that min() lands on (near - e) and max() on (far + e).  An axis with a
  // tiny direction component therefore loosens only its own slab, never the other two.
  float ix, iy, iz;
#if RTB_SLAB_FORM == 0
  float blx, bly, blz, bhx, bhy, bhz;
#elif RTB_SLAB_FORM == 3
  unsigned long long bx2, by2, bz2;  // (b_lo, b_hi) per axis, one FFMA2 operand
#endif
};
// max |d_i|: the filters' error bounds scale with it (one FMNMX3, cheaper than a live register)
__device__ __forceinline__ float ray_dmax(const TravRay& r) { return fmaxf(fabsf(r.dx), fmaxf(fabsf(r.dy), fabsf(r.dz))); }

__device__ __forceinline__ float safe_dir(float d) {
  // Box test only: |d| below 1e-20 behaves like an axis-parallel ray; keeps 1/d and o/d finite.
  return fabsf(d) < 1e-20f ? copysignf(1e-20f, d) : d;
}

// Slab test forms (the box arrives as {lo.x hi.x lo.y hi.y lo.z hi.z}, each axis one aligned 64-bit pair):
//   0  t = fma(plane, inv, b), b = -o*inv pushed outward per axis: 12 FFMA per pair step, 9 constants per ray
//   1  t = (plane - o) * inv, scalar: 12 FADD + 12 FMUL, 3 constants besides the origin the leaf tests keep anyway
//   2  the same on the packed FP32 pipe: FADD2 / FMUL2 take (lo, hi) of an axis as one operand and the ray
//      constant as a broadcast scalar — 6 + 6 instructions per pair step and 6 registers fewer than form 0
//   3  form 0 as 6 FFMA2 with (b_lo, b_hi) as a register pair
#ifndef RTB_FLAT_UNIFORM_AXIS
#define RTB_FLAT_UNIFORM_AXIS 0  // 1: flat loop with one warp-uniform branch on the rect's axis instead of per-lane selects - measured SLOWER (Cornell 34.6 -> 36.5 ms)
#endif
#ifndef RTB_SLAB_BCAST
#define RTB_SLAB_BCAST 1  // 0: packs built outside the asm (the compiler hoists them into register pairs: +6 registers, measured slower)
#endif
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float a, float b) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void unpack2(f32x2 v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) { f32x2 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) { f32x2 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
// pair (op) scalar: the pack sits inside the asm block so that the compiler cannot hoist {s, s} out of the traversal loop as a
// 64-bit loop invariant (two registers per constant); ptxas folds it into the broadcast operand form `FADD2 Rd, Ra, -Rs.F32`
__device__ __forceinline__ f32x2 sub2s(f32x2 a, float s) { f32x2 d; asm("{\n .reg .b64 t;\n mov.b64 t, {%2, %2};\n sub.rn.f32x2 %0, %1, t;\n}" : "=l"(d) : "l"(a), "f"(s)); return d; }
__device__ __forceinline__ f32x2 mul2s(f32x2 a, float s) { f32x2 d; asm("{\n .reg .b64 t;\n mov.b64 t, {%2, %2};\n mul.rn.f32x2 %0, %1, t;\n}" : "=l"(d) : "l"(a), "f"(s)); return d; }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }

__device__ __forceinline__ void slab_axis(float o, float d, float& inv, float& bl, float& bh) {
  inv = __frcp_rn(safe_dir(d));
  // computed t = fma(p, inv, b), b = fl(-o*inv):  t_true = t (1 +- 2u) +- u |o*inv|.  4u covers b's rounding and
  // the rounding of the two sums below; the relative part is applied in slab().
  const float oi = o * inv, e = (4.0f * kU) * fabsf(oi);
  const float near_b = -oi - e, far_b = -oi + e;
  bl = inv >= 0.0f ? near_b : far_b;
  bh = inv >= 0.0f ? far_b : near_b;
}

__device__ __forceinline__ void prepare_ray(TravRay& r, float4 o, float4 d) {
  r.ox = o.x; r.oy = o.y; r.oz = o.z; r.tmin = o.w;
  r.dx = d.x; r.dy = d.y; r.dz = d.z; r.tmax = d.w;
#if RTB_SLAB_FORM == 1 || RTB_SLAB_FORM == 2
  r.ix = __frcp_rn(safe_dir(d.x));
  r.iy = __frcp_rn(safe_dir(d.y));
  r.iz = __frcp_rn(safe_dir(d.z));
#elif RTB_SLAB_FORM == 3
  float bl, bh;
  slab_axis(o.x, d.x, r.ix, bl, bh); r.bx2 = pack2(bl, bh);
  slab_axis(o.y, d.y, r.iy, bl, bh); r.by2 = pack2(bl, bh);
  slab_axis(o.z, d.z, r.iz, bl, bh); r.bz2 = pack2(bl, bh);
#else
  slab_axis(o.x, d.x, r.ix, r.blx, r.bhx);
  slab_axis(o.y, d.y, r.iy, r.bly, r.bhy);
  slab_axis(o.z, d.z, r.iz, r.blz, r.bhz);
#endif
}

// Conservative slab test of one box against [tmin, thi].  Returns the entry distance in tn.
// Forms 0 / 3: per axis min() is a lower bound of the true entry and max() an upper bound of the true exit up to 2^-22
// relative.  Forms 1 / 2: t = fl(fl(plane - o) * inv) with inv = fl(1 / d) is three roundings, all relative to t (<= 3u).
// Either way the comparison widens the exit by 2^-18 relative, so a box the exact (double) test of the reference
// enters is never culled.  fminf/fmaxf drop a NaN operand, which is the conservative side.
__device__ __forceinline__ bool slab(const TravRay& r, f32x2 px, f32x2 py, f32x2 pz, float thi, float& tn) {
  float ax, bx, ay, by, az, bz;
#if RTB_SLAB_FORM == 2
#if RTB_SLAB_BCAST
  unpack2(mul2s(sub2s(px, r.ox), r.ix), ax, bx);
  unpack2(mul2s(sub2s(py, r.oy), r.iy), ay, by);
  unpack2(mul2s(sub2s(pz, r.oz), r.iz), az, bz);
#else
  unpack2(mul2(add2(px, pack2(-r.ox, -r.ox)), pack2(r.ix, r.ix)), ax, bx);
  unpack2(mul2(add2(py, pack2(-r.oy, -r.oy)), pack2(r.iy, r.iy)), ay, by);
  unpack2(mul2(add2(pz, pack2(-r.oz, -r.oz)), pack2(r.iz, r.iz)), az, bz);
#endif
#elif RTB_SLAB_FORM == 3
  unpack2(fma2(px, pack2(r.ix, r.ix), r.bx2), ax, bx);
  unpack2(fma2(py, pack2(r.iy, r.iy), r.by2), ay, by);
  unpack2(fma2(pz, pack2(r.iz, r.iz), r.bz2), az, bz);
#else
  float lx, hx, ly, hy, lz, hz;
  unpack2(px, lx, hx); unpack2(py, ly, hy); unpack2(pz, lz, hz);
#if RTB_SLAB_FORM == 1
  ax = (lx - r.ox) * r.ix; bx = (hx - r.ox) * r.ix;
  ay = (ly - r.oy) * r.iy; by = (hy - r.oy) * r.iy;
  az = (lz - r.oz) * r.iz; bz = (hz - r.oz) * r.iz;
#else
  ax = fmaf(lx, r.ix, r.blx); bx = fmaf(hx, r.ix, r.bhx);
  ay = fmaf(ly, r.iy, r.bly); by = fmaf(hy, r.iy, r.bhy);
  az = fmaf(lz, r.iz, r.blz); bz = fmaf(hz, r.iz, r.bhz);
#endif
#endif
  tn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), r.tmin));
  const float tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), thi));
  return tn <= fmaf(fabsf(tf), 3.8e-6f, tf);  // 2^-18 >> the 4u a borderline pair (|tn| ~ |tf|) needs, any signs
}

// ---------------------------------------------------------------------------------------------
// Exact replays.  Each returns true with the reference's HitRecord::t in `t` when the
// reference's Hit() would return true for the interval (tmin, tmax) of the ORIGINAL ray.

struct ExactRay {
  double ox, oy, oz, dx, dy, dz, tmin, tmax;
};

// geom::Sphere::Hit, src/geom/sphere.h:23-55
__device__ __forceinline__ bool sphere_exact(const double2* __restrict__ table, uint32_t idx, const ExactRay& r, double& t) {
  const double2 q0 = __ldg(&table[2 * (size_t)idx]), q1 = __ldg(&table[2 * (size_t)idx + 1]);
  const double ocx = __dsub_rn(q0.x, r.ox), ocy = __dsub_rn(q0.y, r.oy), ocz = __dsub_rn(q1.x, r.oz);
  // |d|^2 exactly as Vec3::length_squared sums it (vec3.h:47-49)
  const double a = __dadd_rn(__dadd_rn(__dmul_rn(r.dx, r.dx), __dmul_rn(r.dy, r.dy)), __dmul_rn(r.dz, r.dz));
  const double h = __dadd_rn(__dadd_rn(__dmul_rn(r.dx, ocx), __dmul_rn(r.dy, ocy)), __dmul_rn(r.dz, ocz));
  const double oc2 = __dadd_rn(__dadd_rn(__dmul_rn(ocx, ocx), __dmul_rn(ocy, ocy)), __dmul_rn(ocz, ocz));
  const double c = __dsub_rn(oc2, __dmul_rn(q1.y, q1.y));
  const double disc = __dsub_rn(__dmul_rn(h, h), __dmul_rn(a, c));
  if (disc < 0) return false;
  const double sq = __dsqrt_rn(disc);
  double root = __ddiv_rn(__dsub_rn(h, sq), a);
  if (!(r.tmin < root && root < r.tmax)) {
    root = __ddiv_rn(__dadd_rn(h, sq), a);
    if (!(r.tmin < root && root < r.tmax)) return false;
  }
  t = root;
  return true;
}

// geom::Triangle::Hit, src/geom/triangle.h:41-87 (double products, float det/inv_det/u/v/t)
__device__ __forceinline__ bool triangle_exact(const double2* __restrict__ table, uint32_t idx, const ExactRay& r, double& t) {
  const double2* p = table + 5 * (size_t)idx;
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
  const double tx = __dsub_rn(r.ox, ax), ty = __dsub_rn(r.oy, ay), tz = __dsub_rn(r.oz, az);
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
  if (td < r.tmin || td > r.tmax) return false;
  t = td;
  return true;
}

// geom::xy_rect / xz_rect / yz_rect::Hit, src/geom/rect.h:19-40,65-85,109-130
__device__ __forceinline__ bool rect_exact(const double2* __restrict__ table, uint32_t idx, const ExactRay& r, double& t) {
  const double2* p = table + 3 * (size_t)idx;
  const double2 q0 = __ldg(p), q1 = __ldg(p + 1), q2 = __ldg(p + 2);  // (k,a0) (a1,b0) (b1, axis|mat)
  const int axis = (int)(__double_as_longlong(q2.y) & 0xffffffffll);
  // constant axis k, in-plane axes (a,b) = the other two in increasing order
  const double ok = axis == 0 ? r.ox : axis == 1 ? r.oy : r.oz, dk = axis == 0 ? r.dx : axis == 1 ? r.dy : r.dz;
  const double oa = axis == 0 ? r.oy : r.ox, da = axis == 0 ? r.dy : r.dx;
  const double ob = axis == 2 ? r.oy : r.oz, db = axis == 2 ? r.dy : r.dz;
  const double tt = __ddiv_rn(__dsub_rn(q0.x, ok), dk);
  if (!(r.tmin < tt && tt < r.tmax)) return false;
  const double a = __dadd_rn(oa, __dmul_rn(tt, da));
  const double b = __dadd_rn(ob, __dmul_rn(tt, db));
  if (a < q0.y || a > q1.x || b < q1.y || b > q2.x) return false;
  t = tt;
  return true;
}

// One out-of-line entry for all three replays: it is the rare path (see the filters below), so
// it must not bloat the traversal loop's registers or instruction footprint.  Everything travels
// in registers (no references): returns the reference's t, or NaN when the reference rejects.
__device__ __noinline__ double exact_eval(const double2* __restrict__ table, uint32_t type, uint32_t idx, float ox, float oy, float oz,
                                          float dx, float dy, float dz, float tmin, float tmax) {
  ExactRay r{(double)ox, (double)oy, (double)oz, (double)dx, (double)dy, (double)dz, (double)tmin, (double)tmax};
  double t = 0.0;
  bool ok;
  if (type == RTB_PRIM_TRIANGLE) ok = triangle_exact(table, idx, r, t);
  else if (type == RTB_PRIM_SPHERE) ok = sphere_exact(table, idx, r, t);
  else ok = rect_exact(table, idx, r, t);
  return ok ? t : __longlong_as_double(0x7ff8000000000000ll);
}
__device__ __forceinline__ const double2* exact_table(const DevScene& s, uint32_t type) {
  return type == RTB_PRIM_TRIANGLE ? s.triangles : type == RTB_PRIM_SPHERE ? s.spheres : s.rects;
}

// ---------------------------------------------------------------------------------------------
// Filters.  Result classes; for kHit, t/e satisfy |t_reference - t| <= e.
enum { kMiss = 0, kHit = 1, kUnsure = 2 };

// Triangle: FP32 Moeller-Trumbore on the float copies of (a, e1, e2) with a running error bound.
// With u = 2^-24, D = max|d_i|, M1 = max|e1_i|, M2 = max|e2_i|, MA = max|a_i|, MT = max|tvec_i|
// (float inputs carry <= u relative representation error, every FP32 op <= u relative):
//   |P - P^| <= 8u D M2 per component,  |DET - DET^| <= 48u D M1 M2,
//   |T - T^| <= u (MA + MT),            |NU - NU^| <= 60u D M2 (MA+MT),
//   |Q - Q^| <= 8u M1 (MA+MT),          |NV - NV^| <= 48u D M1 (MA+MT),  |NT - NT^| <= 54u M1 M2 (MA+MT).
// All are rounded up to 64u below.  The reference's own float roundings of det/inv_det/u/v/t
// (triangle.h:50-71) add <= 4u relative and are covered by the 8u term in rho.
__device__ __forceinline__ int triangle_filter(const DevScene& s, uint32_t idx, const TravRay& r, float thi, float& t_out, float& e_out) {
  const float4* p = s.triangle_f + 3 * (size_t)idx;
  const float4 A = __ldg(p), E1 = __ldg(p + 1), E2 = __ldg(p + 2);
  const float px = r.dy * E2.z - r.dz * E2.y, py = r.dz * E2.x - r.dx * E2.z, pz = r.dx * E2.y - r.dy * E2.x;
  const float det = E1.x * px + E1.y * py + E1.z * pz;
  const float w = E1.w * E2.w;
  const float dmax = ray_dmax(r);
  const float edet = (64.0f * kU) * w * dmax;
  const float adet = fabsf(det);
  if (adet + edet < 0.999999e-6f) return kMiss;  // triangle.h:55 rejects |det| < 1e-6
  if (!(adet > 8.0f * edet) || !(adet - edet > 1.000001e-6f)) return kUnsure;
  const float tx = r.ox - A.x, ty = r.oy - A.y, tz = r.oz - A.z;
  const float sA = (64.0f * kU) * (fmaxf(fabsf(A.x), fmaxf(fabsf(A.y), fabsf(A.z))) + fmaxf(fabsf(tx), fmaxf(fabsf(ty), fabsf(tz))));
  const float inv = __frcp_rn(det);
  const float g = 1.25f * fabsf(inv);
  const float rho = g * edet + 8.0f * kU;
  const float u = (tx * px + ty * py + tz * pz) * inv;
  const float eu = g * (sA * dmax * E2.w) + fabsf(u) * rho;
  if (u < -eu || u > 1.0f + eu) return kMiss;
  const float qx = ty * E1.z - tz * E1.y, qy = tz * E1.x - tx * E1.z, qz = tx * E1.y - ty * E1.x;
  const float v = (r.dx * qx + r.dy * qy + r.dz * qz) * inv;
  const float ev = g * (sA * dmax * E1.w) + fabsf(v) * rho;
  const float euv = eu + ev + 4.0f * kU;
  if (v < -ev || u + v > 1.0f + euv) return kMiss;
  const float t = (E2.x * qx + E2.y * qy + E2.z * qz) * inv;
  const float et = g * (sA * w) + fabsf(t) * rho;
  if (t + et < r.tmin || t - et > thi) return kMiss;
  t_out = t;
  e_out = et;
  const bool inside = u >= eu && v >= ev && u + v <= 1.0f - euv && t - et >= r.tmin && t + et <= r.tmax;
  return inside ? kHit : kUnsure;
}

// Sphere: the cancellation-prone part (oc, h, c, disc — sphere.h:25-31) in FP64 with FMA, then
// FP32 for the square root and the stable root forms (root1*root2 = c/a).  Relative error of
// each root <= 16u unless disc is within its own rounding noise of 0.
__device__ __forceinline__ int sphere_filter(const DevScene& s, uint32_t idx, const TravRay& r, float thi, float& t_out, float& e_out) {
  const double2 q0 = __ldg(&s.spheres[2 * idx]), q1 = __ldg(&s.spheres[2 * idx + 1]);
  const double ocx = q0.x - (double)r.ox, ocy = q0.y - (double)r.oy, ocz = q1.x - (double)r.oz;
  const double dx = r.dx, dy = r.dy, dz = r.dz;
  const double a = fma(dx, dx, fma(dy, dy, dz * dz));
  const double h = fma(dx, ocx, fma(dy, ocy, dz * ocz));
  const double oc2 = fma(ocx, ocx, fma(ocy, ocy, ocz * ocz));
  const double r2 = q1.y * q1.y;
  const double c = oc2 - r2;
  const double h2 = h * h, ac = a * c;
  const double disc = h2 - ac;
  // double rounding noise of disc (and of the reference's own, differently-ordered, unfused sums)
  const double noise = 1e-14 * (h2 + fabs(ac) + a * (oc2 + r2));
  if (disc < -noise) return kMiss;
  if (!(disc * 1e-3 > noise)) return kUnsure;  // grazing: sqrt amplifies the noise
  // relative error of hf +- sq (same signs: no cancellation).  The ratio is an error BOUND: a float quotient (2 ulp, covered by the
  // factor 1.001) instead of a double division, which was 5 % of the earth scene's instructions; denormal disc keeps the double form
  const float discf = (float)disc;
  const float rel_q = (discf > 1e-30f ? 0.5005f * __fdividef((float)noise, discf) : 0.5f * (float)(noise / disc)) + 4.0f * kU;
  const float hf = (float)h, af = (float)a, cf = (float)c;
  const float c_abs = (float)(4e-16 * (oc2 + r2));               // absolute error of c (origin on the surface: c ~ 0)
  const float sq = sqrtf(discf);
  // roots of a t^2 - 2 h t + c = 0 without cancellation (root1 * root2 = c / a)
  const float qv = hf >= 0.0f ? hf + sq : hf - sq;
  const float iq = __frcp_rn(qv);
  const float td = qv * __frcp_rn(af), tc = cf * iq;  // reciprocal + product: one more rounding than the quotient, in the bound below
  const float ed = fabsf(td) * (rel_q + 4.0f * kU) + 1e-37f;
  const float ec = fabsf(tc) * (rel_q + 4.0f * kU) + c_abs * fabsf(iq) * 1.01f + 1e-37f;
  const float t1 = hf >= 0.0f ? tc : td, e1 = hf >= 0.0f ? ec : ed;
  const float t2 = hf >= 0.0f ? td : tc, e2 = hf >= 0.0f ? ed : ec;
  // sphere.h:37-44: nearest root strictly inside (tmin, tmax), else the far root
  const float lim = fminf(thi, r.tmax);
  float t, e;
  if (t1 - e1 > r.tmin) {
    if (t1 - e1 > thi) return kMiss;  // both roots beyond the current best
    if (!(t1 + e1 < r.tmax)) return t1 - e1 >= r.tmax ? kMiss : kUnsure;
    t = t1; e = e1;
  } else {
    if (!(t1 + e1 < r.tmin)) return kUnsure;  // root1 straddles tmin
    if (t2 + e2 <= r.tmin || t2 - e2 > lim) return kMiss;
    if (!(t2 - e2 > r.tmin && t2 + e2 < r.tmax)) return kUnsure;
    t = t2; e = e2;
  }
  t_out = t;
  e_out = e;
  return kHit;
}

// The rect test proper for a rect whose axis is known at compile time (kAxis 0..2) or per lane (kAxis < 0, read from the record).
// kRayInv: take 1 / d_k from the ray's slab constants (live anyway in the tree loops) instead of computing it; the flat loop, where
// nothing else reads them, computes it (measured: keeping them alive there costs Cornell 7 %).
template <bool kRayInv, int kAxis>
__device__ __forceinline__ int rect_test(const TravRay& r, const float4 q0, const float4 q1, float thi, float& t_out, float& e_out) {
  const int axis = kAxis >= 0 ? kAxis : __float_as_int(q1.y);
  const float ok = axis == 0 ? r.ox : axis == 1 ? r.oy : r.oz, dk = axis == 0 ? r.dx : axis == 1 ? r.dy : r.dz;
  const float oa = axis == 0 ? r.oy : r.ox, da = axis == 0 ? r.dy : r.dx;
  const float ob = axis == 2 ? r.oy : r.oz, db = axis == 2 ? r.dy : r.dz;
  if (dk == 0.0f) return kMiss;  // t = +-inf or NaN: rect.h:22 rejects both
  if (kRayInv && fabsf(dk) < 1e-20f) return kUnsure;  // the slab constant is 1 / +-1e-20 there: the replay decides
  // 1 / d_k is the ray's slab constant (prepare_ray: __frcp_rn(safe_dir(d)) = __frcp_rn(d) for |d| >= 1e-20)
  const float diff = q0.x - ok, inv = kRayInv ? (axis == 0 ? r.ix : axis == 1 ? r.iy : r.iz) : __frcp_rn(dk);
  const float t = diff * inv;
  const float et = 4.0f * kU * ((fabsf(q0.x) + fabsf(diff)) * fabsf(inv) + fabsf(t));
  if (t + et <= r.tmin || t - et > fminf(thi, r.tmax)) return kMiss;
  const float a = fmaf(t, da, oa), b = fmaf(t, db, ob);
  const float ea = et * fabsf(da) + 4.0f * kU * (fabsf(oa) + fabsf(t * da)), eb = et * fabsf(db) + 4.0f * kU * (fabsf(ob) + fabsf(t * db));
  // bounds are float roundings of the doubles: one more ulp each
  const float a0 = q0.y, a1 = q0.z, b0 = q0.w, b1 = q1.x;
  const float sa0 = ea + kU * fabsf(a0), sa1 = ea + kU * fabsf(a1), sb0 = eb + kU * fabsf(b0), sb1 = eb + kU * fabsf(b1);
  if (a < a0 - sa0 || a > a1 + sa1 || b < b0 - sb0 || b > b1 + sb1) return kMiss;
  t_out = t;
  e_out = et;
  const bool inside = a >= a0 + sa0 && a <= a1 - sa1 && b >= b0 + sb0 && b <= b1 - sb1 && t - et > r.tmin && t + et < r.tmax;
  return inside ? kHit : kUnsure;
}

// Rect: FP32 throughout (rect.h:19-40).  |t - t^| <= (u|k| + u|k-o_k|)/|d_k| + 3u|t|.
// kRayInv = true: the tree loops (every lane stands at its own rect: the axis is selected per lane).  false: the flat loop, where all
// lanes test the SAME rect - the axis is warp-uniform, one branch picks the body specialised for it (no selects).
template <bool kRayInv = true>
__device__ __forceinline__ int rect_filter(const DevScene& s, uint32_t idx, const TravRay& r, float thi, float& t_out, float& e_out) {
  const float4 q0 = __ldg(&s.rect_f[2 * (size_t)idx]), q1 = __ldg(&s.rect_f[2 * (size_t)idx + 1]);  // (k,a0,a1,b0) (b1,axis,mat,-)
  if (kRayInv) return rect_test<true, -1>(r, q0, q1, thi, t_out, e_out);
#if RTB_FLAT_UNIFORM_AXIS
  const int axis = __float_as_int(q1.y);
  if (axis == 0) return rect_test<false, 0>(r, q0, q1, thi, t_out, e_out);
  if (axis == 1) return rect_test<false, 1>(r, q0, q1, thi, t_out, e_out);
  return rect_test<false, 2>(r, q0, q1, thi, t_out, e_out);
#else
  return rect_test<false, -1>(r, q0, q1, thi, t_out, e_out);
#endif
}

__device__ __forceinline__ uint32_t rank_of(const DevScene& s, uint32_t ref) {
  const uint32_t ty = RTB_REF_TYPE(ref), i = RTB_REF_INDEX(ref);
  return ty == RTB_PRIM_SPHERE ? __ldg(&s.sphere_rank[i]) : ty == RTB_PRIM_TRIANGLE ? __ldg(&s.triangle_rank[i]) : __ldg(&s.rect_rank[i]);
}

// Does candidate (t, ref) replace (best_t, best_ref), both EXACT?  Strictly smaller t wins.  On an
// exact tie the reference's slot order decides (bvh.h:95-106 visits slots in rank order with the
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
// Current best hit of a lane: the reference's t of primitive `ref` lies in [tf - ef, tf + ef]
// (ef == 0: tf IS the reference's t, which happens for every exactly-replayed triangle because
// Triangle::Hit rounds t to float, triangle.h:71).  hi >= tf + ef is the culling bound of the slab
// tests.  With nothing hit yet ref == kNoHit and tf = hi = the ray's tmax.  The exact double is
// deliberately NOT carried: the rare contested case replays both contenders.
struct Best {
  uint32_t ref;
  float tf, ef;
  float hi;
};

__device__ __forceinline__ void best_set_exact(Best& b, uint32_t ref, double t) {
  b.ref = ref;
  b.tf = (float)t;
  b.ef = (double)b.tf == t ? 0.0f : fabsf(b.tf) * 1.2e-7f;  // one ulp covers the rounding of a double t
  b.hi = __double2float_ru(t);
}

// Rare path of consider(): candidate (type, idx) cannot be ordered against the best by the FP32
// filters alone.  Replays the reference's own arithmetic for both and applies its tie rule.
__device__ __forceinline__ void consider_exact(const DevScene& s, float ox, float oy, float oz, float dx, float dy, float dz, float tmin,
                                            float tmax, uint32_t type, uint32_t idx, Best& b, HitCounters* cnt) {
#ifdef RTB_COUNTERS
  if (cnt) atomicAdd(&cnt->exact_evals, 1ull);
#else
  (void)cnt;
#endif
  double tb = (double)tmax;
  if (b.ref != kNoHit) {
    const uint32_t bty = RTB_REF_TYPE(b.ref);
    tb = exact_eval(exact_table(s, bty), bty, RTB_REF_INDEX(b.ref), ox, oy, oz, dx, dy, dz, tmin, tmax);
    if (!(tb == tb)) {  // a filter HIT the reference rejects would be a bug in the bounds; keep the interval midpoint
      tb = (double)b.tf;
#ifdef RTB_COUNTERS
      if (cnt) atomicAdd(&cnt->filter_violations, 1ull);
#endif
    }
    best_set_exact(b, b.ref, tb);
  }
  const double tc = exact_eval(exact_table(s, type), type, idx, ox, oy, oz, dx, dy, dz, tmin, tmax);
  if (!(tc == tc)) return;
  const uint32_t ref = RTB_REF(type, idx);
  if (replaces(s, tc, ref, tb, b.ref)) best_set_exact(b, ref, tc);
}

// Fold one primitive into the lane's best hit.
template <bool kRayInv = true>
__device__ __forceinline__ void consider(const DevScene& s, const TravRay& r, uint32_t type, uint32_t idx, Best& b, HitCounters* cnt) {
  float t = 0.f, e = 0.f;
  int cls;
  RTB_CHECK(kChkPrim, idx < (type == RTB_PRIM_TRIANGLE ? s.n_triangles : type == RTB_PRIM_SPHERE ? s.n_spheres : s.n_rects));
  if (type == RTB_PRIM_TRIANGLE) cls = triangle_filter(s, idx, r, b.hi, t, e);
  else if (type == RTB_PRIM_SPHERE) cls = sphere_filter(s, idx, r, b.hi, t, e);
  else cls = rect_filter<kRayInv>(s, idx, r, b.hi, t, e);
  if (cls == kMiss) return;
  if (cls == kHit) {
    if (t - e > b.hi) return;                                  // certainly farther than the best
    if (b.ref == kNoHit || t + e < b.tf - b.ef) {              // certainly nearer: take it, still inexact
      b.ref = RTB_REF(type, idx); b.tf = t; b.ef = e;
      b.hi = t + e;
      return;
    }
  }
  consider_exact(s, r.ox, r.oy, r.oz, r.dx, r.dy, r.dz, r.tmin, r.tmax, type, idx, b, cnt);
}

}  // namespace rtb
