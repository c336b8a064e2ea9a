// rtb_shade.cuh — device versions of the reference's camera, material and texture code.
// Every function cites the reference lines it restates; arithmetic is FP32 (the reference is
// double), which is inside the Monte-Carlo tolerance the image-parity tests state.
#pragma once
#include "rtb_device.cuh"

namespace rtb {

struct f3 { float x, y, z; };
__device__ __forceinline__ f3 mk3(float x, float y, float z) { return {x, y, z}; }
__device__ __forceinline__ f3 operator+(f3 a, f3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ f3 operator-(f3 a, f3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ f3 operator*(f3 a, f3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
__device__ __forceinline__ f3 operator*(float s, f3 a) { return {s * a.x, s * a.y, s * a.z}; }
__device__ __forceinline__ f3 operator-(f3 a) { return {-a.x, -a.y, -a.z}; }
__device__ __forceinline__ float dot3(f3 a, f3 b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z)); }
__device__ __forceinline__ f3 cross3(f3 a, f3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
// core::Normalize, src/core/math_utils.h:93-97 (zero vector stays zero)
__device__ __forceinline__ f3 normalize3(f3 v) {
  const float l2 = dot3(v, v);
  return l2 == 0.0f ? mk3(0, 0, 0) : rsqrtf(l2) * v;
}

// Uniform direction on the unit sphere.  Same distribution as core::RandomUnitVector's rejection
// loop (math_utils.h:62-70), sampled directly (z uniform, phi uniform) so the warp never diverges.
// Shading arithmetic is Monte-Carlo sampling in FP32: the special-function units' accuracy (MUFU.RSQ / SIN / COS,
// ~1e-6 absolute on [-pi, pi]) is far inside the estimator's noise, and saves the ~30 instructions per bounce that the
// correctly rounded sqrtf / sincospif / division sequences cost (RTB_FAST_SHADE_MATH=0 restores them).
#ifndef RTB_FAST_SHADE_MATH
#define RTB_FAST_SHADE_MATH 1
#endif
__device__ __forceinline__ float shade_sqrt(float x) {  // x >= 0
#if RTB_FAST_SHADE_MATH
  const float t = fmaxf(x, 1e-30f);
  return t * rsqrtf(t);
#else
  return sqrtf(x);
#endif
}
__device__ __forceinline__ float shade_rcp(float x) {
#if RTB_FAST_SHADE_MATH
  return __fdividef(1.0f, x);
#else
  return 1.0f / x;
#endif
}
// sine and cosine of 2 pi u, u in [0, 1)
__device__ __forceinline__ void shade_sincos2pi(float u, float* s, float* c) {
#if RTB_FAST_SHADE_MATH
  __sincosf(fmaf(6.283185307179586f, u, -3.14159265358979f), s, c);  // the same uniform angle, shifted into MUFU's accurate range
#else
  sincospif(2.0f * u, s, c);
#endif
}
__device__ __forceinline__ f3 unit_vector(float u1, float u2) {
  const float z = 1.0f - 2.0f * u1;
  const float rxy = shade_sqrt(fmaxf(0.0f, 1.0f - z * z));
  float s, c;
  shade_sincos2pi(u2, &s, &c);
  return {rxy * c, rxy * s, z};
}

// core::Reflect / core::Refract, math_utils.h:14-29
__device__ __forceinline__ f3 reflect3(f3 v, f3 n) { return v - (2.0f * dot3(v, n)) * n; }
__device__ __forceinline__ f3 refract3(f3 uv, f3 n, float eta) {
  const float cos_theta = fminf(dot3(-uv, n), 1.0f);
  const f3 perp = eta * (uv + cos_theta * n);
  const f3 par = (-shade_sqrt(fabsf(1.0f - dot3(perp, perp)))) * n;
  return perp + par;
}

// Camera::GetPixel miss branch (src/scene/camera.h:171-173) == WavefrontRenderer::background
// (src/renderer/wavefront.cc:33-38)
__device__ __forceinline__ f3 sky(f3 dir) {
  const f3 u = normalize3(dir);
  const float a = 0.5f * (u.y + 1.0f);
  return (1.0f - a) * mk3(1.0f, 1.0f, 1.0f) + a * mk3(0.5f, 0.7f, 1.0f);
}

// Sphere::get_sphere_uv, src/geom/sphere.h:73-79
__device__ __forceinline__ void sphere_uv(f3 n, float& u, float& v) {
  const float kPi = 3.14159265358979323846f;
  const float theta = acosf(fminf(fmaxf(-n.y, -1.0f), 1.0f));
  const float phi = atan2f(-n.z, n.x) + kPi;
  u = phi / (2.0f * kPi);
  v = theta / kPi;
}

// Texture::Value: SolidColor (texture.h:24-26), CheckerTexture (texture.h:39-47, nesting followed
// iteratively), ImageTexture (texture.h:59-73 + Image::PixelData clamp image.cc:50-67).
// (u, v) of a sphere are only needed by image textures: they are computed here, on demand, from the outward
// normal (sphere_n, valid when on_sphere) instead of for every sphere hit as the reference does (sphere.h:50).
__device__ __forceinline__ f3 texture_value(const DevScene& s, int tex, float u, float v, f3 p, bool on_sphere, f3 sphere_n) {
  for (int guard = 0; guard < 8; ++guard) {
    RTB_CHECK(kChkMaterial, tex >= 0 && (uint32_t)tex < s.n_textures);
    const RtbTexture t = s.textures[tex];
    if (t.type == RTB_TEX_SOLID) return mk3(t.color[0], t.color[1], t.color[2]);
    if (t.type == RTB_TEX_CHECKER) {
      const int xi = (int)floorf(t.inv_scale * p.x), yi = (int)floorf(t.inv_scale * p.y), zi = (int)floorf(t.inv_scale * p.z);
      tex = ((xi + yi + zi) % 2 == 0) ? t.even : t.odd;
      continue;
    }
    // image
    const int2 dim = s.image_dim[t.image];
    if (dim.y <= 0) return mk3(0.0f, 1.0f, 1.0f);  // texture.h:61
    if (on_sphere) sphere_uv(sphere_n, u, v);
    u = fminf(fmaxf(u, 0.0f), 1.0f);
    v = 1.0f - fminf(fmaxf(v, 0.0f), 1.0f);
    int i = (int)(u * dim.x), j = (int)(v * dim.y);
    i = i < dim.x ? i : dim.x - 1;
    j = j < dim.y ? j : dim.y - 1;
    const uchar4 px = tex2D<uchar4>(s.image_tex[t.image], i + 0.5f, j + 0.5f);
    const float k = 1.0f / 255.0f;
    return mk3(k * px.x, k * px.y, k * px.z);
  }
  return mk3(0, 0, 0);
}

// Surface frame at a hit: outward normal, (u,v), material.  Sphere: sphere.h:47-51;
// Triangle: triangle.h:82-84 (geometric normal; u,v are not set by the reference -> 0);
// rects: rect.h:30-38.
__device__ __forceinline__ void surface_at(const DevScene& s, uint32_t ref, f3 p, f3& n_out, float& u, float& v, uint32_t& mat) {
  const uint32_t ty = RTB_REF_TYPE(ref), idx = RTB_REF_INDEX(ref);
  u = 0.0f;
  v = 0.0f;
  if (ty == RTB_PRIM_SPHERE) {
    const float4 c = __ldg(&s.sphere_f[idx]);
    const float inv_r = shade_rcp(c.w);
    n_out = inv_r * mk3(p.x - c.x, p.y - c.y, p.z - c.z);
    mat = __ldg(&s.sphere_mat[idx]);
  } else if (ty == RTB_PRIM_TRIANGLE) {
    const float4 a = __ldg(&s.triangle_f[3 * (size_t)idx]), e1 = __ldg(&s.triangle_f[3 * (size_t)idx + 1]), e2 = __ldg(&s.triangle_f[3 * (size_t)idx + 2]);
    n_out = normalize3(cross3(mk3(e1.x, e1.y, e1.z), mk3(e2.x, e2.y, e2.z)));
    mat = __float_as_uint(a.w);
  } else {
    const float4 q0 = __ldg(&s.rect_f[2 * (size_t)idx]), q1 = __ldg(&s.rect_f[2 * (size_t)idx + 1]);  // (k,a0,a1,b0) (b1,axis,mat,-)
    const int axis = __float_as_int(q1.y);
    mat = __float_as_uint(q1.z);
    n_out = mk3(axis == 0 ? 1.0f : 0.0f, axis == 1 ? 1.0f : 0.0f, axis == 2 ? 1.0f : 0.0f);
    const float pa = axis == 0 ? p.y : p.x, pb = axis == 2 ? p.y : p.z;
    u = (pa - q0.y) * shade_rcp(q0.z - q0.y);
    v = (pb - q0.w) * shade_rcp(q1.x - q0.w);
  }
}

// Dielectric::Reflectance (Schlick), src/material/material.cc:258-262
__device__ __forceinline__ float schlick(float cosine, float ref_idx) {
  float r0 = (1.0f - ref_idx) / (1.0f + ref_idx);
  r0 = r0 * r0;
  const float m = 1.0f - cosine;
  const float m2 = m * m;
  return r0 + (1.0f - r0) * (m2 * m2 * m);
}

}  // namespace rtb
