// oracle/ref_harness.cc — TEST INFRASTRUCTURE, never linked into or called by the product.
//
// Thin C ABI around the UNMODIFIED reference sources, compiled in place from $REF_DIR
// (/root/reference) by oracle/Makefile into oracle/_ref/.  Everything numerical here is the
// reference's own code: geom::Bvh (SAH build + Hit), Aabb::Hit, Sphere/Triangle/xy|xz|yz_rect::Hit,
// scene::Camera::{Initialize,GetRay,GetPixel}, the material::* classes, ImageTexture/stb_image.
// The harness only adds
//   (1) scene construction from the product's export stream (so both sides see one scene),
//   (2) the id-recovering restatement of the Bvh::Hit loop (src/geom/bvh.h:84-116) over the
//       public accessors nodes()/prim_indices()/primitives() (bvh.h:134-136) — the reference's
//       HitRecord carries no primitive id (src/geom/hittable.h:18-30) — checked here against
//       Bvh::Hit itself by ora_intersect_world(),
//   (3) the MegaKernel/DefaultSampler pixel loop (src/renderer/mega_kernel.h:15-54,
//       src/integrator/sampler.h:26-34) with float outputs (mean and M2) instead of 8-bit PPM,
//       seeded per thread with core::SeedRng (src/core/random.h:46-48), and timed,
//   (4) ora_render_b: the reference's WavefrontRenderer::Render (src/renderer/wavefront.cc:40-242) run AS IS with
//       its CPURayIntegrator; std::cout / std::clog are captured, the P3 bytes and the per-pass "queue=" sizes
//       (= pixels not yet converged, wavefront.cc:62-82) are returned.
#include <omp.h>

#include <chrono>
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include <iostream>
#include <sstream>

#include "core/random.h"
#include "integrator/cpu_ray_integrator.h"
#include "integrator/pixel_state.h"
#include "renderer/wavefront.h"
#include "geom/bvh.h"
#include "geom/rect.h"
#include "geom/sphere.h"
#include "geom/triangle.h"
#include "material/material.h"
#include "material/texture.h"
#include "scene/camera.h"
#include "scene/scene.h"

using namespace rt;

namespace {

enum { kSphere = 0, kTriangle = 1, kRect = 2, kBvhBegin = 10, kBvhEnd = 11, kListBegin = 12, kListEnd = 13 };
constexpr int kRec = 16;

struct OraScene {
  scene::Scene world;
  std::vector<std::shared_ptr<geom::Bvh>> bvhs;
  std::unordered_map<const geom::Hittable*, int> gid;
  std::vector<std::shared_ptr<material::Texture>> texs;
  std::vector<std::shared_ptr<material::Material>> mats;
};

struct Reader {
  const double* prims; size_t n_prims;
  const double* mats; size_t n_mats;
  const double* texs; size_t n_texs;
  std::vector<std::string> images;
  OraScene* s;
  size_t pos = 0;
  int next_gid = 0;

  std::shared_ptr<material::Texture> Tex(int id) {
    if (s->texs[id]) return s->texs[id];
    const double* r = texs + 8 * size_t(id);
    switch (int(r[0])) {
      case 0: s->texs[id] = std::make_shared<material::SolidColor>(core::Color(r[4], r[5], r[6])); break;
      case 1: s->texs[id] = std::make_shared<material::CheckerTexture>(r[7], Tex(int(r[2])), Tex(int(r[3]))); break;
      default: s->texs[id] = std::make_shared<material::ImageTexture>(images.at(size_t(r[1])).c_str()); break;
    }
    return s->texs[id];
  }
  std::shared_ptr<material::Material> Mat(int id) {
    if (s->mats[id]) return s->mats[id];
    const double* r = mats + 8 * size_t(id);
    switch (int(r[0])) {
      case 0: s->mats[id] = std::make_shared<material::Lambertian>(Tex(int(r[1]))); break;
      case 1: s->mats[id] = std::make_shared<material::Metal>(core::Color(r[2], r[3], r[4]), r[5]); break;
      case 2: s->mats[id] = std::make_shared<material::Dielectric>(r[6]); break;
      default: s->mats[id] = std::make_shared<material::DiffuseLight>(Tex(int(r[1]))); break;
    }
    return s->mats[id];
  }
  void Keep(std::vector<std::shared_ptr<geom::Hittable>>& dst, std::shared_ptr<geom::Hittable> h) {
    s->gid[h.get()] = next_gid++;
    dst.push_back(std::move(h));
  }
  void ReadInto(std::vector<std::shared_ptr<geom::Hittable>>& dst, int end_kind) {
    while (pos < n_prims) {
      const double* r = prims + kRec * pos++;
      const int kind = int(r[0]);
      if (kind == end_kind) return;
      const int m = int(r[1]);
      if (kind == kSphere) {
        Keep(dst, std::make_shared<geom::Sphere>(core::Point3(r[2], r[3], r[4]), r[5], Mat(m)));
      } else if (kind == kTriangle) {
        Keep(dst, std::make_shared<geom::Triangle>(core::Point3(r[2], r[3], r[4]), core::Point3(r[5], r[6], r[7]),
                                                   core::Point3(r[8], r[9], r[10]), Mat(m)));
      } else if (kind == kRect) {
        const int axis = int(r[2]);
        if (axis == 2) Keep(dst, std::make_shared<geom::xy_rect>(r[4], r[5], r[6], r[7], r[3], Mat(m)));
        else if (axis == 1) Keep(dst, std::make_shared<geom::xz_rect>(r[4], r[5], r[6], r[7], r[3], Mat(m)));
        else Keep(dst, std::make_shared<geom::yz_rect>(r[4], r[5], r[6], r[7], r[3], Mat(m)));
      } else if (kind == kBvhBegin) {
        std::vector<std::shared_ptr<geom::Hittable>> inner;
        ReadInto(inner, kBvhEnd);
        auto b = std::make_shared<geom::Bvh>(inner);
        s->bvhs.push_back(b);
        dst.push_back(b);
      } else if (kind == kListBegin) {
        std::vector<std::shared_ptr<geom::Hittable>> inner;
        ReadInto(inner, kListEnd);
        auto list = std::make_shared<scene::Scene>();
        for (auto& o : inner) list->Add(o);
        dst.push_back(list);
      }
    }
  }
};

// Closest hit with the primitive recovered.  Statement-for-statement the traversal of
// Bvh::Hit (bvh.h:71-119) and Scene::Hit (scene.h:47-61); box and primitive tests are the
// reference's own member functions.
bool HitId(const OraScene& s, const geom::Hittable* obj, const core::Ray& r, core::Interval ray_t, geom::HitRecord& rec,
           int& id) {
  if (auto* bvh = dynamic_cast<const geom::Bvh*>(obj)) {
    const auto& nodes = bvh->nodes();
    if (nodes.empty()) return false;
    bool any = false;
    double closest = ray_t.max_;
    geom::HitRecord tmp;
    int tmp_id = -1;
    int stack[64];
    int sp = 0;
    stack[sp++] = 0;  // flatten() puts the root first (bvh.h:347-350)
    while (sp > 0) {
      const geom::BvhNodeGPU& node = nodes[stack[--sp]];
      if (!node.bbox.Hit(r, core::Interval(ray_t.min_, closest))) continue;
      if (node.isLeaf) {
        for (uint32_t i = 0; i < node.right_pCnt; ++i) {
          const auto& prim = bvh->primitives()[bvh->prim_indices()[node.left_pIdx + i]];
          if (HitId(s, prim.get(), r, core::Interval(ray_t.min_, closest), tmp, tmp_id)) {
            any = true;
            closest = tmp.t;
            rec = tmp;
            id = tmp_id;
          }
        }
      } else {
        stack[sp++] = int(node.right_pCnt);
        stack[sp++] = int(node.left_pIdx);
      }
    }
    return any;
  }
  if (auto* list = dynamic_cast<const scene::Scene*>(obj)) {
    bool any = false;
    double closest = ray_t.max_;
    geom::HitRecord tmp;
    int tmp_id = -1;
    for (const auto& o : list->objects_) {
      if (HitId(s, o.get(), r, core::Interval(ray_t.min_, closest), tmp, tmp_id)) {
        any = true;
        closest = tmp.t;
        rec = tmp;
        id = tmp_id;
      }
    }
    return any;
  }
  if (obj->Hit(r, ray_t, rec)) {
    id = s.gid.at(obj);
    return true;
  }
  return false;
}

core::Ray MakeRay(const float* f) {
  return core::Ray(core::Point3(f[0], f[1], f[2]), core::Vec3(f[4], f[5], f[6]));
}

thread_local uint64_t g_hit_calls = 0;
// Counts closest-hit queries: Camera::GetPixel passes `world` down its recursion (camera.h:158-164).
class CountingWorld : public geom::Hittable {
 public:
  explicit CountingWorld(const scene::Scene& w) : w_(w) {}
  bool Hit(const core::Ray& r, core::Interval t, geom::HitRecord& rec) const override {
    ++g_hit_calls;
    return w_.Hit(r, t, rec);
  }
  geom::Aabb BoundingBox() const override { return w_.BoundingBox(); }
  int TypeId() const override { return -1; }
  int ObjectIndex() const override { return -1; }
  void set_object_index(int) override {}
 private:
  const scene::Scene& w_;
};

}  // namespace

extern "C" {

void* ora_scene_create(const double* prims, size_t n_prims, const double* mats, size_t n_mats, const double* texs,
                       size_t n_texs, const char* image_paths) {
  auto s = std::make_unique<OraScene>();
  s->texs.resize(n_texs);
  s->mats.resize(n_mats);
  Reader rd{prims, n_prims, mats, n_mats, texs, n_texs, {}, s.get()};
  if (image_paths) {
    std::string cur;
    for (const char* p = image_paths; *p; ++p) {
      if (*p == '\n') { rd.images.push_back(cur); cur.clear(); } else cur.push_back(*p);
    }
    if (!cur.empty()) rd.images.push_back(cur);
  }
  if (n_prims == 0 || int(prims[0]) != kListBegin) return nullptr;
  rd.pos = 1;
  std::vector<std::shared_ptr<geom::Hittable>> top;
  rd.ReadInto(top, kListEnd);
  for (auto& o : top) s->world.Add(o);
  return s.release();
}

void ora_scene_free(void* h) { delete static_cast<OraScene*>(h); }

int ora_num_threads(void) { return omp_get_max_threads(); }

// rays: 8 floats each {ox,oy,oz,tmin,dx,dy,dz,tmax}, widened to double exactly.
void ora_intersect(void* h, const float* rays, size_t n, double* t_out, int32_t* gid_out, int threads) {
  const OraScene& s = *static_cast<OraScene*>(h);
  if (threads > 0) omp_set_num_threads(threads);
#pragma omp parallel for schedule(dynamic, 1024)
  for (size_t i = 0; i < n; ++i) {
    const float* f = rays + 8 * i;
    geom::HitRecord rec;
    int id = -1;
    const bool ok = HitId(s, &s.world, MakeRay(f), core::Interval(f[3], f[7]), rec, id);
    t_out[i] = ok ? rec.t : 0.0;
    gid_out[i] = ok ? id : -1;
  }
}

// The whole HitRecord the reference fills (hittable.h:18-34; sphere.h:44-51, triangle.h:78-84, rect.h:26-38) for the
// closest hit of each ray over (0.001f, +inf), the interval CPURayIntegrator::IntersectBatch uses
// (cpu_ray_integrator.h:21-22): out13 = {hit, t, p[3], normal[3], front_face, u, v, id, has_mat}.
void ora_intersect_records(void* h, const float* rays, size_t n, double* out13, int threads) {
  const OraScene& s = *static_cast<OraScene*>(h);
  if (threads > 0) omp_set_num_threads(threads);
#pragma omp parallel for schedule(dynamic, 1024)
  for (size_t i = 0; i < n; ++i) {
    const float* f = rays + 8 * i;
    geom::HitRecord rec;
    int id = -1;
    const bool ok = HitId(s, &s.world, MakeRay(f), core::Interval(0.001f, core::kInfinity), rec, id);
    double* o = out13 + 13 * i;
    for (int k = 0; k < 13; ++k) o[k] = 0.0;
    o[0] = ok;
    if (!ok) continue;
    o[1] = rec.t;
    for (int c = 0; c < 3; ++c) { o[2 + c] = rec.p[c]; o[5 + c] = rec.normal[c]; }
    o[8] = rec.front_face;
    o[9] = rec.u;
    o[10] = rec.v;
    o[11] = id;
    o[12] = rec.mat != nullptr;
  }
}

// The reference's own entry (scene::Scene::Hit -> Bvh::Hit), no ids: validates ora_intersect.
// Also what CPURayIntegrator::IntersectBatch does per ray (cpu_ray_integrator.h:25-44).
double ora_intersect_world(void* h, const float* rays, size_t n, double* t_out, uint8_t* hit_out, int threads) {
  const OraScene& s = *static_cast<OraScene*>(h);
  if (threads > 0) omp_set_num_threads(threads);
  const auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(dynamic, 1024)
  for (size_t i = 0; i < n; ++i) {
    const float* f = rays + 8 * i;
    geom::HitRecord rec;
    const bool ok = s.world.Hit(MakeRay(f), core::Interval(f[3], f[7]), rec);
    if (t_out) t_out[i] = ok ? rec.t : 0.0;
    if (hit_out) hit_out[i] = ok;
  }
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

int ora_bvh_count(void* h) { return int(static_cast<OraScene*>(h)->bvhs.size()); }
void ora_bvh_size(void* h, int k, size_t* n_nodes, size_t* n_prims) {
  const auto& b = *static_cast<OraScene*>(h)->bvhs[k];
  *n_nodes = b.nodes().size();
  *n_prims = b.prim_indices().size();
}
void ora_bvh_dump(void* h, int k, double* nodes9, int32_t* prim_indices) {
  const auto& b = *static_cast<OraScene*>(h)->bvhs[k];
  for (size_t i = 0; i < b.nodes().size(); ++i) {
    const auto& n = b.nodes()[i];
    double* o = nodes9 + 9 * i;
    o[0] = n.bbox.x.min_; o[1] = n.bbox.x.max_; o[2] = n.bbox.y.min_; o[3] = n.bbox.y.max_; o[4] = n.bbox.z.min_; o[5] = n.bbox.z.max_;
    o[6] = n.left_pIdx; o[7] = n.right_pCnt; o[8] = n.isLeaf;
  }
  for (size_t i = 0; i < b.prim_indices().size(); ++i) prim_indices[i] = b.prim_indices()[i];
}

// cfg: aspect, width, spp, max_depth, vfov, lookfrom[3], lookat[3], vup[3], defocus_angle, focus_dist (16 doubles)
static void SetupCamera(scene::Camera& cam, const double* cfg) {
  scene::CameraConfig c;
  c.aspect_ratio = cfg[0];
  c.image_width = int(cfg[1]);
  c.samples_per_pixel = int(cfg[2]);
  c.max_depth = int(cfg[3]);
  c.vfov = cfg[4];
  c.lookfrom = core::Vec3(cfg[5], cfg[6], cfg[7]);
  c.lookat = core::Vec3(cfg[8], cfg[9], cfg[10]);
  c.vup = core::Vec3(cfg[11], cfg[12], cfg[13]);
  c.defocus_angle = cfg[14];
  c.focus_dist = cfg[15];
  cam.SetFromConfig(c);
  cam.Initialize();
}

void ora_camera_dims(const double* cfg, int* w, int* h) {
  scene::ColorCamera cam;
  SetupCamera(cam, cfg);
  *w = cam.get_image_width();
  *h = cam.get_image_height();
}

// Camera::Initialize's derived frame, for checking the host mirror: 6 vectors x 3 doubles are
// not reachable (private), so expose what GetRay reveals: primary rays for given (i,j) with the
// RNG forced to a known state.  n pixel pairs in ij -> 6 doubles each (origin, direction).
void ora_camera_rays(const double* cfg, const int32_t* ij, size_t n, unsigned seed, double* out6) {
  scene::ColorCamera cam;
  SetupCamera(cam, cfg);
  core::SeedRng(seed);
  for (size_t k = 0; k < n; ++k) {
    const core::Ray r = cam.GetRay(ij[2 * k], ij[2 * k + 1]);
    for (int c = 0; c < 3; ++c) { out6[6 * k + c] = r.origin()[c]; out6[6 * k + 3 + c] = r.direction()[c]; }
  }
}

// Path A (MegaKernel + DefaultSampler + Camera::GetPixel).  mean/m2: W*H*3 doubles (either may
// be null).  y range [y0,y1) lets a bounded sample of the image be timed.  Returns seconds spent
// in the pixel loop; *rays_out = closest-hit queries issued.
double ora_render_a(void* h, const double* cfg, int spp, int max_depth, unsigned seed, int threads, int y0, int y1,
                    double* mean, double* m2, uint64_t* rays_out, int depth_camera) {
  const OraScene& s = *static_cast<OraScene*>(h);
  scene::ColorCamera color_cam;
  scene::DepthCamera depth_cam;
  scene::Camera& cam = depth_camera ? static_cast<scene::Camera&>(depth_cam) : static_cast<scene::Camera&>(color_cam);
  SetupCamera(cam, cfg);
  const int width = cam.get_image_width(), height = cam.get_image_height();
  if (y1 < 0 || y1 > height) y1 = height;
  if (y0 < 0) y0 = 0;
  if (threads > 0) omp_set_num_threads(threads);
  CountingWorld world(s.world);
  uint64_t rays = 0;
  const auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel reduction(+ : rays)
  {
    core::SeedRng(seed * 7919u + 104729u * unsigned(omp_get_thread_num()) + 1u);
    g_hit_calls = 0;
#pragma omp for schedule(dynamic)
    for (int y = y0; y < y1; y++) {
      for (int x = 0; x < width; x++) {
        core::Color sum(0, 0, 0), mu(0, 0, 0), acc2(0, 0, 0);
        for (int k = 0; k < spp; k++) {
          const core::Ray r = cam.GetRay(x, y);
          const core::Color c = cam.GetPixel(r, max_depth, world);
          sum += c;
          for (int ch = 0; ch < 3; ++ch) {  // Welford, as pixel_state.h:22-39
            const double d = c[ch] - mu[ch];
            mu[ch] += d / (k + 1);
            acc2[ch] += d * (c[ch] - mu[ch]);
          }
        }
        const size_t o = (size_t(y) * width + x) * 3;
        for (int ch = 0; ch < 3; ++ch) {
          if (mean) mean[o + ch] = sum[ch] / spp;
          if (m2) m2[o + ch] = acc2[ch];
        }
      }
    }
    rays += g_hit_calls;
  }
  const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  if (rays_out) *rays_out = rays;
  return secs;
}

// Path B exactly as the reference's main() runs it (main.cc:188-194): WavefrontRenderer over CPURayIntegrator.
// rgb8: W*H*3 bytes parsed from the P3 text it writes; queue_sizes[s] = rays generated in pass s (one per
// pixel that has not converged), up to n_queue entries.  Returns seconds, < 0 on a parse failure.
double ora_render_b(void* h, const double* cfg, int max_depth, int max_samples, unsigned seed, int threads, uint8_t* rgb8,
                    int32_t* queue_sizes, int n_queue) {
  const OraScene& s = *static_cast<OraScene*>(h);
  scene::ColorCamera cam;
  SetupCamera(cam, cfg);
  if (threads > 0) omp_set_num_threads(threads);
#pragma omp parallel
  core::SeedRng(seed * 7919u + 104729u * unsigned(omp_get_thread_num()) + 1u);
  integrator::CPURayIntegrator integ(&s.world);
  renderer::WavefrontRenderer wf(s.world, cam, integ, max_depth, max_samples, 8192);
  std::ostringstream out, log;
  std::streambuf* old_out = std::cout.rdbuf(out.rdbuf());
  std::streambuf* old_log = std::clog.rdbuf(log.rdbuf());
  const auto t0 = std::chrono::steady_clock::now();
  wf.Render();
  const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  std::cout.rdbuf(old_out);
  std::clog.rdbuf(old_log);
  std::istringstream in(out.str());
  std::string magic;
  int w = 0, hh = 0, maxv = 0;
  in >> magic >> w >> hh >> maxv;
  if (magic != "P3" || w != cam.get_image_width() || hh != cam.get_image_height()) return -1.0;
  for (size_t i = 0; i < size_t(w) * hh * 3; ++i) {
    int v = -1;
    in >> v;
    if (v < 0 || v > 255) return -2.0;
    rgb8[i] = uint8_t(v);
  }
  const std::string text = log.str();
  size_t pos = 0;
  for (int k = 0; k < n_queue; ++k) {
    pos = text.find("queue=", pos);
    if (pos == std::string::npos) { queue_sizes[k] = -1; continue; }
    pos += 6;
    queue_sizes[k] = int32_t(std::strtol(text.c_str() + pos, nullptr, 10));
  }
  return secs;
}

// integrator::RecordSample + IsConverged (src/integrator/pixel_state.h:22-72) driven the way WavefrontRenderer::Render
// drives them (wavefront.cc:120-127: record, then mark converged; a converged pixel gets no further samples,
// wavefront.cc:70-71).  samples: n_seq x len x 3 doubles.  stop_n[i] = PixelState::samples when the pixel stopped
// (len if it never converged); sum / m2: the PixelState fields at that point.
void ora_pixel_state_run(const double* samples, size_t n_seq, int len, double rel_threshold, int min_spp, int32_t* stop_n, double* sum,
                         double* m2) {
  for (size_t i = 0; i < n_seq; ++i) {
    integrator::PixelState ps;
    for (int k = 0; k < len && !ps.converged; ++k) {
      const double* c = samples + (i * size_t(len) + size_t(k)) * 3;
      integrator::RecordSample(ps, core::Color(c[0], c[1], c[2]));
      if (!ps.converged && integrator::IsConverged(ps, rel_threshold, min_spp)) ps.converged = true;
    }
    stop_n[i] = ps.samples;
    for (int ch = 0; ch < 3; ++ch) { sum[3 * i + ch] = ps.sum[ch]; m2[3 * i + ch] = ps.m2[ch]; }
  }
}

// Texture::Value and Material::Emitted probes (ImageTexture pipeline check).
void ora_texture_value(void* h, int tex_id, const double* uvp, size_t n, double* rgb) {
  const OraScene& s = *static_cast<OraScene*>(h);
  for (size_t i = 0; i < n; ++i) {
    const double* q = uvp + 5 * i;
    const core::Color c = s.texs[size_t(tex_id)]->Value(q[0], q[1], core::Point3(q[2], q[3], q[4]));
    rgb[3 * i] = c[0]; rgb[3 * i + 1] = c[1]; rgb[3 * i + 2] = c[2];
  }
}

}  // extern "C"
