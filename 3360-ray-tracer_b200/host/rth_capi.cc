// rth_* — C wrappers over the host C++ scene API so that the Python tests / bench can build the
// named scenes, flatten them to an RtbSceneDesc and export them for the oracle.  Declared in
// include/rth.h.  Pure host code: no CUDA in this translation unit, and librth.so does not link librtb200.so — the
// device library's tree compiler reaches it through rth_set_tree_compiler (so a process that only needs the host scene
// API, e.g. bench.py --impl reference feeding the oracle, never maps the CUDA library).
#include <rth.h>

#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>

#include "rt/b200/flatten.h"
#include "rt/b200/scenes.h"
#include "rt/core/png.h"
#include "rt/scene/camera.h"

using namespace rt;

namespace {
RthTreeCompileFn g_compile = nullptr;
RthTreeDestroyFn g_destroy = nullptr;
RthLastErrorFn g_compile_error = nullptr;
}  // namespace

struct RthScene {
  b200::BuiltScene built;
  b200::FlatScene flat;
  RtbSceneDesc desc;
  RtbCompiledTree* tree = nullptr;  // device-native traversal tree, built once with the scene (host work)
  std::string image_paths;  // '\n'-joined
  RthTreeDestroyFn destroy = nullptr;  // the destroyer that belongs to the compiler which made `tree`
  ~RthScene() {
    if (tree && destroy) destroy(tree);
  }
};

namespace {
thread_local std::string g_error;
int Fail(const std::exception& e) {
  g_error = e.what();
  return RTB_EINVAL;
}
double Param(const double* p, int n, int i, double dflt) { return (p && i < n) ? p[i] : dflt; }

void Finish(RthScene& s, uint32_t top_budget, uint32_t flags = 0) {
  if (flags & RTH_BUILD_DEVICE_TREE) {
    s.flat = b200::Flattener(top_budget).RunSoup(s.built.world);
    s.desc = s.flat.desc();
  } else {
    s.flat = b200::Flattener(top_budget).Run(s.built.world);
    s.desc = s.flat.desc();
    if (g_compile) {
      if (g_compile(&s.desc, &s.tree) != RTB_OK) throw std::runtime_error(g_compile_error ? g_compile_error() : "tree compile failed");
      s.destroy = g_destroy;
      s.desc.compiled = s.tree;
    }
  }
  s.image_paths.clear();
  for (const auto& p : s.flat.image_paths) {
    if (!s.image_paths.empty()) s.image_paths += '\n';
    s.image_paths += p;
  }
}

// Re-create host objects from an export stream (the inverse of Flattener::Export).
struct RecordReader {
  const double* prims; size_t n_prims;
  const double* mats; size_t n_mats;
  const double* texs; size_t n_texs;
  std::vector<std::string> images;
  std::vector<std::shared_ptr<material::Texture>> tex_objs;
  std::vector<std::shared_ptr<material::Material>> mat_objs;
  b200::BuiltScene* out;
  size_t pos = 0;

  std::shared_ptr<material::Texture> Tex(int id) {
    if (id < 0 || static_cast<size_t>(id) >= n_texs) throw std::runtime_error("records: texture id out of range");
    if (tex_objs[id]) return tex_objs[id];
    const double* r = texs + 8 * static_cast<size_t>(id);
    switch (static_cast<int>(r[0])) {
      case RTB_TEX_SOLID: tex_objs[id] = std::make_shared<material::SolidColor>(core::Color(r[4], r[5], r[6])); break;
      case RTB_TEX_CHECKER: tex_objs[id] = std::make_shared<material::CheckerTexture>(r[7], Tex(int(r[2])), Tex(int(r[3]))); break;
      case RTB_TEX_IMAGE: {
        const int img = static_cast<int>(r[1]);
        if (img < 0 || static_cast<size_t>(img) >= images.size()) throw std::runtime_error("records: image id out of range");
        tex_objs[id] = std::make_shared<material::ImageTexture>(images[img].c_str());
        break;
      }
      default: throw std::runtime_error("records: unknown texture type");
    }
    return tex_objs[id];
  }
  std::shared_ptr<material::Material> Mat(int id) {
    if (id < 0 || static_cast<size_t>(id) >= n_mats) throw std::runtime_error("records: material id out of range");
    if (mat_objs[id]) return mat_objs[id];
    const double* r = mats + 8 * static_cast<size_t>(id);
    switch (static_cast<int>(r[0])) {
      case RTB_MAT_LAMBERTIAN: mat_objs[id] = std::make_shared<material::Lambertian>(Tex(int(r[1]))); break;
      case RTB_MAT_METAL: mat_objs[id] = std::make_shared<material::Metal>(core::Color(r[2], r[3], r[4]), r[5]); break;
      case RTB_MAT_DIELECTRIC: mat_objs[id] = std::make_shared<material::Dielectric>(r[6]); break;
      case RTB_MAT_DIFFUSE_LIGHT: mat_objs[id] = std::make_shared<material::DiffuseLight>(Tex(int(r[1]))); break;
      default: throw std::runtime_error("records: unknown material type");
    }
    return mat_objs[id];
  }
  // Reads objects until the matching end marker (or the end of the stream at top level).
  void ReadInto(std::vector<std::shared_ptr<geom::Hittable>>& dst, int end_kind) {
    while (pos < n_prims) {
      const double* r = prims + b200::kExportRecord * pos++;
      const int kind = static_cast<int>(r[0]);
      if (kind == end_kind) return;
      const int m = static_cast<int>(r[1]);
      switch (kind) {
        case b200::kExSphere:
          dst.push_back(std::make_shared<geom::Sphere>(core::Point3(r[2], r[3], r[4]), r[5], Mat(m)));
          break;
        case b200::kExTriangle:
          dst.push_back(std::make_shared<geom::Triangle>(core::Point3(r[2], r[3], r[4]), core::Point3(r[5], r[6], r[7]),
                                                         core::Point3(r[8], r[9], r[10]), Mat(m)));
          break;
        case b200::kExRect: {
          const int axis = static_cast<int>(r[2]);
          if (axis == 2) dst.push_back(std::make_shared<geom::xy_rect>(r[4], r[5], r[6], r[7], r[3], Mat(m)));
          else if (axis == 1) dst.push_back(std::make_shared<geom::xz_rect>(r[4], r[5], r[6], r[7], r[3], Mat(m)));
          else if (axis == 0) dst.push_back(std::make_shared<geom::yz_rect>(r[4], r[5], r[6], r[7], r[3], Mat(m)));
          else throw std::runtime_error("records: bad rect axis");
          break;
        }
        case b200::kExBvhBegin: {
          std::vector<std::shared_ptr<geom::Hittable>> inner;
          ReadInto(inner, b200::kExBvhEnd);
          dst.push_back(out->AddBvh(inner));
          break;
        }
        case b200::kExListBegin: {
          auto list = std::make_shared<scene::Scene>();
          std::vector<std::shared_ptr<geom::Hittable>> inner;
          ReadInto(inner, b200::kExListEnd);
          for (auto& o : inner) list->Add(o);
          dst.push_back(list);
          break;
        }
        default: throw std::runtime_error("records: unexpected record kind");
      }
    }
    if (end_kind >= 0) throw std::runtime_error("records: unterminated group");
  }
};
}  // namespace

extern "C" {

const char* rth_last_error(void) { return g_error.c_str(); }

void rth_set_tree_compiler(RthTreeCompileFn compile, RthTreeDestroyFn destroy, RthLastErrorFn last_error) {
  g_compile = compile;
  g_destroy = destroy;
  g_compile_error = last_error;
}

int rth_scene_build(const char* name, const double* params, int n_params, const char* asset_dir, uint32_t top_budget,
                    RthScene** out) {
  return rth_scene_build_ex(name, params, n_params, asset_dir, top_budget, 0u, out);
}

int rth_scene_build_ex(const char* name, const double* params, int n_params, const char* asset_dir, uint32_t top_budget,
                       uint32_t flags, RthScene** out) {
  if (!name || !out) return RTB_EINVAL;
  struct DeferGuard {  // host geom::Bvh constructors skip their SAH build while the device will build the tree
    bool old;
    explicit DeferGuard(bool on) : old(geom::Bvh::DeferBuild()) { geom::Bvh::DeferBuild() = on; }
    ~DeferGuard() { geom::Bvh::DeferBuild() = old; }
  } guard((flags & RTH_BUILD_DEVICE_TREE) != 0);
  try {
    auto s = std::make_unique<RthScene>();
    const std::string n(name), assets(asset_dir ? asset_dir : "assets");
    const std::string earth_jpg = assets + "/textures/earthmap.jpg", bunny_obj = assets + "/models/stanford-bunny.obj";
    if (n == "cornell") {
      b200::cornell_box(s->built);
    } else if (n == "rtiow") {  // params: extent, seed, with_earth, intro
      b200::rtiow_spheres(s->built, int(Param(params, n_params, 0, 11)), unsigned(Param(params, n_params, 1, 1)),
                          Param(params, n_params, 2, 0) != 0, Param(params, n_params, 3, 0) != 0, earth_jpg);
    } else if (n == "checkered_spheres") {
      b200::checkered_spheres(s->built);
    } else if (n == "earth") {
      b200::earth(s->built, earth_jpg);
    } else if (n == "earth_bubbles") {
      b200::earth_bubbles(s->built, earth_jpg);
    } else if (n == "bunny") {  // params: scale
      b200::bunny(s->built, bunny_obj, Param(params, n_params, 0, 50));
    } else if (n == "stress") {  // params: nx, nz, ns, half, seed
      b200::stress(s->built, int(Param(params, n_params, 0, 1415)), int(Param(params, n_params, 1, 1414)),
                   int(Param(params, n_params, 2, 1000)), Param(params, n_params, 3, 500), unsigned(Param(params, n_params, 4, 2)));
    } else {
      throw std::runtime_error("unknown scene '" + n + "'");
    }
    Finish(*s, top_budget ? top_budget : 1024, flags);
    *out = s.release();
    return RTB_OK;
  } catch (const std::exception& e) {
    return Fail(e);
  }
}

int rth_scene_from_records(const double* prims, size_t n_prims, const double* mats, size_t n_mats, const double* texs,
                           size_t n_texs, const char* image_paths, uint32_t top_budget, RthScene** out) {
  if (!prims || !out) return RTB_EINVAL;
  try {
    auto s = std::make_unique<RthScene>();
    RecordReader rd{prims, n_prims, mats, n_mats, texs, n_texs, {}, {}, {}, &s->built};
    rd.tex_objs.resize(n_texs);
    rd.mat_objs.resize(n_mats);
    if (image_paths) {
      std::string all(image_paths), cur;
      for (char c : all) {
        if (c == '\n') { rd.images.push_back(cur); cur.clear(); } else cur.push_back(c);
      }
      if (!cur.empty()) rd.images.push_back(cur);
    }
    // the outermost list record is the world itself
    if (n_prims == 0 || static_cast<int>(prims[0]) != b200::kExListBegin) throw std::runtime_error("records: stream must start with a list");
    rd.pos = 1;
    std::vector<std::shared_ptr<geom::Hittable>> top;
    rd.ReadInto(top, b200::kExListEnd);
    for (auto& o : top) s->built.world.Add(o);
    Finish(*s, top_budget ? top_budget : 1024);
    *out = s.release();
    return RTB_OK;
  } catch (const std::exception& e) {
    return Fail(e);
  }
}

const RtbSceneDesc* rth_scene_desc(const RthScene* s) { return s ? &s->desc : nullptr; }

int rth_scene_export(const RthScene* s, const double** prims, size_t* n_prims, const double** mats, size_t* n_mats,
                     const double** texs, size_t* n_texs, const char** image_paths) {
  if (!s) return RTB_EINVAL;
  if (prims) *prims = s->flat.prim_records.data();
  if (n_prims) *n_prims = s->flat.prim_records.size() / b200::kExportRecord;
  if (mats) *mats = s->flat.material_records.data();
  if (n_mats) *n_mats = s->flat.material_records.size() / 8;
  if (texs) *texs = s->flat.texture_records.data();
  if (n_texs) *n_texs = s->flat.texture_records.size() / 8;
  if (image_paths) *image_paths = s->image_paths.c_str();
  return RTB_OK;
}

int rth_scene_gid_map(const RthScene* s, const uint32_t** gid_to_ref, size_t* n) {
  if (!s) return RTB_EINVAL;
  if (gid_to_ref) *gid_to_ref = s->flat.gid_to_ref.data();
  if (n) *n = s->flat.gid_to_ref.size();
  return RTB_OK;
}

int rth_scene_bvh_count(const RthScene* s) { return s ? static_cast<int>(s->built.bvhs.size()) : 0; }

int rth_scene_bvh_size(const RthScene* s, int k, size_t* n_nodes, size_t* n_prims) {
  if (!s || k < 0 || k >= static_cast<int>(s->built.bvhs.size())) return RTB_EINVAL;
  if (n_nodes) *n_nodes = s->built.bvhs[k]->nodes().size();
  if (n_prims) *n_prims = s->built.bvhs[k]->prim_indices().size();
  return RTB_OK;
}

int rth_scene_bvh_dump(const RthScene* s, int k, double* nodes9, int32_t* prim_indices) {
  if (!s || k < 0 || k >= static_cast<int>(s->built.bvhs.size())) return RTB_EINVAL;
  const geom::Bvh& b = *s->built.bvhs[k];
  if (nodes9) {
    for (size_t i = 0; i < b.nodes().size(); ++i) {
      const geom::BvhNodeGPU& n = b.nodes()[i];
      double* o = nodes9 + 9 * i;
      o[0] = n.bbox.x.min_; o[1] = n.bbox.x.max_; o[2] = n.bbox.y.min_; o[3] = n.bbox.y.max_; o[4] = n.bbox.z.min_; o[5] = n.bbox.z.max_;
      o[6] = n.left_pIdx; o[7] = n.right_pCnt; o[8] = n.isLeaf;
    }
  }
  if (prim_indices)
    for (size_t i = 0; i < b.prim_indices().size(); ++i) prim_indices[i] = b.prim_indices()[i];
  return RTB_OK;
}

void rth_scene_free(RthScene* s) { delete s; }

int rth_camera_init(const RthCameraConfig* cfg, int depth_camera, RtbCamera* out) {
  if (!cfg || !out) return RTB_EINVAL;
  scene::CameraConfig c;
  c.aspect_ratio = cfg->aspect_ratio;
  c.image_width = cfg->image_width;
  c.samples_per_pixel = cfg->samples_per_pixel;
  c.max_depth = cfg->max_depth;
  c.vfov = cfg->vfov;
  c.lookfrom = core::Vec3(cfg->lookfrom[0], cfg->lookfrom[1], cfg->lookfrom[2]);
  c.lookat = core::Vec3(cfg->lookat[0], cfg->lookat[1], cfg->lookat[2]);
  c.vup = core::Vec3(cfg->vup[0], cfg->vup[1], cfg->vup[2]);
  c.defocus_angle = cfg->defocus_angle;
  c.focus_dist = cfg->focus_dist;
  scene::ColorCamera color;
  scene::DepthCamera depth;
  scene::Camera& cam = depth_camera ? static_cast<scene::Camera&>(depth) : static_cast<scene::Camera&>(color);
  cam.SetFromConfig(c);
  cam.Initialize();
  *out = cam.ToDevice();
  return RTB_OK;
}

int rth_camera_preset(const char* json_path, const char* name, RthCameraConfig* out) {
  if (!json_path || !name || !out) return RTB_EINVAL;
  try {
    auto cams = scene::loadCameras(json_path);
    auto it = cams.find(name);
    if (it == cams.end()) throw std::runtime_error(std::string("camera '") + name + "' not found");
    const scene::CameraConfig& c = it->second;
    out->aspect_ratio = c.aspect_ratio;
    out->image_width = c.image_width;
    out->samples_per_pixel = c.samples_per_pixel;
    out->max_depth = c.max_depth;
    out->vfov = c.vfov;
    for (int i = 0; i < 3; ++i) { out->lookfrom[i] = c.lookfrom[i]; out->lookat[i] = c.lookat[i]; out->vup[i] = c.vup[i]; }
    out->defocus_angle = c.defocus_angle;
    out->focus_dist = c.focus_dist;
    return RTB_OK;
  } catch (const std::exception& e) {
    return Fail(e);
  }
}

int rth_encode_png(const uint8_t* rgb8, int width, int height, uint8_t* out, size_t capacity, size_t* n_bytes) {
  if (!rgb8 || !n_bytes || width <= 0 || height <= 0) return RTB_EINVAL;
  try {
    const std::string file = core::EncodePng(rgb8, width, height);
    *n_bytes = file.size();
    if (!out) return RTB_OK;  // size query
    if (capacity < file.size()) throw std::runtime_error("rth_encode_png: buffer of " + std::to_string(capacity) + " bytes, need " + std::to_string(file.size()));
    std::memcpy(out, file.data(), file.size());
    return RTB_OK;
  } catch (const std::exception& e) {
    return Fail(e);
  }
}

}  // extern "C"
