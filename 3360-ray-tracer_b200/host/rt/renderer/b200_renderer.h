// rt::renderer — the drop-in boundary for the render loop (seam 1).
//
// B200Renderer renders scene+camera on one B200 through the C ABI and writes the reference's
// output: ASCII PPM "P3" on stdout (src/renderer/mega_kernel.h:48-51, src/core/color.h:18-33),
// progress on stderr.  MegaKernel and WavefrontRenderer keep the reference's constructor
// signatures (mega_kernel.h:12-13, wavefront.h:24-33) and select estimator A / B.
#pragma once
#include <rtb.h>

#include <cstdint>
#include <iostream>
#include <stdexcept>
#include <string>
#include <vector>

#include "../b200/flatten.h"
#include "../core/timer.h"
#include "../integrator/ray_integrator.h"
#include "../integrator/sampler.h"
#include "../scene/camera.h"
#include "../scene/scene.h"

namespace rt::renderer {

struct B200Options {
  int mode = RTB_MODE_A;
  int samples_per_pixel = 0;  // 0: camera.samples_per_pixel_
  int max_depth = -1;         // <0: camera.max_depth_
  uint64_t seed = 0x3360;
  int device = 0;
  uint32_t pool_size = 0;
  bool emit_ppm = true;       // write P3 to `out`
  std::ostream* out = &std::cout;
  std::ostream* log = &std::clog;
};

class B200Renderer {
 public:
  B200Renderer(const scene::Scene& world, const scene::Camera& cam, B200Options opt = {}) : world_(world), cam_(cam), opt_(opt) {}

  void Render() {
    scene::Camera& cam = const_cast<scene::Camera&>(cam_);
    cam.Initialize();  // MegaKernel::Render does the same first (mega_kernel.h:16)
    const int w = cam.get_image_width(), h = cam.get_image_height();
    const int spp = opt_.samples_per_pixel > 0 ? opt_.samples_per_pixel : cam.samples_per_pixel_;
    const int depth = opt_.max_depth >= 0 ? opt_.max_depth : cam.max_depth_;
    core::Timer clock;
    if (rtb_init(opt_.device) != RTB_OK) throw std::runtime_error(std::string("B200Renderer: ") + rtb_last_error());
    b200::FlatScene flat = b200::Flattener().Run(world_);
    const RtbSceneDesc desc = flat.desc();
    RtbScene* scene = nullptr;
    if (rtb_scene_create(&desc, &scene) != RTB_OK) throw std::runtime_error(std::string("B200Renderer: ") + rtb_last_error());
    const double t_upload = clock.elapsed();
    const RtbCamera dc = cam.ToDevice();
    RtbRenderParams prm{};
    prm.spp_begin = 0;
    prm.spp_end = spp;
    prm.max_depth = depth;
    prm.mode = opt_.mode;
    prm.seed = opt_.seed;
    prm.pool_size = opt_.pool_size;
    sum_.assign(static_cast<size_t>(w) * h * 3, 0.0f);
    const int rc = rtb_render(scene, &dc, &prm, sum_.data(), &stats_);
    std::vector<uint8_t> rgb8(sum_.size());
    const int rc2 = rc == RTB_OK ? rtb_resolve(sum_.data(), static_cast<size_t>(w) * h, spp, rgb8.data()) : rc;
    rtb_scene_destroy(scene);
    if (rc2 != RTB_OK) throw std::runtime_error(std::string("B200Renderer: ") + rtb_last_error());
    width_ = w;
    height_ = h;
    spp_ = spp;
    if (opt_.log) {
      *opt_.log << "B200: " << w << "x" << h << " spp=" << spp << " depth=" << depth << " mode=" << (opt_.mode == RTB_MODE_A ? 'A' : 'B')
                << "  flatten+upload " << t_upload << "s  render " << stats_.ms_total * 1e-3 << "s  "
                << stats_.samples / (stats_.ms_total * 1e3) << " Msamples/s  " << stats_.rays / (stats_.ms_total * 1e3) << " Mrays/s\n";
    }
    if (opt_.emit_ppm && opt_.out) {
      std::string text;
      text.reserve(rgb8.size() * 4 + 32);
      text += "P3\n" + std::to_string(w) + ' ' + std::to_string(h) + "\n255\n";
      for (size_t i = 0; i < rgb8.size(); i += 3) {
        text += std::to_string(rgb8[i]);
        text += ' ';
        text += std::to_string(rgb8[i + 1]);
        text += ' ';
        text += std::to_string(rgb8[i + 2]);
        text += '\n';
      }
      opt_.out->write(text.data(), static_cast<std::streamsize>(text.size()));
    }
    if (opt_.log) *opt_.log << "\rDone. Total samples: " << stats_.samples << "Per pixel: " << stats_.samples / (static_cast<uint64_t>(w) * h) << "\n";
  }

  const std::vector<float>& radiance_sum() const { return sum_; }
  const RtbStats& stats() const { return stats_; }
  int width() const { return width_; }
  int height() const { return height_; }
  int samples_per_pixel() const { return spp_; }

 private:
  const scene::Scene& world_;
  const scene::Camera& cam_;
  B200Options opt_;
  std::vector<float> sum_;
  RtbStats stats_{};
  int width_ = 0, height_ = 0, spp_ = 0;
};

// renderer::MegaKernel(scene, camera, sampler).Render()  — estimator A (mega_kernel.h:10-59)
class MegaKernel {
 public:
  MegaKernel(scene::Scene& scene, scene::Camera& camera, integrator::Sampler& sampler) : sampler_(sampler), world_(scene), cam_(camera) {}
  void Render() {
    B200Options o;
    o.mode = RTB_MODE_A;
    o.samples_per_pixel = sampler_.max_samples();
    if (sampler_.adaptive()) std::clog << "note: AdaptiveSampler runs at its max_samples on the device (adaptive stop: SURVEY 8f-2)\n";
    B200Renderer(world_, cam_, o).Render();
  }
 private:
  integrator::Sampler& sampler_;
  scene::Scene& world_;
  scene::Camera& cam_;
};

// renderer::WavefrontRenderer(world, cam, integrator, max_depth, max_samples, batch).Render()
// — estimator B (wavefront.cc:40-242).  `integrator` and `batch_size` are accepted for source
// compatibility; intersection runs inside the device wavefront, whose batch is the path pool.
class WavefrontRenderer {
 public:
  WavefrontRenderer(const scene::Scene& world, const scene::Camera& cam, integrator::RayIntegrator& integrator, int max_depth = 10,
                    int max_samples = 128, int batch_size = 8192)
      : world_(world), cam_(cam), integrator_(integrator), max_depth_(max_depth), max_ssp_(max_samples), batch_size_(batch_size) {}
  void Render() {
    B200Options o;
    o.mode = RTB_MODE_B;
    o.samples_per_pixel = max_ssp_;
    o.max_depth = max_depth_;
    B200Renderer(world_, cam_, o).Render();
  }
 private:
  const scene::Scene& world_;
  const scene::Camera& cam_;
  integrator::RayIntegrator& integrator_;
  int max_depth_, max_ssp_, batch_size_;
};

}  // namespace rt::renderer
